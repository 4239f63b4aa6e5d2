"""CPU-side checks: the C-ABI library loads and exports every symbol the header declares; the host
tables (mel, DCT, spline operator, segment bounds) agree with the oracle; no product code imports oracle/."""
import ctypes
import os
import re

import numpy as np
import pytest

from cvml_goalnet_b200 import _lib, dsp
from oracle import goalnet_oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    src = open(os.path.join(ROOT, "include", "goalnet_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gn_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    _lib.build()
    lib = ctypes.CDLL(_lib.LIB_PATH)
    names = _header_symbols()
    assert len(names) >= 17
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/goalnet_b200.h but not exported"
        assert n in _lib.SIGNATURES, f"{n} has no ctypes prototype in _lib.SIGNATURES"
    assert sorted(_lib.SIGNATURES) == names


def test_abi_version_and_status_strings():
    lib = _lib.load()
    assert lib.gn_abi_version() == 1
    assert lib.gn_status_string(0) == b"ok"
    assert b"argument" in lib.gn_status_string(-1)
    assert b"shape" in lib.gn_status_string(-2)


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    lib = _lib.load()
    p = ctypes.c_void_p()
    assert lib.gn_create(ctypes.byref(p), 0) == -4          # GN_E_NODEVICE
    from cvml_goalnet_b200 import utils as U
    with pytest.raises(AssertionError, match="no CPU fallback"):
        U.extract_condensed_frame_tensor_array(np.zeros((1, 8, 8, 3), np.uint8))
    with pytest.raises(AssertionError):
        U.knapsack([1, 2], [1, 2], 2)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "cvml_goalnet_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", text, flags=re.M), f


def test_mel_dct_tables_match_oracle():
    for sr in (16000, 22050):
        np.testing.assert_allclose(dsp.mel_filterbank(sr), O.mel_filterbank(sr), rtol=1e-5, atol=1e-9)
        start, count, offset, w = dsp.mel_filterbank_csr(sr)
        fb = np.zeros((128, 1025), np.float32)
        for m in range(128):
            fb[m, start[m]:start[m] + count[m]] = w[offset[m]:offset[m] + count[m]]
        np.testing.assert_array_equal(fb, dsp.mel_filterbank(sr))
    from scipy.fftpack import dct
    eye = np.eye(128)
    np.testing.assert_allclose(dsp.dct_matrix(), dct(eye, axis=0, type=2, norm="ortho")[:30], atol=1e-7)


def test_cubic_operator_matches_scipy_interp1d():
    for q, B in ((16, 30), (15, 30), (22, 30), (4, 7), (44, 60)):
        np.testing.assert_allclose(dsp.cubic_resample_matrix(q, B), O.cubic_resample_matrix(q, B), atol=2e-6)
    with pytest.raises(AssertionError):
        dsp.cubic_resample_matrix(3, 30)


def test_segment_bounds_match_oracle():
    for n_samples, n_frames in ((32000, 4), (38888, 5), (100003, 7), (8000, 1)):
        s, l = dsp.segment_bounds(n_samples, n_frames)
        ref = O.segment_bounds(n_samples, n_frames)
        assert [(int(a), int(a + b)) for a, b in zip(s, l)] == ref


def test_scale_problem_is_exact_divisor_compression():
    from cvml_goalnet_b200.utils import _scale_problem
    w, cap, rem = _scale_problem([30, 45, 0, 60], 100, 5)
    assert (w, cap, rem) == ([2, 3, 0, 4], 6, 50)       # unit = 75: 500 // 75 = 6, 500 % 75 = 50
    w, cap, rem = _scale_problem([1.5, 2.5], 3.3, 5)      # int(7.5)=7, int(12.5)=12, int(16.5)=16
    assert (w, cap, rem) == ([7, 12], 16, 0)


def test_expand_array_host():
    from cvml_goalnet_b200.utils import expand_array
    for arr, r, L in (([1, 2, 3], 2, 5), ([1, 2, 3], 2, 9), ([4, 5], 3, 2), ([7], 5, 5)):
        assert list(expand_array(list(arr), r, L)) == list(O.expand_array(list(arr), r, L))


def test_fft_passes_host_check(tmp_path):
    """gn_fft2.cuh (the MFCC kernel's 2048-point FFT: 128 threads, 16 x 16 x 8, two transforms in the packed fp32 lanes)
    run for every thread on the CPU against a direct DFT in double."""
    import shutil
    import subprocess
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not available")
    exe = str(tmp_path / "fft_host_check")
    subprocess.run([nvcc, "-std=c++17", "-O2", "-Wno-deprecated-gpu-targets", "-o", exe,
                    os.path.join(ROOT, "tests", "native", "fft2_host_check.cu")], check=True)
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
