"""ctypes binding of the C-ABI in include/goalnet_b200.h (libgoalnet_b200.so, built in-tree).

This is the only place the package touches the native library.  There is NO CPU fallback: if the
shared library is missing, or no sm_100 device is present, the compute entry points raise.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
# GN_LIB_PATH: development override to A/B two builds of the same ABI on one box (never a different backend)
LIB_PATH = os.environ.get("GN_LIB_PATH") or os.path.join(_HERE, "libgoalnet_b200.so")
CSRC = os.path.join(_HERE, "csrc")
SOURCES = ["gn_preprocess.cu", "gn_mfcc.cu", "gn_front.cu", "gn_front_tc.cu", "gn_tc_gemm.cu", "gn_model.cu", "gn_select.cu", "gn_train.cu", "gn_train_avm.cu"]
NVCC_FLAGS = ["-shared", "-Xcompiler", "-fPIC", "-std=c++17", "-O3", "-lineinfo",
              "-gencode", "arch=compute_100a,code=sm_100a"]

GN_OK = 0


class GoalNetError(AssertionError):
    """Non-zero status from the native library.  Subclasses AssertionError because the reference
    signals every error with a bare `assert ... 'E: ...'` (utils.py:45,49,345,560,609)."""

    def __init__(self, status: int, where: str, text: str):
        super().__init__(f"E: {where} failed with status {status} ({text})")
        self.status = status


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile csrc/*.cu for sm_100a into libgoalnet_b200.so (nvcc cross-compiles without a GPU)."""
    srcs = [os.path.join(CSRC, s) for s in SOURCES]
    deps = srcs + [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cuh")]
    deps.append(os.path.join(os.path.dirname(_HERE), "include", "goalnet_b200.h"))
    if not force and os.path.exists(LIB_PATH):
        newest = max(os.path.getmtime(d) for d in deps)
        if os.path.getmtime(LIB_PATH) >= newest:
            return LIB_PATH
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + NVCC_FLAGS + ["-o", LIB_PATH] + srcs
    if verbose:
        print(" ".join(cmd))
    subprocess.run(cmd, check=True, cwd=CSRC)
    return LIB_PATH


class gn_avm_state(C.Structure):
    _names = ["conv1_w", "conv1_b", "conv2_w", "conv2_b", "conv3_w", "conv3_b",
              "bn1_w", "bn1_b", "bn1_m", "bn1_v", "bn2_w", "bn2_b", "bn2_m", "bn2_v",
              "bn3_w", "bn3_b", "bn3_m", "bn3_v", "linear5_w", "linear5_b",
              "aconv1_w", "aconv1_b", "aconv2_w", "aconv2_b", "alinear3_w", "alinear3_b",
              "f0_w", "f0_b", "f3_w", "f3_b", "f6_w", "f6_b", "f9_w", "f9_b", "f12_w", "f12_b"]
    _fields_ = [(n, C.c_void_p) for n in _names]


# field of gn_avm_state -> reference state_dict key
STATE_KEYS = {
    "conv1_w": "visbl.conv1.weight", "conv1_b": "visbl.conv1.bias",
    "conv2_w": "visbl.conv2.weight", "conv2_b": "visbl.conv2.bias",
    "conv3_w": "visbl.conv3.weight", "conv3_b": "visbl.conv3.bias",
    "linear5_w": "visbl.linear5.weight", "linear5_b": "visbl.linear5.bias",
    "aconv1_w": "audbl.conv1.weight", "aconv1_b": "audbl.conv1.bias",
    "aconv2_w": "audbl.conv2.weight", "aconv2_b": "audbl.conv2.bias",
    "alinear3_w": "audbl.linear3.weight", "alinear3_b": "audbl.linear3.bias",
    "f0_w": "fusion.0.weight", "f0_b": "fusion.0.bias", "f3_w": "fusion.3.weight", "f3_b": "fusion.3.bias",
    "f6_w": "fusion.6.weight", "f6_b": "fusion.6.bias", "f9_w": "fusion.9.weight", "f9_b": "fusion.9.bias",
    "f12_w": "fusion.12.weight", "f12_b": "fusion.12.bias",
}
for _i in (1, 2, 3):
    STATE_KEYS[f"bn{_i}_w"] = f"visbl.bnorm{_i}.weight"
    STATE_KEYS[f"bn{_i}_b"] = f"visbl.bnorm{_i}.bias"
    STATE_KEYS[f"bn{_i}_m"] = f"visbl.bnorm{_i}.running_mean"
    STATE_KEYS[f"bn{_i}_v"] = f"visbl.bnorm{_i}.running_var"

class gn_score_sink(C.Structure):
    _fields_ = [("n_dst", C.c_int32), ("reserved", C.c_int32), ("dst", C.c_void_p * 16), ("rows_per_match", C.c_int64),
                ("dst_match_stride", C.c_int64), ("dst_offset", C.c_int64), ("row0", C.c_int64)]


class gn_fusion_state(C.Structure):
    _fields_ = [("weight", C.c_void_p * 5), ("bias", C.c_void_p * 5)]


_P = C.c_void_p
_I = C.c_int
_L = C.c_int64

# name -> (restype, argtypes); must list every symbol include/goalnet_b200.h declares
SIGNATURES = {
    "gn_abi_version": (_I, []),
    "gn_status_string": (C.c_char_p, [_I]),
    "gn_create": (_I, [C.POINTER(_P), _I]),
    "gn_destroy": (_I, [_P]),
    "gn_launch_count": (_L, [_P]),
    "gn_preprocess_u8_bgr": (_I, [_P, _P, _L, _I, _I, _I, _P, _P, _P]),
    "gn_mfcc_resampled": (_I, [_P, _P, _L, _P, _P, _P, _P, _L, _I, _I, _P, _P, _P, _P, _I, _P, _P, _P]),
    "gn_model_create": (_I, [_P, C.POINTER(_P), C.POINTER(gn_avm_state), _I, _I, _I, _P]),
    "gn_model_destroy": (_I, [_P]),
    "gn_model_set_classifier_head": (_I, [_P, _P, _P, _I, _P]),
    "gn_avm_workspace_bytes": (_L, [_P, _L]),
    "gn_avm_forward": (_I, [_P, _P, _P, _L, _P, _P, _P, _L, _P]),
    "gn_avm_forward_scatter": (_I, [_P, _P, _P, _L, C.POINTER(gn_score_sink), _P, _L, _P]),
    "gn_unshard_rows": (_I, [_P, _P, _I, _L, _L, _L, _P, _P]),
    "gn_model_set_timing": (_I, [_P, _I]),
    "gn_model_get_timing": (_I, [_P, C.POINTER(C.c_float), _I]),
    "gn_clip_information": (_I, [_P, _P, _L, _I, _L, _P, _I, _P, _P, _P]),
    "gn_knapsack_workspace_bytes": (_L, [_I, _L]),
    "gn_knapsack": (_I, [_P, _P, _P, _I, _L, _I, _I, _P, _P, _P, _L, _P]),
    "gn_knapsack_static": (_I, [_P, _P, _P, _I, _L, _I, _I, _P, _P, _P]),
    "gn_summary_indicator": (_I, [_P, _P, _P, _I, _L, _P, _P]),
    "gn_fscore_counts": (_I, [_P, _P, _P, _I, _L, _P, _P]),
    "gn_fusion_trainer_create": (_I, [_P, C.POINTER(_P), C.POINTER(gn_fusion_state), _I, _L, _P]),
    "gn_fusion_trainer_create_head": (_I, [_P, C.POINTER(_P), C.POINTER(gn_fusion_state), _I, _L, _I, _P]),
    "gn_fusion_trainer_destroy": (_I, [_P]),
    "gn_fusion_param_count": (_L, [_P]),
    "gn_fusion_set_hyper": (_I, [_P, C.c_float, C.c_float, C.c_float, C.c_float]),
    "gn_fusion_get_params": (_I, [_P, _P, _P]),
    "gn_fusion_forward_backward": (_I, [_P, _P, _P, _L, _I, C.c_float, C.POINTER(_P), C.c_uint64, _P, _P, _P, _P]),
    "gn_fusion_adam_step": (_I, [_P, _P, C.c_float, _P]),
    "gn_dbg_gemm_tf32": (_I, [_P, _P, _L, _I, _P, _I, _P, _P]),
    "gn_avm_trainer_create": (_I, [_P, C.POINTER(_P), C.POINTER(gn_avm_state), _I, _I, _I, _L, _P]),
    "gn_avm_trainer_destroy": (_I, [_P]),
    "gn_avm_trainer_param_count": (_L, [_P]),
    "gn_avm_trainer_set_hyper": (_I, [_P, C.c_float, C.c_float, C.c_float, C.c_float]),
    "gn_avm_trainer_forward_backward": (_I, [_P, _P, _P, _P, _L, _I, C.c_float, C.POINTER(_P), C.c_uint64, _P, _P, _P]),
    "gn_avm_trainer_grads": (_I, [_P, C.POINTER(_P)]),
    "gn_avm_trainer_adam_step": (_I, [_P, C.c_float, _P]),
    "gn_avm_train_step": (_I, [_P, _P, _P, _P, _L, _I, C.c_float, C.POINTER(_P), C.c_uint64, _P, _P, _P]),
    "gn_avm_trainer_export": (_I, [_P, C.POINTER(gn_avm_state), _P]),
    "gn_fusion_train_step": (_I, [_P, _P, _P, _L, _I, C.c_float, C.POINTER(_P), C.c_uint64, _P, _P, _P, _P]),
    "gn_dbg_avm_forward_simt": (_I, [_P, _P, _P, _L, _P, _P, _P, _L, _P]),
    "gn_dbg_read_activation": (_I, [_P, _I, _L, _P, _P, _P]),
}

_lib = None
_lock = threading.Lock()


def load():
    """dlopen the library and declare every prototype.  Raises (never falls back) when it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise GoalNetError(-4, "load", f"{LIB_PATH} not built; run `python -c 'import __graft_entry__ as g; "
                                               f"g.build()'` -- there is no CPU fallback")
            lib = C.CDLL(LIB_PATH)
            for name, (res, args) in SIGNATURES.items():
                fn = getattr(lib, name)
                fn.restype = res
                fn.argtypes = args
            _lib = lib
    return _lib


def check(status: int, where: str):
    if status != GN_OK:
        text = load().gn_status_string(int(status))
        raise GoalNetError(int(status), where, text.decode() if text else "?")


_handles = {}


def handle(device_index: int = 0):
    """Per-device gn_handle, created on first use."""
    h = _handles.get(device_index)
    if h is None:
        lib = load()
        p = _P()
        check(lib.gn_create(C.byref(p), int(device_index)), "gn_create")
        h = _handles[device_index] = p
    return h


_graph_launches = {}      # device index -> kernel launches executed by CUDA-graph replays (not seen by the library)


def add_graph_launches(device_index: int, n: int):
    _graph_launches[device_index] = _graph_launches.get(device_index, 0) + int(n)


def launch_count(device_index: int = 0) -> int:
    """Kernel launches of this library on the device: issued directly (counted by the library) plus those executed by
    replays of captured CUDA graphs (counted per replay from the number of launches recorded during the capture)."""
    return int(load().gn_launch_count(handle(device_index))) + _graph_launches.get(device_index, 0)
