"""Host-side constant tables for the audio kernel (mel filterbank, DCT rows, cubic resampling
operator, segment bounds).  Pure NumPy, float64 internally; no SciPy / librosa dependency.

These are the constants `librosa.feature.mfcc` (reference utils.py:333) and the `interp1d(kind='cubic')`
loop (utils.py:336-343) evaluate implicitly; the CUDA kernel consumes them as device arrays.
"""
from __future__ import annotations

from functools import lru_cache

import numpy as np

N_FFT = 2048
HOP = 512
N_MELS = 128
N_MFCC = 30
MAX_Q = 128         # GN_MFCC_MAX_Q: STFT columns per frame slot the kernel sizes its shared-memory tiles for


def _hz_to_mel(f):
    """Slaney mel scale (librosa htk=False): linear below 1 kHz, log above."""
    f = np.asarray(f, dtype=np.float64)
    f_sp = 200.0 / 3.0
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    with np.errstate(divide="ignore"):
        log_part = min_log_mel + np.log(np.maximum(f, 1e-300) / min_log_hz) / logstep
    return np.where(f >= min_log_hz, log_part, f / f_sp)


def _mel_to_hz(m):
    m = np.asarray(m, dtype=np.float64)
    f_sp = 200.0 / 3.0
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    return np.where(m >= min_log_mel, min_log_hz * np.exp(logstep * (m - min_log_mel)), f_sp * m)


@lru_cache(maxsize=8)
def mel_filterbank(sr: int) -> np.ndarray:
    """(128, 1025) float32 triangular filters, Slaney-normalised, fmin 0, fmax sr/2."""
    freqs = np.arange(1 + N_FFT // 2, dtype=np.float64) * (float(sr) / N_FFT)
    edges = _mel_to_hz(np.linspace(_hz_to_mel(0.0), _hz_to_mel(sr / 2.0), N_MELS + 2))
    width = np.diff(edges)
    fb = np.zeros((N_MELS, freqs.size), dtype=np.float64)
    for m in range(N_MELS):
        rise = (freqs - edges[m]) / width[m]
        fall = (edges[m + 2] - freqs) / width[m + 1]
        fb[m] = np.maximum(0.0, np.minimum(rise, fall))
    fb *= (2.0 / (edges[2:] - edges[:-2]))[:, None]
    return fb.astype(np.float32)


@lru_cache(maxsize=8)
def mel_filterbank_csr(sr: int):
    """Band m covers bins [start[m], start[m]+count[m]) with weights w[offset[m]:offset[m]+count[m]]."""
    fb = mel_filterbank(sr)
    start = np.zeros(N_MELS, np.int32)
    count = np.zeros(N_MELS, np.int32)
    offset = np.zeros(N_MELS, np.int32)
    w = []
    for m in range(N_MELS):
        nz = np.nonzero(fb[m])[0]
        if nz.size:
            start[m], count[m] = nz[0], nz[-1] - nz[0] + 1
        offset[m] = len(w)
        w.extend(fb[m, start[m]:start[m] + count[m]].tolist())
    if not w:
        w = [0.0]
    return start, count, offset, np.asarray(w, dtype=np.float32)


@lru_cache(maxsize=2)
def dct_matrix() -> np.ndarray:
    """(30, 128) float32 rows of the orthonormal DCT-II (scipy.fftpack.dct type 2, norm='ortho')."""
    n = np.arange(N_MELS, dtype=np.float64)
    k = np.arange(N_MFCC, dtype=np.float64)[:, None]
    d = np.cos(np.pi * k * (2.0 * n + 1.0) / (2.0 * N_MELS)) * np.sqrt(2.0 / N_MELS)
    d[0] *= np.sqrt(0.5)
    return d.astype(np.float32)


@lru_cache(maxsize=64)
def cubic_resample_matrix(q: int, B: int) -> np.ndarray:
    """(B, q) float32 operator of a not-a-knot cubic spline through (0..q-1, y) evaluated at
    linspace(0, q-1, B) -- what interp1d(kind='cubic') computes at utils.py:337-343."""
    if q < 4:
        raise AssertionError("E: cubic interpolation needs at least 4 STFT columns per frame slot")
    # second derivatives m solve A m = R y (unit spacing); rows 0 and q-1 are the not-a-knot conditions
    A = np.zeros((q, q))
    R = np.zeros((q, q))
    for i in range(1, q - 1):
        A[i, i - 1:i + 2] = (1.0, 4.0, 1.0)
        R[i, i - 1:i + 2] = (6.0, -12.0, 6.0)
    A[0, :3] = (1.0, -2.0, 1.0)
    A[q - 1, q - 3:] = (1.0, -2.0, 1.0)
    M = np.linalg.solve(A, R)                      # m = M y
    x = np.linspace(0.0, q - 1.0, B)
    i = np.minimum(np.floor(x).astype(np.int64), q - 2)
    t = x - i
    W = np.zeros((B, q))
    rows = np.arange(B)
    W[rows, i] += 1.0 - t
    W[rows, i + 1] += t
    W += (((1.0 - t) ** 3 - (1.0 - t)) / 6.0)[:, None] * M[i]
    W += ((t ** 3 - t) / 6.0)[:, None] * M[i + 1]
    return W.astype(np.float32)


def segment_bounds(n_samples: int, n_frames: int):
    """utils.py:322-330: start = round(i*spf), end = min(round(start+spf), n) with Python's half-even
    round; returns (start int64[n_frames], length int32[n_frames])."""
    spf = n_samples / n_frames
    start = np.empty(n_frames, np.int64)
    length = np.empty(n_frames, np.int32)
    for i in range(n_frames):
        s = round(i * spf)
        e = round(s + spf)
        if e > n_samples:
            e = n_samples
        start[i] = s
        length[i] = max(e - s, 0)
    return start, length
