"""Drop-in mirror of the reference's `utils.py` for the per-frame scoring path.

Same names, argument meaning and error convention (`AssertionError('E: ...')`) as the functions
`main.py:11` / `baseline.py:9` import from the reference's `utils`; the arithmetic runs in the
sm_100a CUDA kernels behind the C-ABI of include/goalnet_b200.h.  PyTorch is used for device memory
and streams only.  File-based signatures keep their reference form and delegate, after the I/O, to an
array-based twin (`*_array`) -- the twins are what the tests, the benchmark and multi-GPU code call.

There is no CPU fallback: every function here needs the native library and a B200.
"""
from __future__ import annotations

import ctypes as C
import math
from functools import lru_cache

import numpy as np
import torch
import torch.nn as nn

from . import _lib, dsp

__all__ = [
    "AVM", "VisBl", "AudBl", "dataloader", "extract_condensed_frame_tensor", "extract_condensed_frame_tensor_array",
    "extract_audio_features", "extract_audio_features_array", "expand_array", "get_clip_information", "knapsack",
    "postprocess", "postprocess_array", "get_fscore", "postprocess_and_get_fscores",
    "postprocess_and_get_fscores_array", "get_frame_tensor",
]


# --------------------------------------------------------------------------------------------------
# plumbing
# --------------------------------------------------------------------------------------------------
def _device(device=None) -> torch.device:
    if not torch.cuda.is_available():
        raise _lib.GoalNetError(-4, "device", "no CUDA device: this package has no CPU fallback")
    if device is None:
        return torch.device("cuda", torch.cuda.current_device())
    device = torch.device(device)
    if device.type != "cuda":
        raise _lib.GoalNetError(-4, "device", "a CUDA device is required")
    return torch.device("cuda", device.index if device.index is not None else torch.cuda.current_device())


def _to_dev(x, dtype, device) -> torch.Tensor:
    if isinstance(x, np.ndarray):
        x = torch.from_numpy(np.ascontiguousarray(x))
    elif not torch.is_tensor(x):
        x = torch.as_tensor(x)
    return x.to(device=device, dtype=dtype, non_blocking=True).contiguous()


def _ptr(t) -> C.c_void_p:
    return C.c_void_p(t.data_ptr() if t is not None else 0)


def _stream(device) -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _h(device):
    return _lib.handle(device.index)


# --------------------------------------------------------------------------------------------------
# (a1) frames  -- reference utils.py:274-292
# --------------------------------------------------------------------------------------------------
def extract_condensed_frame_tensor_array(frames_u8, S: int = 40, device=None, out=None, as_numpy: bool = False):
    """Array twin of `extract_condensed_frame_tensor`: uint8 BGR (N,H,W,3) already decoded and
    subsampled -> float32 (N,3,S,S) on the device (utils.py:284-285,292)."""
    dev = _device(device if device is not None else (frames_u8.device if torch.is_tensor(frames_u8) and frames_u8.is_cuda else None))
    fr = _to_dev(frames_u8, torch.uint8, dev)
    assert fr.dim() == 4 and fr.shape[-1] == 3, "E: frames must have shape (N, H, W, 3)"
    n, H, W, _ = fr.shape
    if out is None:
        out = torch.empty((n, 3, S, S), dtype=torch.float32, device=dev)
    if n == 0:
        return out.cpu().numpy() if as_numpy else out
    with torch.cuda.device(dev):
        _lib.check(_lib.load().gn_preprocess_u8_bgr(_h(dev), _ptr(fr), n, H, W, S, _ptr(out), None, _stream(dev)),
                   "gn_preprocess_u8_bgr")
    return out.cpu().numpy() if as_numpy else out


def extract_condensed_frame_tensor(fp: str, skip_frames: int):
    """utils.py:274-292, same signature and return: (np.ndarray (N,3,40,40) float32, full_n_frames).
    Decoding is host I/O through OpenCV exactly as in the reference; the arithmetic is on the GPU."""
    import cv2
    video = cv2.VideoCapture(fp)
    count = 0
    success = True
    frames = []
    while success:
        success, image = video.read()
        if count % skip_frames == 0 and success:
            frames.append(image)
        count += 1
    full_n_frames = count - 1
    video.release()
    if not frames:
        return np.zeros((0, 3, 40, 40), np.float32), full_n_frames
    return extract_condensed_frame_tensor_array(np.stack(frames), 40, as_numpy=True), full_n_frames


def get_frame_tensor(fp: str):
    """utils.py:294-305 (pure I/O, kept for `infer`)."""
    import cv2
    video = cv2.VideoCapture(fp)
    frames = []
    while True:
        success, image = video.read()
        if not success:
            break
        frames.append(image)
    return np.array(frames)


# --------------------------------------------------------------------------------------------------
# (a2) audio  -- reference utils.py:313-349
# --------------------------------------------------------------------------------------------------
@lru_cache(maxsize=8)
def _audio_tables(sr: int, dev_index: int):
    dev = torch.device("cuda", dev_index)
    start, count, offset, w = dsp.mel_filterbank_csr(int(sr))
    dct_t = np.zeros((dsp.N_MELS, 32), np.float32)              # transposed + padded: a warp's 32 coefficients are one line
    dct_t[:, :dsp.N_MFCC] = dsp.dct_matrix().T
    return dict(start=torch.from_numpy(start).to(dev), count=torch.from_numpy(count).to(dev),
                offset=torch.from_numpy(offset).to(dev), w=torch.from_numpy(w).to(dev), n_w=int(w.size),
                dct=torch.from_numpy(dct_t).to(dev))


@lru_cache(maxsize=32)
def _segment_plan(n_samples: int, n_frames: int, B: int, dev_index: int):
    dev = torch.device("cuda", dev_index)
    start, length = dsp.segment_bounds(n_samples, n_frames)
    q = 1 + length // dsp.HOP
    assert n_frames == 0 or int(q.min()) >= 4, "E: Shape mismatch"   # interp1d(kind='cubic') needs >= 4 columns
    assert n_frames == 0 or int(q.max()) <= dsp.MAX_Q, "E: audio segment too long for one frame slot"
    offs, mats, table = np.zeros(n_frames, np.int32), [], {}
    pos = 0
    for qq in np.unique(q):
        table[int(qq)] = pos
        mats.append(dsp.cubic_resample_matrix(int(qq), B).reshape(-1))
        pos += mats[-1].size
    for i in range(n_frames):
        offs[i] = table[int(q[i])]
    wcat = np.concatenate(mats) if mats else np.zeros(1, np.float32)
    return dict(start=torch.from_numpy(start).to(dev), length=torch.from_numpy(length).to(dev),
                w_offset=torch.from_numpy(offs).to(dev), w=torch.from_numpy(wcat).to(dev),
                max_q=int(q.max()) if n_frames else 4, start_host=start, length_host=length)


def audio_slot_sample_range(n_samples: int, n_frames: int, a: int, b: int):
    """Sample range [lo, hi) that the frame slots a..b-1 of utils.py:322-330 read."""
    start, length = dsp.segment_bounds(int(n_samples), int(n_frames)) if n_frames else (np.zeros(0), np.zeros(0))
    if b <= a:
        return 0, 0
    return int(start[a]), int((start[a:b] + length[a:b]).max())


def segment_plan_from_bounds(start, length, B: int, device=None):
    """Device tables for `extract_audio_features_array(..., plan=...)` from explicit slot bounds (sample index of the
    first sample and sample count of every frame slot inside the buffer that will be passed as `y`).  Used where the
    buffer is not one whole track: e.g. a rank's shard of several matches laid end to end."""
    dev = _device(device)
    start = np.ascontiguousarray(start, dtype=np.int64)
    length = np.ascontiguousarray(length, dtype=np.int32)
    n_frames = int(start.shape[0])
    q = 1 + length // dsp.HOP
    assert n_frames == 0 or int(q.min()) >= 4, "E: Shape mismatch"
    assert n_frames == 0 or int(q.max()) <= dsp.MAX_Q, "E: audio segment too long for one frame slot"
    table, mats, pos = {}, [], 0
    for qq in np.unique(q):
        table[int(qq)] = pos
        mats.append(dsp.cubic_resample_matrix(int(qq), B).reshape(-1))
        pos += mats[-1].size
    offs = np.asarray([table[int(v)] for v in q], dtype=np.int32) if n_frames else np.zeros(0, np.int32)
    wcat = np.concatenate(mats) if mats else np.zeros(1, np.float32)
    return dict(start=torch.from_numpy(start).to(dev), length=torch.from_numpy(length).to(dev),
                w_offset=torch.from_numpy(offs).to(dev), w=torch.from_numpy(wcat).to(dev),
                max_q=int(q.max()) if n_frames else 4, start_host=start, length_host=length)


def extract_audio_features_array(y, sr: int, n_frames: int, bin_length: int, device=None, out=None,
                                 as_numpy: bool = False, slots=None, plan=None):
    """Array twin of `extract_audio_features` (utils.py:313-349) with `(y, sr)` replacing librosa.load:
    float32 (N,30,B) MFCC features, cubic-resampled to B columns per frame slot.
    `slots=(a, b)` computes only the frame slots a..b-1 (rows a..b-1 of `out`), for chunked pipelines.
    `plan` (from `segment_plan_from_bounds`) replaces the slot bounds utils.py:322-330 derives from len(y)."""
    dev = _device(device if device is not None else (y.device if torch.is_tensor(y) and y.is_cuda else None))
    yd = _to_dev(y, torch.float32, dev).reshape(-1)
    B = int(bin_length)
    if int(n_frames) == 0:
        out = torch.empty((0, dsp.N_MFCC, B), dtype=torch.float32, device=dev)
        return out.cpu().numpy() if as_numpy else out
    if plan is None:
        plan = _segment_plan(int(yd.numel()), int(n_frames), B, dev.index)
    else:
        assert int(plan["start"].shape[0]) == int(n_frames), "E: Shape mismatch"
    tab = _audio_tables(int(sr), dev.index)
    if out is None:
        out = torch.empty((n_frames, dsp.N_MFCC, B), dtype=torch.float32, device=dev)
    a, b = (0, int(n_frames)) if slots is None else (int(slots[0]), int(slots[1]))
    assert 0 <= a <= b <= n_frames, "E: Shape mismatch"
    if b > a:
        with torch.cuda.device(dev):
            _lib.check(_lib.load().gn_mfcc_resampled(
                _h(dev), _ptr(yd), yd.numel(), _ptr(plan["start"][a:]), _ptr(plan["length"][a:]),
                _ptr(plan["w_offset"][a:]), _ptr(plan["w"]), b - a, B, plan["max_q"], _ptr(tab["start"]),
                _ptr(tab["count"]), _ptr(tab["offset"]), _ptr(tab["w"]), tab["n_w"], _ptr(tab["dct"]), _ptr(out[a:]),
                _stream(dev)),
                "gn_mfcc_resampled")
    return out.cpu().numpy() if as_numpy else out


def extract_audio_features(audio_fp: str, n_frames: int, bin_length: int = 30) -> np.ndarray:
    """utils.py:313-349.  `bin_length` defaults to the paper's B = 30 so that `infer`'s two-argument call
    (main.py:321, quirk Q2) works.  Loading/resampling the file is librosa's job, as in the reference."""
    try:
        import librosa
    except ImportError as e:  # pragma: no cover - librosa is not in this image
        raise AssertionError("E: librosa is required to load audio files; use extract_audio_features_array(y, sr, ...)") from e
    y, sr = librosa.load(audio_fp)
    feats = extract_audio_features_array(y, sr, n_frames, bin_length, as_numpy=True)
    assert feats.shape[-1] == bin_length, "E: Shape mismatch"
    return feats


# --------------------------------------------------------------------------------------------------
# (a3-a5) model  -- reference utils.py:145-272
# --------------------------------------------------------------------------------------------------
class VisBl(nn.Module):
    """Parameter container with the reference's names/shapes (utils.py:145-170), materialised for SxS frames."""

    def __init__(self, S: int = 40):
        super().__init__()
        c1 = (S + 6 - 3) // 3 + 1
        p3 = c1 - 6
        self.conv1 = nn.Conv2d(3, 64, 3, stride=3, padding=3)
        self.bnorm1 = nn.BatchNorm2d(64)
        self.conv2 = nn.Conv2d(64, 256, 3, stride=1, padding=1)
        self.bnorm2 = nn.BatchNorm2d(256)
        self.conv3 = nn.Conv2d(256, 512, 3, stride=1, padding=1)
        self.bnorm3 = nn.BatchNorm2d(512)
        self.linear5 = nn.Linear(512 * p3 * p3, 512)


class AudBl(nn.Module):
    """utils.py:197-212 parameter container, materialised for B audio bins."""

    def __init__(self, B: int = 30):
        super().__init__()
        l1 = (B + 2 - 3) // 2 + 1
        l2 = (l1 + 2 - 3) // 2 + 1
        self.conv1 = nn.Conv1d(30, 64, 3, stride=2, padding=1)
        self.conv2 = nn.Conv1d(64, 128, 3, stride=2, padding=1)
        self.linear3 = nn.Linear(128 * l2, 128)


class AVM(nn.Module):
    """Audio-visual frame-importance model (utils.py:229-272), eval-mode forward on the B200.

    `AVM(audio_included)`; `forward(audio_input, visual_input)` (audio FIRST, utils.py:260) returns a
    (N,1) float32 CUDA tensor in [1,5].  `state_dict()` / `load_state_dict()` use the reference's key
    names and tensor layouts, so checkpoints written by the reference's `torch.save(model.state_dict())`
    (main.py:263,282) load unchanged.  Dropout is the identity and BatchNorm uses running statistics
    (parity is defined in eval mode).
    """

    def __init__(self, audio_included: bool, S: int = 40, B: int = 30):
        super().__init__()
        self.audio_included = bool(audio_included)
        self.S, self.B = S, B
        self.visbl = VisBl(S)
        if self.audio_included:
            self.audbl = AudBl(B)
        fin = 512 + (128 if self.audio_included else 0)
        # same Sequential indices as the reference (ReLU / Dropout occupy 1,2,4,5,7,8,10,11,13)
        self.fusion = nn.Sequential(
            nn.Linear(fin, 512), nn.ReLU(), nn.Dropout(0.2), nn.Linear(512, 512), nn.ReLU(), nn.Dropout(0.2),
            nn.Linear(512, 256), nn.ReLU(), nn.Dropout(0.2), nn.Linear(256, 128), nn.ReLU(), nn.Dropout(0.2),
            nn.Linear(128, 1), nn.Sigmoid())
        self._native = {}       # device index -> (gn_model*, packed state version)
        self._version = 0
        self._workspace = {}
        self.last_visual_features = None

    # -- weights ----------------------------------------------------------------------------------
    def load_state_dict(self, state_dict, strict: bool = True, assign: bool = False):
        sd = {k: (torch.as_tensor(np.asarray(v)) if not torch.is_tensor(v) else v) for k, v in state_dict.items()}
        res = super().load_state_dict(sd, strict=strict)
        self._version += 1
        return res

    def invalidate(self):
        """Call after mutating parameters in place so that the packed device copy is rebuilt."""
        self._version += 1

    def _release(self):
        lib = _lib.load()
        for (mptr, _) in self._native.values():
            lib.gn_model_destroy(mptr)
        self._native = {}

    def __del__(self):
        try:
            self._release()
        except Exception:
            pass

    def _model(self, dev):
        cur = self._native.get(dev.index)
        if cur is not None and cur[1] == self._version:
            return cur[0]
        lib = _lib.load()
        if cur is not None:
            lib.gn_model_destroy(cur[0])
            del self._native[dev.index]
        sd = self.state_dict()
        keep, st = [], _lib.gn_avm_state()
        for field, key in _lib.STATE_KEYS.items():
            if key not in sd:
                setattr(st, field, None)
                continue
            t = sd[key].detach().to(device=dev, dtype=torch.float32).contiguous()
            keep.append(t)
            setattr(st, field, t.data_ptr())
        mptr = C.c_void_p()
        with torch.cuda.device(dev):
            _lib.check(lib.gn_model_create(_h(dev), C.byref(mptr), C.byref(st), int(self.audio_included), self.S,
                                           self.B, _stream(dev)), "gn_model_create")
            self._after_create(lib, mptr, sd, dev)
        del keep
        self._native[dev.index] = (mptr, self._version)
        return mptr

    OUT_COLS = 1          # columns of forward()'s result

    def _after_create(self, lib, mptr, sd, dev):
        """Hook for head variants (CAVM)."""

    def _ws(self, dev, mptr, n):
        """Workspace view for n frames, 1024-byte aligned (TMA / swizzle atoms), grown on demand."""
        need = int(_lib.load().gn_avm_workspace_bytes(mptr, n))
        ws = self._workspace.get(dev.index)
        if ws is None or ws.numel() < need + 1024:
            ws = self._workspace[dev.index] = torch.empty(need + 1024, dtype=torch.uint8, device=dev)
        off = (-ws.data_ptr()) % 1024
        return ws[off:off + need], need

    # -- forward ------------------------------------------------------------------------------------
    def forward(self, audio_input, visual_input, out=None, return_features: bool = False, _simt: bool = False):
        dev = _device(visual_input.device if torch.is_tensor(visual_input) and visual_input.is_cuda else None)
        vis = _to_dev(visual_input, torch.float32, dev)
        assert vis.dim() == 4 and tuple(vis.shape[1:]) == (3, self.S, self.S), "E: Invalid shape for visual input"
        n = vis.shape[0]
        aud = None
        if self.audio_included:
            aud = _to_dev(audio_input, torch.float32, dev)
            assert tuple(aud.shape) == (n, 30, self.B), "E: Inconsistency in data loader definition"
        mptr = self._model(dev)
        ws, need = self._ws(dev, mptr, n)
        scores = out if out is not None else torch.empty((n, self.OUT_COLS), dtype=torch.float32, device=dev)
        assert scores.is_cuda and scores.dtype == torch.float32 and scores.numel() == n * self.OUT_COLS and \
            scores.is_contiguous(), "E: Invalid output buffer"
        feats = torch.empty((n, 512), dtype=torch.float32, device=dev) if return_features else None
        lib = _lib.load()
        fn = lib.gn_dbg_avm_forward_simt if _simt else lib.gn_avm_forward
        with torch.cuda.device(dev):
            _lib.check(fn(mptr, _ptr(vis), _ptr(aud), n, _ptr(scores), _ptr(feats), _ptr(ws), need, _stream(dev)),
                       "gn_avm_forward")
        self.last_visual_features = feats
        return scores

    def forward_scatter(self, audio_input, visual_input, dst_ptrs, rows_per_match: int, dst_match_stride: int,
                        dst_offset: int, row0: int = 0):
        """Forward whose scores are stored straight into the (M, n) score matrix of every rank from the last GEMM's
        epilogue (`gn_avm_forward_scatter`; BASELINE config 3 / SURVEY 8e).  `dst_ptrs`: device addresses of the
        destination matrices (peer-mapped for other GPUs).  Inputs must already be float32 CUDA tensors."""
        dev = visual_input.device
        n = int(visual_input.shape[0])
        assert visual_input.is_cuda and visual_input.dtype == torch.float32 and visual_input.is_contiguous() and \
            tuple(visual_input.shape[1:]) == (3, self.S, self.S), "E: Invalid shape for visual input"
        if self.audio_included:
            assert audio_input.is_cuda and audio_input.dtype == torch.float32 and audio_input.is_contiguous() and \
                tuple(audio_input.shape) == (n, 30, self.B), "E: Inconsistency in data loader definition"
        assert 1 <= len(dst_ptrs) <= 16, "E: too many score destinations"
        mptr = self._model(dev)
        ws, need = self._ws(dev, mptr, n)
        sink = _lib.gn_score_sink()
        sink.n_dst = len(dst_ptrs)
        for i, p in enumerate(dst_ptrs):
            sink.dst[i] = int(p)
        sink.rows_per_match, sink.dst_match_stride = int(rows_per_match), int(dst_match_stride)
        sink.dst_offset, sink.row0 = int(dst_offset), int(row0)
        with torch.cuda.device(dev):
            _lib.check(_lib.load().gn_avm_forward_scatter(mptr, _ptr(visual_input), _ptr(audio_input if self.audio_included else None),
                                                          n, C.byref(sink), _ptr(ws), need, _stream(dev)),
                       "gn_avm_forward_scatter")

    STAGES = ("visbl_block1", "audbl", "conv2_block", "conv3_block", "linear5", "fusion_mlp")

    def set_timing(self, enable: bool, device=None):
        """Record CUDA events around each stage of the forward (read them with `stage_ms`)."""
        dev = _device(device)
        _lib.check(_lib.load().gn_model_set_timing(self._model(dev), int(bool(enable))), "gn_model_set_timing")

    def stage_ms(self, device=None) -> dict:
        dev = _device(device)
        buf = (C.c_float * len(self.STAGES))()
        _lib.check(_lib.load().gn_model_get_timing(self._model(dev), buf, len(self.STAGES)), "gn_model_get_timing")
        return {k: float(buf[i]) for i, k in enumerate(self.STAGES)}

    def read_activation(self, which: int, n: int, device=None):
        """Validation helper: block outputs 1..3 as (n,C,P,P) float32, 4 = fusion input (n,Fin)."""
        dev = _device(device)
        mptr = self._model(dev)
        ws, _ = self._ws(dev, mptr, n)
        p1 = (self.S + 3) // 3 - 1          # conv1 k3 s3 p3 -> (S+3)//3 + 1, then MaxPool(3,1) per block
        shapes = {1: (n, 64, p1, p1), 2: (n, 256, p1 - 2, p1 - 2), 3: (n, 512, p1 - 4, p1 - 4),
                  4: (n, 512 + (128 if self.audio_included else 0))}
        out = torch.empty(shapes[which], dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            _lib.check(_lib.load().gn_dbg_read_activation(mptr, which, n, _ptr(ws), _ptr(out), _stream(dev)),
                       "gn_dbg_read_activation")
        return out


# --------------------------------------------------------------------------------------------------
# (a6-a9) keyshot selection  -- reference utils.py:396-410, 445-510, 606-643
# --------------------------------------------------------------------------------------------------
class CAVM(AVM):
    """Classifier variant of AVM ("CAVM" / "CVM" in the report's Table 1).

    The reference keeps it as commented lines: `# nn.Softmax(dim = 1),` (utils.py:257), `# criterion =
    nn.CrossEntropyLoss()` (main.py:69) and `# batch_predictions = torch.argmax(batch_predictions, axis = 1) + 1`
    (main.py:97,112,190,224).  fusion.12 is Linear(128, 5), fusion.13 is Softmax(dim=1) and there is no 4y+1:
    `forward` returns (N,5) class probabilities, `predict` the (N,1) importance argmax+1 that `postprocess` takes.
    Same state_dict key names as AVM (fusion.12.weight is (5,128)).
    """

    OUT_COLS = 5

    def __init__(self, audio_included: bool, S: int = 40, B: int = 30):
        super().__init__(audio_included, S, B)
        self.fusion[12] = nn.Linear(128, self.OUT_COLS)
        self.fusion[13] = nn.Softmax(dim=1)

    def _after_create(self, lib, mptr, sd, dev):
        w = sd["fusion.12.weight"].detach().to(device=dev, dtype=torch.float32).contiguous()
        b = sd["fusion.12.bias"].detach().to(device=dev, dtype=torch.float32).contiguous()
        assert tuple(w.shape) == (self.OUT_COLS, 128) and tuple(b.shape) == (self.OUT_COLS,), "E: Invalid classifier head"
        _lib.check(lib.gn_model_set_classifier_head(mptr, _ptr(w), _ptr(b), self.OUT_COLS, _stream(dev)),
                   "gn_model_set_classifier_head")
        torch.cuda.current_stream(dev).synchronize()      # w, b may be temporaries

    def forward(self, audio_input, visual_input, out=None, return_features: bool = False, _simt: bool = False):
        assert not _simt, "E: the validation path has the regression head only"
        return super().forward(audio_input, visual_input, out=out, return_features=return_features)

    def predict(self, audio_input, visual_input):
        """argmax + 1 (main.py:97, commented) as float32 (N,1), the shape `postprocess` expects."""
        probs = self.forward(audio_input, visual_input)
        return (torch.argmax(probs, dim=1, keepdim=True) + 1).to(torch.float32)


def expand_array(arr, expansion_rate, length):
    """utils.py:396-410 (list plumbing; on the hot path it is fused into gn_clip_information)."""
    if len(arr) == length:
        return arr
    expanded_arr = [el for el in arr for _ in range(expansion_rate)]
    if len(expanded_arr) >= length:
        return expanded_arr[:length]
    return expanded_arr + [expanded_arr[-1]] * (length - len(expanded_arr))


def _clip_information_device(scores_dev, n, skip_frames, n_raw, clips_dev, dev, values=None, lengths=None):
    n_clips = clips_dev.shape[0]
    if values is None:
        values = torch.empty(n_clips, dtype=torch.int32, device=dev)
        lengths = torch.empty(n_clips, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        _lib.check(_lib.load().gn_clip_information(_h(dev), _ptr(scores_dev), n, skip_frames, n_raw, _ptr(clips_dev),
                                                   n_clips, _ptr(values), _ptr(lengths), _stream(dev)),
                   "gn_clip_information")
    return values, lengths


def get_clip_information(clip_intervals, importances, device=None):
    """utils.py:445-464: per clip [a,b]: value = sum(imp[a:b]), length = len(imp[a:b]) (end-exclusive)."""
    dev = _device(device)
    imp = np.asarray(importances)
    assert imp.ndim == 1, "E: importances must be a flat per-frame list"
    n_raw = int(imp.shape[0])
    ci = np.asarray(clip_intervals, dtype=np.int64).reshape(-1, 2)
    if ci.shape[0] == 0:
        return [], [], clip_intervals
    assert n_raw > 0 and np.all(np.abs(imp) <= 127) and np.all(imp == np.round(imp)), "E: importances must be int8 values"
    assert ci.min() >= 0, "E: negative clip bounds are not supported"
    v, l = _clip_information_device(_to_dev(imp, torch.float32, dev), n_raw, 1, n_raw,
                                    _to_dev(np.minimum(ci, 2 ** 31 - 1), torch.int32, dev), dev)
    return v.cpu().tolist(), l.cpu().tolist(), clip_intervals


class SelectBuffers:
    """Preallocated device buffers of the selection stage for `batch` problems of up to `n_clips` clips over `n_raw`
    raw frames: with these, `postprocess_array` allocates nothing and launches no library fill inside a scoring step.
    Problems with fewer clips are padded with (value 0, weight 0) items, which the reference's table never selects
    (utils.py:487-492: K[i][w] = max(0 + K[i-1][w], K[i-1][w]); the backtrack at :500 skips them)."""

    def __init__(self, n_clips: int, n_raw: int, batch: int = 1, device=None):
        dev = _device(device)
        self.dev, self.n_clips, self.n_raw, self.batch = dev, int(n_clips), int(n_raw), int(batch)
        nc = max(self.n_clips, 1)
        self.values = torch.zeros((batch, nc), dtype=torch.int32, device=dev)
        self.lengths = torch.zeros((batch, nc), dtype=torch.int32, device=dev)
        self.sel = torch.zeros((batch, nc), dtype=torch.uint8, device=dev)
        self.best = torch.zeros(batch, dtype=torch.int32, device=dev)
        self.cap = int(0.15 * self.n_raw)                                          # utils.py:629
        self.ws_bytes = int(_lib.load().gn_knapsack_workspace_bytes(self.n_clips, self.cap)) * batch
        self.ws = torch.empty(max(self.ws_bytes, 256), dtype=torch.uint8, device=dev)
        self.ind = torch.zeros((batch, self.n_raw), dtype=torch.uint8, device=dev)
        self.counts = None


def _knapsack_device(values_dev, weights_dev, n_items, cap_cols, cap_rem, batch, dev, bufs: "SelectBuffers" = None):
    lib = _lib.load()
    need = int(lib.gn_knapsack_workspace_bytes(n_items, cap_cols)) * batch
    if bufs is not None:
        assert bufs.ws_bytes >= need and bufs.sel.shape[0] >= batch and bufs.sel.shape[1] == max(n_items, 1), \
            "E: selection buffers do not fit this problem"
        ws, sel, best = bufs.ws, bufs.sel, bufs.best
    else:
        ws = torch.empty(max(need, 256), dtype=torch.uint8, device=dev)
        sel = torch.empty((batch, max(n_items, 1)), dtype=torch.uint8, device=dev)
        best = torch.empty(batch, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        _lib.check(lib.gn_knapsack(_h(dev), _ptr(values_dev), _ptr(weights_dev), n_items, cap_cols, cap_rem, batch,
                                   _ptr(sel), _ptr(best), _ptr(ws), need, _stream(dev)), "gn_knapsack")
    return sel[:batch, :n_items], best


def _scale_problem(weights, capacity, scale_factor):
    """utils.py:478-479 on the host, plus the exact common-divisor compression of the table columns."""
    w = [int(x * scale_factor) for x in weights]
    cap = int(capacity * scale_factor)
    assert all(x >= 0 for x in w), "E: negative clip weights are not supported"
    unit = 0
    for x in w:
        unit = math.gcd(unit, x)
    unit = max(unit, 1)
    if cap < 0:
        cap = -1
    return [x // unit for x in w], cap // unit if cap >= 0 else -1, (cap % unit) if cap >= 0 else 0


def knapsack(values, weights, capacity, scale_factor=5, device=None):
    """utils.py:466-510: indices of the selected clips, ascending.  Bit-exact with the reference."""
    n = len(values)
    assert len(weights) == n, "E: values and weights differ in length"
    if n == 0:
        return []
    vals = [int(v) for v in values]
    assert all(v == fv for v, fv in zip(vals, values)), "E: clip values must be integers"
    assert sum(abs(v) for v in vals) < 2 ** 31, "E: clip values overflow int32"
    w, cap_cols, cap_rem = _scale_problem(weights, capacity, scale_factor)
    if cap_cols < 0:
        return []          # K[n][capacity] is not reachable with a negative capacity in the reference either
    dev = _device(device)
    sel, _ = _knapsack_device(_to_dev(np.asarray(vals, np.int32), torch.int32, dev),
                              _to_dev(np.minimum(np.asarray(w, np.int64), 2 ** 31 - 1).astype(np.int32), torch.int32, dev),
                              n, cap_cols, cap_rem, 1, dev)
    return torch.nonzero(sel[0]).flatten().cpu().tolist()


def postprocess_array(batch_importances, skip_frames, full_n_frames, change_points, full_frames=None, device=None,
                      return_device: bool = False, bufs: SelectBuffers = None):
    """HDF5-free `postprocess` (utils.py:608-613,627-643): scores (N,1)|(N,) -> (summarized_video|None,
    uint8 indicator over the raw frames).  `change_points` is the (n_clips,2) inclusive KTS array the
    reference reads from the h5 file (utils.py:625).  Everything from the rounding on is integer work on
    the GPU and bit-exact with the reference.  `bufs` (a `SelectBuffers` for exactly this clip count and
    N_raw) makes the call allocation-free; the returned indicator is then `bufs.ind[0]`."""
    dev = _device(device if device is not None else
                  (batch_importances.device if torch.is_tensor(batch_importances) and batch_importances.is_cuda else None))
    sc = _to_dev(batch_importances, torch.float32, dev)
    if sc.dim() != 1:
        assert sc.dim() == 2 and sc.shape[-1] == 1, "E: Invalid shape for importance tensor"
        sc = sc[:, 0].contiguous()
    n = int(sc.shape[0])
    n_raw = int(full_n_frames)
    if torch.is_tensor(change_points):
        cps = _to_dev(change_points.reshape(-1, 2), torch.int32, dev)
    else:
        cps = _to_dev(np.asarray(change_points).reshape(-1, 2).astype(np.int32), torch.int32, dev)
    n_clips = int(cps.shape[0])
    lib = _lib.load()
    if bufs is not None:
        assert bufs.n_clips == n_clips and bufs.n_raw == n_raw and bufs.dev == dev, "E: selection buffers do not fit"
        ind, sel = bufs.ind[0], bufs.sel[0]
    else:
        ind = torch.zeros(n_raw, dtype=torch.uint8, device=dev)
        sel = torch.zeros(max(n_clips, 1), dtype=torch.uint8, device=dev)
    if n_clips > 0 and n > 0:
        values, lengths = _clip_information_device(sc, n, int(skip_frames), n_raw, cps, dev,
                                                   bufs.values[0] if bufs else None, bufs.lengths[0] if bufs else None)
        # weights = int(len * 5), capacity = int(int(0.15 * N_raw) * 5) (utils.py:478-479, 629): both are
        # multiples of 5, so the table compresses exactly to int(0.15 * N_raw) + 1 columns (unit = 5).
        cap = int(0.15 * n_raw)
        sel2, _ = _knapsack_device(values, lengths, n_clips, cap, 0, 1, dev, bufs)
        sel = sel2[0].contiguous()
        with torch.cuda.device(dev):
            _lib.check(lib.gn_summary_indicator(_h(dev), _ptr(cps), _ptr(sel), n_clips, n_raw, _ptr(ind), _stream(dev)),
                       "gn_summary_indicator")
    elif bufs is not None:
        ind.zero_()
    summarized_video = None
    if full_frames is not None:
        cps_host = cps.cpu().numpy()
        chosen = torch.nonzero(sel[:n_clips]).flatten().cpu().tolist()
        parts = [full_frames[int(cps_host[j][0]):int(cps_host[j][1])] for j in chosen]      # utils.py:634
        summarized_video = np.concatenate(parts, axis=0) if parts else None
    if return_device:
        return summarized_video, ind
    return summarized_video, ind.cpu().numpy()


def gt_summaries_array(annotator_importances, skip_frames, full_n_frames, change_points, device=None,
                       return_device: bool = False):
    """Ground-truth summaries of one video, as `get_dataloaders` builds them (utils.py:102-118): one
    `postprocess` per annotator row of `annotator_importances` (U, n).  The U knapsacks share the clip
    lengths and the capacity, so they run as ONE batched DP launch (a cluster per annotator) instead of U
    sequential ones.  Returns uint8 (U, N_raw), bit-exact with U calls of `postprocess_array`."""
    dev = _device(device)
    imp = _to_dev(annotator_importances, torch.float32, dev)
    assert imp.dim() == 2, "E: Invalid shape for importance tensor"
    users, n = int(imp.shape[0]), int(imp.shape[1])
    n_raw = int(full_n_frames)
    if torch.is_tensor(change_points):
        cps = _to_dev(change_points.reshape(-1, 2), torch.int32, dev)
    else:
        cps = _to_dev(np.asarray(change_points).reshape(-1, 2).astype(np.int32), torch.int32, dev)
    n_clips = int(cps.shape[0])
    ind = torch.zeros((users, n_raw), dtype=torch.uint8, device=dev)
    if users and n_clips > 0 and n > 0:
        lib = _lib.load()
        values = torch.empty((users, n_clips), dtype=torch.int32, device=dev)
        lengths = torch.empty((users, n_clips), dtype=torch.int32, device=dev)
        with torch.cuda.device(dev):
            for u in range(users):
                _lib.check(lib.gn_clip_information(_h(dev), _ptr(imp[u]), n, int(skip_frames), n_raw, _ptr(cps), n_clips,
                                                   _ptr(values[u]), _ptr(lengths[u]), _stream(dev)), "gn_clip_information")
        sel, _ = _knapsack_device(values, lengths, n_clips, int(0.15 * n_raw), 0, users, dev)
        sel = sel.contiguous()
        with torch.cuda.device(dev):
            for u in range(users):
                _lib.check(lib.gn_summary_indicator(_h(dev), _ptr(cps), _ptr(sel[u]), n_clips, n_raw, _ptr(ind[u]),
                                                    _stream(dev)), "gn_summary_indicator")
    return ind if return_device else ind.cpu().numpy()


def _read_change_points(video_id, h5_file_path, mat_file_path):
    """utils.py:615-625 (HDF5 I/O, unchanged semantics; needs h5py, which this image lacks)."""
    try:
        import h5py
    except ImportError as e:  # pragma: no cover
        raise AssertionError("E: h5py is required to read change points; use postprocess_array(...)") from e
    with h5py.File(h5_file_path, "r") as f5:
        h5 = [[vid, f5[str(vid)]["change_points"][-1][1]] for vid in f5.keys()]
    with h5py.File(mat_file_path, "r") as fm:
        def deref(refs, cast):
            return [cast(fm[r]) for ra in refs for r in ra]
        names = deref(fm["tvsum50"]["video"][:], lambda d: "".join(chr(c[0]) for c in d))
        counts = [int(c[0]) for ra in fm["tvsum50"]["nframes"][:] for r in ra for c in fm[r]]
    id_map = {}
    for vm, cm in zip(names, counts):
        for vh, ch in h5:
            if cm == ch + 1:
                id_map[vm] = vh
    with h5py.File(h5_file_path, "r") as f5:
        return f5[id_map[video_id]]["change_points"][:]


def postprocess(video_id, h5_file_path, mat_file_path, batch_importances, skip_frames, full_n_frames,
                full_frames=None):
    """utils.py:606-643, same signature; the HDF5 reads stay host I/O, the rest runs on the GPU."""
    cps = _read_change_points(video_id, h5_file_path, mat_file_path)
    return postprocess_array(batch_importances, skip_frames, full_n_frames, cps, full_frames)


# --------------------------------------------------------------------------------------------------
# (a10) evaluation  -- reference utils.py:552-604
# --------------------------------------------------------------------------------------------------
def fscore_counts_device(gd_summary_indices, predicted_summary_indices, device=None, out=None):
    """Device counts behind utils.py:571-573 as an int64 (2U+1,) CUDA tensor {overlap[u].., sum(G_u).., sum(S)};
    no host synchronisation (`out` = a preallocated tensor to fill)."""
    dev = _device(device)
    G = _to_dev(gd_summary_indices, torch.uint8, dev)
    S = _to_dev(predicted_summary_indices, torch.uint8, dev)
    assert G.dim() == 2 and G.shape[1] == S.shape[0]
    n_users, n_raw = int(G.shape[0]), int(G.shape[1])
    counts = out if out is not None else torch.empty(2 * n_users + 1, dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        _lib.check(_lib.load().gn_fscore_counts(_h(dev), _ptr(S), _ptr(G), n_users, n_raw, _ptr(counts), _stream(dev)),
                   "gn_fscore_counts")
    return counts


def fscore_counts(gd_summary_indices, predicted_summary_indices, device=None):
    """Device counts behind utils.py:571-573: (overlap[u], sum(G_u)[u], sum(S)) as Python ints."""
    counts = fscore_counts_device(gd_summary_indices, predicted_summary_indices, device)
    n_users = (counts.numel() - 1) // 2
    c = counts.cpu().tolist()
    return c[:n_users], c[n_users:2 * n_users], c[2 * n_users]


def fscore_from_counts(overlap, gsum, ssum):
    """The float64 tail of utils.py:571-580 on integer counts: (mean, max) F-score over the annotators."""
    f_scores = []
    for ov, sg in zip(overlap, gsum):
        precision = ov / ssum if ssum != 0 else 0
        recall = ov / sg if sg != 0 else 0
        f_scores.append(2 * precision * recall / (precision + recall) if (precision + recall) != 0 else 0)
    acc = 0.0
    for f in f_scores:          # plain left-to-right adds, like sum() over np.float64 scalars (utils.py:580)
        acc = acc + f
    return acc / len(f_scores), max(f_scores)


def get_fscore(gd_summary_indices, predicted_summary_indices, device=None):
    """utils.py:552-580: (mean, max) F-score over the annotators.  Counts are integer sums on the GPU
    (NumPy-1 semantics); the float64 tail follows the reference's operation order."""
    overlap, gsum, ssum = fscore_counts(gd_summary_indices, predicted_summary_indices, device)
    return fscore_from_counts(overlap, gsum, ssum)


def postprocess_and_get_fscores_array(batch_predictions, full_n_batch_frames, gd_summarized_video_frame_indices,
                                      skip_frames, change_points, device=None):
    """Array twin of utils.py:587-604."""
    _, ind = postprocess_array(batch_predictions, skip_frames, full_n_batch_frames, change_points, device=device,
                               return_device=True)
    return get_fscore(gd_summarized_video_frame_indices, ind, device=ind.device)


def postprocess_and_get_fscores(video_id, batch_predictions, full_n_batch_frames, gd_summarized_video_frame_indices,
                                h5_file_path, mat_file_path, skip_frames):
    """utils.py:587-604, same signature."""
    cps = _read_change_points(video_id, h5_file_path, mat_file_path)
    return postprocess_and_get_fscores_array(batch_predictions, full_n_batch_frames,
                                             gd_summarized_video_frame_indices, skip_frames, cps)


# --------------------------------------------------------------------------------------------------
# dataset container  -- reference utils.py:16-76 (tuple order only; TSV titles are optional here)
# --------------------------------------------------------------------------------------------------
class dataloader(torch.utils.data.Dataset):
    def __init__(self, fps, frames, full_n_frames, audios, labels=None, gd_summarized_video_frame_indices=None,
                 titles=None):
        self.fps = fps
        self.video_ids = [fp.split("/")[-1].split(".")[0] for fp in fps]
        self.frames = [torch.as_tensor(f, dtype=torch.float32) for f in frames]
        self.full_n_frames = full_n_frames
        self.audios = audios if audios[0][0] is None else [torch.as_tensor(a, dtype=torch.float32) for a in audios]
        if labels is None:
            self.labels = self.gd_summarized_video_frame_indices = [None for _ in range(len(frames))]
            assert len(self.frames) == len(self.audios), "E: Inconsistency in data loader definition"
        else:
            self.labels = [torch.as_tensor(l, dtype=torch.float32) for l in labels]
            self.gd_summarized_video_frame_indices = gd_summarized_video_frame_indices
            assert len(self.frames) == len(self.audios) == len(self.labels), "E: Inconsistency in data loader definition"
        self.titles = titles if titles is not None else list(self.video_ids)
        self.N = len(self.frames)

    def __len__(self):
        return self.N

    def __getitem__(self, video_idx):
        self.title = self.titles[video_idx]
        self.full_n_frames_ = self.full_n_frames[video_idx]
        return (self.video_ids[video_idx], self.frames[video_idx], self.audios[video_idx], self.labels[video_idx],
                self.gd_summarized_video_frame_indices[video_idx])
