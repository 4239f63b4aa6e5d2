"""CPU oracle: a restatement of CVML-GoalNet's per-frame scoring path.

TEST INFRASTRUCTURE ONLY.  Only `tests/`, `__graft_entry__.smoke()` and the
`cpu_baseline` / `--impl reference` legs of `bench.py` may import this module, and
only as the checker or the timed CPU baseline -- never as the product path.  The
product package (`cvml_goalnet_b200`) must not import anything from `oracle/`.

Every function cites the reference file:line it restates (paths are relative to
the reference repository root).  NumPy / torch-CPU fp32, eval-mode semantics
(SURVEY.md quirk Q4).

PARITY PINNING.  The reference ships no tests, fixtures or weights, so its own
code *run in the build container* is what pins this oracle: `oracle/gen_golden.py`
imports the unmodified `/root/reference/utils.py` (four absent I/O modules
stubbed, see `oracle/ref_import.py`), runs `AVM.forward`, the arithmetic of
`extract_condensed_frame_tensor`, `expand_array`, `get_clip_information`,
`knapsack`, `get_fscore` (under NumPy-1 integer semantics, quirk Q6) and the
literal `interp1d` loop on seeded inputs, and commits the outputs under
`tests/golden/`.  `tests/test_oracle_golden.py` checks every function below against
those vectors.  ONE stage stays "parity unpinned": `librosa.feature.mfcc`
(utils.py:333).  librosa is a third-party dependency that is not vendored, not
version-pinned (no requirements file; the report dates it to librosa 0.10.x) and
absent from this image, so `mfcc()` below restates librosa 0.10's published
algorithm and *is* the definition for that stage.
"""
from __future__ import annotations

import numpy as np

# ----------------------------------------------------------------------------------------------
# (a1) frame preprocessing -- reference utils.py:274-292
# ----------------------------------------------------------------------------------------------


def _bilinear_taps(n_src: int, n_dst: int):
    """cv2.resize INTER_LINEAR source taps for one axis (half-pixel centres, no anti-alias).

    Restates what `cv2.resize(image, (40, 40))` at utils.py:285 does per axis:
    src = (dst + 0.5) * (1 / (n_dst / n_src)) - 0.5 in double; i0 = floor(src); frac = float(src - i0);
    clamp: i0 < 0 -> (0, frac 0); i0 >= n_src - 1 -> (n_src - 1, frac 0).
    (Probed against cv2 4.13 in the build container: the fraction is formed in double and only then
    narrowed to float; forming it in float is off by up to 3e-6.)
    """
    scale = 1.0 / (np.float64(n_dst) / np.float64(n_src))
    d = np.arange(n_dst, dtype=np.float64)
    src = (d + 0.5) * scale - 0.5
    i0 = np.floor(src).astype(np.int64)
    frac = (src - i0).astype(np.float32)
    lo = i0 < 0
    i0[lo] = 0
    frac[lo] = 0.0
    hi = i0 >= n_src - 1
    i0[hi] = n_src - 1
    frac[hi] = 0.0
    i1 = np.minimum(i0 + 1, n_src - 1)
    return i0, i1, frac


def preprocess_frame(img_u8: np.ndarray, S: int = 40) -> np.ndarray:
    """One frame: global min-max normalise (utils.py:284) then bilinear S x S (utils.py:285).

    Closed form: bilinear interpolation commutes with the affine normalisation, so sample the
    uint8 image and normalise the S*S*3 samples.  Returns float32 (S,S,3), channel order kept
    (BGR, quirk Q8).
    """
    H, W, _ = img_u8.shape
    mn = np.float64(img_u8.min())
    mx = np.float64(img_u8.max())
    y0, y1, fy = _bilinear_taps(H, S)
    x0, x1, fx = _bilinear_taps(W, S)
    img = ((img_u8.astype(np.float64) - mn) / (mx - mn + 1e-7)).astype(np.float32)
    fy = fy[:, None, None]
    fx = fx[None, :, None]
    top = img[y0][:, x0] * (1 - fx) + img[y0][:, x1] * fx
    bot = img[y1][:, x0] * (1 - fx) + img[y1][:, x1] * fx
    return (top * (1 - fy) + bot * fy).astype(np.float32)


def preprocess_frame_literal(img_u8: np.ndarray, S: int = 40) -> np.ndarray:
    """utils.py:284-285 executed literally through OpenCV (cv2 is in the image on both boxes)."""
    import cv2
    image = ((img_u8 - img_u8.min()) / (img_u8.max() - img_u8.min() + 1e-7)).astype(np.float32)
    return cv2.resize(image, (S, S))


def extract_condensed_frame_tensor_array(frames_u8: np.ndarray, S: int = 40, literal: bool = False) -> np.ndarray:
    """Array twin of utils.py:274-292 for already-decoded, already-subsampled frames.

    frames_u8: (N,H,W,3) uint8 BGR.  Returns float32 (N,3,S,S) (the transpose at utils.py:292).
    """
    fn = preprocess_frame_literal if literal else preprocess_frame
    out = np.stack([fn(f, S) for f in frames_u8]) if len(frames_u8) else np.zeros((0, S, S, 3), np.float32)
    return np.ascontiguousarray(np.transpose(out, (0, 3, 1, 2)))


# ----------------------------------------------------------------------------------------------
# (a2) audio features -- reference utils.py:313-349; librosa.feature.mfcc restated (0.10 defaults)
# ----------------------------------------------------------------------------------------------

N_FFT = 2048
HOP = 512
N_MELS = 128
N_MFCC = 30


def _hz_to_mel(f):
    f = np.asanyarray(f, dtype=np.float64)
    f_sp = 200.0 / 3
    mels = f / f_sp
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    return np.where(f >= min_log_hz, min_log_mel + np.log(np.maximum(f, 1e-300) / min_log_hz) / logstep, mels)


def _mel_to_hz(m):
    m = np.asanyarray(m, dtype=np.float64)
    f_sp = 200.0 / 3
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    return np.where(m >= min_log_mel, min_log_hz * np.exp(logstep * (m - min_log_mel)), f_sp * m)


def mel_filterbank(sr: int, n_fft: int = N_FFT, n_mels: int = N_MELS) -> np.ndarray:
    """librosa.filters.mel(sr, n_fft, n_mels=128, fmin=0, fmax=sr/2, htk=False, norm='slaney') -> float32."""
    fftfreqs = np.fft.rfftfreq(n=n_fft, d=1.0 / sr)
    mel_f = _mel_to_hz(np.linspace(_hz_to_mel(0.0), _hz_to_mel(sr / 2.0), n_mels + 2))
    fdiff = np.diff(mel_f)
    ramps = mel_f[:, None] - fftfreqs[None, :]
    weights = np.zeros((n_mels, 1 + n_fft // 2), dtype=np.float64)
    for i in range(n_mels):
        lower = -ramps[i] / fdiff[i]
        upper = ramps[i + 2] / fdiff[i + 1]
        weights[i] = np.maximum(0, np.minimum(lower, upper))
    enorm = 2.0 / (mel_f[2:n_mels + 2] - mel_f[:n_mels])
    weights *= enorm[:, None]
    return weights.astype(np.float32)


def hann_periodic(n: int = N_FFT) -> np.ndarray:
    """scipy.signal.get_window('hann', n, fftbins=True) (float64)."""
    return 0.5 - 0.5 * np.cos(2.0 * np.pi * np.arange(n) / n)


_MEL_CACHE = {}


def mfcc(y: np.ndarray, sr: int, n_mfcc: int = N_MFCC) -> np.ndarray:
    """Restatement of `librosa.feature.mfcc(y=y, sr=sr, n_mfcc=30)` as called at utils.py:333.

    librosa 0.10 defaults: STFT n_fft=2048 hop=512 periodic hann, center=True with zero padding
    (pad_mode='constant'), power 2; mel 128 Slaney; power_to_db(ref=1, amin=1e-10, top_db=80)
    with the max taken over the whole (128,q) matrix of THIS segment; DCT-II ortho over the mel
    axis, first n_mfcc rows.  STFT is evaluated in float64 (float64 window x float32 signal) and
    stored as complex64; everything after is float32, as in librosa.  Returns float32 (n_mfcc, q),
    q = 1 + len(y)//512.
    """
    from scipy.fftpack import dct
    y = np.asarray(y, dtype=np.float32)
    q = 1 + len(y) // HOP
    ypad = np.concatenate([np.zeros(N_FFT // 2, np.float32), y, np.zeros(N_FFT // 2, np.float32)])
    idx = np.arange(N_FFT)[None, :] + HOP * np.arange(q)[:, None]
    frames = ypad[idx]                                     # (q, 2048) float32
    spec = np.fft.rfft(hann_periodic()[None, :] * frames, axis=1).astype(np.complex64)  # (q,1025)
    power = (np.abs(spec) ** 2).astype(np.float32)          # float32
    key = (int(sr),)
    if key not in _MEL_CACHE:
        _MEL_CACHE[key] = mel_filterbank(int(sr))
    mel = np.einsum("tf,mf->mt", power, _MEL_CACHE[key]).astype(np.float32)  # (128,q)
    log_spec = (10.0 * np.log10(np.maximum(np.float32(1e-10), mel))).astype(np.float32)
    log_spec = np.maximum(log_spec, log_spec.max() - np.float32(80.0))
    return dct(log_spec, axis=0, type=2, norm="ortho")[:n_mfcc].astype(np.float32)


def segment_bounds(n_samples: int, n_frames: int):
    """utils.py:322-330: start=round(i*spf), end=min(round(start+spf), len(y)) (Python half-even)."""
    spf = n_samples / n_frames
    out = []
    for i in range(n_frames):
        s = round(i * spf)
        e = round(s + spf)
        if e > n_samples:
            e = n_samples
        out.append((s, e))
    return out


def cubic_resample_matrix(q: int, B: int) -> np.ndarray:
    """The linear operator of utils.py:336-343: not-a-knot cubic interp1d on grid 0..q-1 evaluated
    at linspace(0, q-1, B).  Obtained by pushing the identity through scipy.interpolate.interp1d."""
    from scipy.interpolate import interp1d
    f = interp1d(np.arange(q), np.eye(q), kind="cubic", axis=0, fill_value="extrapolate")
    return np.ascontiguousarray(f(np.linspace(0, q - 1, B)))  # (B, q) float64


def extract_audio_features_array(y: np.ndarray, sr: int, n_frames: int, bin_length: int) -> np.ndarray:
    """Array twin of utils.py:313-349 (audio already loaded: `y, sr` replace librosa.load)."""
    from scipy.interpolate import interp1d
    B = bin_length
    feats = []
    for (s, e) in segment_bounds(len(y), n_frames):
        m = mfcc(y[s:e], sr)
        rows = []
        for f_idx in range(m.shape[0]):  # literal loop of utils.py:336-343
            interpolator = interp1d(np.arange(m.shape[1]), m[f_idx, :], kind="cubic", fill_value="extrapolate")
            rows.append(interpolator(np.linspace(0, m.shape[1] - 1, B)))
        rows = np.array(rows)
        assert rows.shape[-1] == B, "E: Shape mismatch"
        feats.append(rows)
    return np.array(feats)  # (N, 30, B) float64


# ----------------------------------------------------------------------------------------------
# (a3-a5) model forward, eval mode -- reference utils.py:145-272
# ----------------------------------------------------------------------------------------------


def avm_forward(sd: dict, audio, visual, audio_included: bool = True, return_features: bool = False,
                classifier: bool = False):
    """Functional fp32 restatement of AVM.forward (utils.py:260-272) in eval mode.

    sd: state_dict-keyed tensors/arrays.  audio (N,30,B), visual (N,3,S,S).  Dropout is identity
    and BatchNorm uses running statistics (eps 1e-5) -- quirk Q4.
    """
    import torch
    import torch.nn.functional as F
    t = lambda k: torch.as_tensor(np.asarray(sd[k]) if not torch.is_tensor(sd[k]) else sd[k]).float()
    with torch.no_grad():
        x = torch.as_tensor(visual).float()
        feats = {}
        for i, (stride, pad) in zip((1, 2, 3), ((3, 3), (1, 1), (1, 1))):     # utils.py:174-187
            x = F.conv2d(x, t(f"visbl.conv{i}.weight"), t(f"visbl.conv{i}.bias"), stride=stride, padding=pad)
            x = F.max_pool2d(F.relu(x), kernel_size=3, stride=1)
            x = F.batch_norm(x, t(f"visbl.bnorm{i}.running_mean"), t(f"visbl.bnorm{i}.running_var"),
                             t(f"visbl.bnorm{i}.weight"), t(f"visbl.bnorm{i}.bias"), False, 0.1, 1e-5)
            feats[f"vis{i}"] = x
        x = x.flatten(1)                                                          # utils.py:189 (C,H,W order)
        vis = F.relu(F.linear(x, t("visbl.linear5.weight"), t("visbl.linear5.bias")))  # utils.py:191-192
        feats["visual_features"] = vis
        if audio_included:
            a = torch.as_tensor(audio).float()
            a = F.relu(F.conv1d(a, t("audbl.conv1.weight"), t("audbl.conv1.bias"), stride=2, padding=1))
            a = F.relu(F.conv1d(a, t("audbl.conv2.weight"), t("audbl.conv2.bias"), stride=2, padding=1))
            a = F.relu(F.linear(a.flatten(1), t("audbl.linear3.weight"), t("audbl.linear3.bias")))
            feats["audio_features"] = a
            h = torch.cat((a, vis), dim=-1)                                       # utils.py:266, audio first
        else:
            h = vis
        for idx in (0, 3, 6, 9):                                                  # utils.py:242-254
            h = F.relu(F.linear(h, t(f"fusion.{idx}.weight"), t(f"fusion.{idx}.bias")))
        if classifier:
            # the reference's commented classifier variant: utils.py:257 `# nn.Softmax(dim = 1),` in place of
            # nn.Sigmoid() with a 5-wide fusion.12; no 4y+1 (the prediction is argmax + 1, main.py:97)
            out = torch.softmax(F.linear(h, t("fusion.12.weight"), t("fusion.12.bias")), dim=1)
        else:
            h = torch.sigmoid(F.linear(h, t("fusion.12.weight"), t("fusion.12.bias")))
            out = 4 * h + 1                                                       # utils.py:270
    if return_features:
        return out, feats
    return out


# ----------------------------------------------------------------------------------------------
# (a6-a9) post-processing: integers from here on -- reference utils.py:396-410, 445-510, 606-643
# ----------------------------------------------------------------------------------------------


def round_importances(batch_importances) -> list:
    """utils.py:608-611: squeeze (N,1)->(N,), torch.round (half-to-even) -> int8 -> list."""
    import torch
    b = torch.as_tensor(np.asarray(batch_importances))
    if len(b.shape) != 1:
        assert len(b.shape) == 2 and b.shape[-1] == 1, "E: Invalid shape for importance tensor"
        b = b[:, 0]
    return torch.round(b).type(torch.int8).tolist()


def expand_array(arr, expansion_rate, length):
    """utils.py:396-410."""
    if len(arr) == length:
        return arr
    expanded = []
    for el in arr:
        expanded += [el] * expansion_rate
    if len(expanded) >= length:
        expanded = expanded[:length]
    else:
        expanded += [expanded[-1]] * (length - len(expanded))
    return expanded


def get_clip_information(clip_intervals, importances):
    """utils.py:445-464: value = sum(imp[a:b]), length = len(imp[a:b]) (end-EXCLUSIVE, quirk Q7)."""
    vals, lens = [], []
    for ci in clip_intervals:
        sl = importances[int(ci[0]):int(ci[1])]
        vals.append(int(sum(sl)))
        lens.append(len(sl))
    return vals, lens, clip_intervals


def knapsack(values, weights, capacity, scale_factor=5):
    """utils.py:466-510, literal pure-Python table (small cases only)."""
    weights = [int(w * scale_factor) for w in weights]
    capacity = int(capacity * scale_factor)
    n = len(values)
    K = [[0] * (capacity + 1) for _ in range(n + 1)]
    for i in range(1, n + 1):
        wt, v = weights[i - 1], values[i - 1]
        Kp, Kc = K[i - 1], K[i]
        for w in range(1, capacity + 1):
            if wt <= w:
                a = v + Kp[w - wt]
                Kc[w] = a if a > Kp[w] else Kp[w]
            else:
                Kc[w] = Kp[w]
    res = K[n][capacity]
    w = capacity
    sel = []
    for i in range(n, 0, -1):
        if res <= 0:
            break
        if res == K[i - 1][w]:
            continue
        sel.append(i - 1)
        res -= values[i - 1]
        w -= weights[i - 1]
    sel.reverse()
    return sel


def knapsack_rows(values, weights, capacity, scale_factor=5):
    """Same table and backtrack as `knapsack` (utils.py:466-510), with each table row evaluated as one
    NumPy vector expression so that full-size matches finish in seconds.  The FULL (n+1) x (5*cap+1)
    table is kept -- no compression -- so this is the reference algorithm, only vectorised."""
    weights = [int(w * scale_factor) for w in weights]
    capacity = int(capacity * scale_factor)
    n = len(values)
    K = np.zeros((n + 1, capacity + 1), dtype=np.int64)
    for i in range(1, n + 1):
        wt, v = weights[i - 1], int(values[i - 1])
        K[i] = K[i - 1]
        if wt <= capacity:
            lo = max(wt, 1)                       # column 0 is forced to 0 (utils.py:487-488, quirk Q11)
            cand = K[i - 1, lo - wt:capacity + 1 - wt] + v
            np.maximum(K[i, lo:], cand, out=K[i, lo:])
    res = int(K[n, capacity])
    w = capacity
    sel = []
    for i in range(n, 0, -1):
        if res <= 0:
            break
        if res == K[i - 1, w]:
            continue
        sel.append(i - 1)
        res -= int(values[i - 1])
        w -= weights[i - 1]
    sel.reverse()
    return sel


def postprocess_array(batch_importances, skip_frames, full_n_frames, change_points, fast=True):
    """HDF5-free restatement of utils.py:606-643 (change_points passed in; quirk Q10).

    Returns (selected clip indices, uint8 indicator over raw frames marking a..b INCLUSIVE).
    """
    imp = round_importances(batch_importances)                                    # :608-611
    expanded = expand_array(arr=imp, expansion_rate=skip_frames, length=full_n_frames)  # :613
    clip_intervals = [[int(a), int(b)] for a, b in np.asarray(change_points).reshape(-1, 2)]
    vals, lens, clip_intervals = get_clip_information(clip_intervals, expanded)    # :627
    cap = int(0.15 * full_n_frames)                                               # :629
    sel = (knapsack_rows if fast else knapsack)(vals, lens, cap)                  # :631
    ind = np.zeros(shape=(full_n_frames,), dtype=np.uint8)                        # :638
    for ci in (clip_intervals[j] for j in sel):
        for f in range(ci[0], ci[1] + 1):                                         # :640 inclusive
            ind[f] = 1
    return sel, ind


# ----------------------------------------------------------------------------------------------
# (a10) evaluation -- reference utils.py:552-580
# ----------------------------------------------------------------------------------------------


def fscore_counts(gd: np.ndarray, pred: np.ndarray):
    """Integer counts behind utils.py:571-573 with int64 sums (NumPy-1 semantics, quirk Q6):
    per user (|S and G_u|, |G_u|) and |S|."""
    S = np.asarray(pred)
    G = np.asarray(gd)
    assert G.shape[1] == len(S)
    overlap = np.logical_and(S[None, :], G).sum(axis=1, dtype=np.int64)
    return overlap, G.sum(axis=1, dtype=np.int64), int(S.sum(dtype=np.int64))


def get_fscore(gd_summary_indices: np.ndarray, predicted_summary_indices: np.ndarray):
    """utils.py:552-580 with int64 counts; float64 tail in the reference's operation order."""
    overlap, gsum, ssum = fscore_counts(gd_summary_indices, predicted_summary_indices)
    f_scores = []
    for u in range(gd_summary_indices.shape[0]):
        ov, sg = int(overlap[u]), int(gsum[u])
        precision = ov / ssum if ssum != 0 else 0
        recall = ov / sg if sg != 0 else 0
        f = 2 * precision * recall / (precision + recall) if (precision + recall) != 0 else 0
        f_scores.append(f)
    # utils.py:580 `sum(f_scores)`: the reference adds np.float64 scalars one by one (plain IEEE adds).
    # Python >= 3.12's sum() would use compensated summation on exact floats, so spell the loop out.
    acc = 0.0
    for f in f_scores:
        acc = acc + f
    return acc / len(f_scores), max(f_scores)


def postprocess_and_get_fscores_array(batch_predictions, full_n_batch_frames, gd, skip_frames, change_points):
    """utils.py:587-604 with change_points passed in."""
    _, ind = postprocess_array(batch_predictions, skip_frames, full_n_batch_frames, change_points)
    return get_fscore(gd, ind)


# ----------------------------------------------------------------------------------------------
# whole path (the CPU baseline that bench.py times)
# ----------------------------------------------------------------------------------------------


def score_and_summarize(sd, frames_u8, y, sr, skip_frames, full_n_frames, change_points, gd=None,
                        B: int = 30, S: int = 40):
    """frames + audio -> scores -> keyshot indicator (-> F-scores): main.py:315-345 without file I/O."""
    vis = extract_condensed_frame_tensor_array(frames_u8, S)
    aud = extract_audio_features_array(y, sr, len(frames_u8), B)
    scores = avm_forward(sd, aud.astype(np.float32), vis).numpy()
    sel, ind = postprocess_array(scores, skip_frames, full_n_frames, change_points)
    fs = get_fscore(gd, ind) if gd is not None else None
    return scores, sel, ind, fs


# ------------------------------------------------------------------------------------------------
# fusion-network training step (BASELINE config 4) -- restates main.py:68-70,187-193 for AVM.fusion
# ------------------------------------------------------------------------------------------------
FUSION_KEYS = ["fusion.0", "fusion.3", "fusion.6", "fusion.9", "fusion.12"]


def fusion_flat(sd: dict) -> np.ndarray:
    """Parameters of AVM.fusion as one float32 vector in `model.fusion.parameters()` order (utils.py:242-258)."""
    parts = []
    for k in FUSION_KEYS:
        parts.append(np.asarray(sd[k + ".weight"], np.float32).ravel())
        parts.append(np.asarray(sd[k + ".bias"], np.float32).ravel())
    return np.concatenate(parts)


def fusion_train_steps(sd: dict, features, labels, steps: int = 1, fixed_loss: bool = False, masks=None,
                       dropout_p: float = 0.2, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                       classifier: bool = False):
    """`steps` iterations of  optimizer.zero_grad(); pred = fusion(x); loss = MSELoss()(pred, labels);
    loss.backward(); optimizer.step()  (main.py:187-193) with optim.Adam(lr) (main.py:70) in float32 on the CPU.
    As written pred is (N,1) and labels (N,), so MSELoss broadcasts to N x N (quirk Q5); fixed_loss
    compares element-wise.  Dropout(0.2) (utils.py:245-254): `masks` = per step a list of four uint8
    keep-masks applied as x * mask / (1 - p); None = dropout off (eval-mode forward).
    classifier=True is the reference's commented classifier variant: fusion.12 = Linear(128,5) followed by
    nn.Softmax(dim=1) (utils.py:257) and `criterion = nn.CrossEntropyLoss()` (main.py:69) called as
    `criterion(batch_predictions, (batch_labels-1).long())` (main.py:96), i.e. ON THE SOFTMAX OUTPUT as written;
    fixed_loss then applies the criterion to the logits instead.  Predictions are the (N,5) probabilities.
    Returns (losses [steps], grads of the FIRST step (flat), predictions of the first step, final flat params)."""
    import torch
    import warnings
    ws = [torch.tensor(np.asarray(sd[k + ".weight"], np.float32), requires_grad=True) for k in FUSION_KEYS]
    bs = [torch.tensor(np.asarray(sd[k + ".bias"], np.float32), requires_grad=True) for k in FUSION_KEYS]
    params = [p for pair in zip(ws, bs) for p in pair]
    opt = torch.optim.Adam(params, lr=lr, betas=betas, eps=eps)
    crit = torch.nn.CrossEntropyLoss() if classifier else torch.nn.MSELoss()
    x0 = torch.as_tensor(np.asarray(features, np.float32))
    lab = torch.as_tensor(np.asarray(labels, np.float32))
    losses, grads0, pred0 = [], None, None
    for s in range(steps):
        opt.zero_grad()
        h = x0
        for k in range(4):
            h = torch.relu(h @ ws[k].T + bs[k])
            if masks is not None:
                h = h * torch.as_tensor(np.asarray(masks[s][k], np.float32)) / (1.0 - dropout_p)
        if classifier:
            logits = h @ ws[4].T + bs[4]
            pred = torch.softmax(logits, dim=1)                      # nn.Softmax(dim = 1), utils.py:257
            loss = crit(logits if fixed_loss else pred, (lab - 1).long())     # main.py:96
        else:
            pred = 4 * torch.sigmoid(h @ ws[4].T + bs[4]) + 1        # (N,1), utils.py:256,270
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")                      # the reference's broadcasting UserWarning
                loss = crit(pred, lab) if not fixed_loss else crit(pred[:, 0], lab)
        loss.backward()
        if s == 0:
            grads0 = np.concatenate([p.grad.detach().numpy().ravel() for p in params])
            pred0 = pred.detach().numpy().copy() if classifier else pred.detach().numpy()[:, 0].copy()
        opt.step()
        losses.append(float(loss.item()))
    flat = np.concatenate([p.detach().numpy().ravel() for p in params])
    return np.array(losses), grads0, pred0, flat


# ------------------------------------------------------------------------------------------------
# full-AVM training sub-step -- restates main.py:187-193 for the whole network in TRAIN mode
# ------------------------------------------------------------------------------------------------
AVM_PARAM_KEYS = ["visbl.conv1", "visbl.bnorm1", "visbl.conv2", "visbl.bnorm2", "visbl.conv3", "visbl.bnorm3",
                  "visbl.linear5", "audbl.conv1", "audbl.conv2", "audbl.linear3",
                  "fusion.0", "fusion.3", "fusion.6", "fusion.9", "fusion.12"]


def avm_train_steps(sd: dict, audio, visual, labels, steps: int = 1, audio_included: bool = True, masks=None,
                    dropout_p: float = 0.2, fixed_loss: bool = False, lr: float = 1e-3, betas=(0.9, 0.999),
                    eps: float = 1e-8):
    """`steps` iterations of the reference's train sub-step (main.py:187-193: zero_grad, forward, MSELoss, backward,
    Adam step; optim.Adam(lr) of main.py:70) on ONE sub-batch, float32 on the CPU, model in train mode as the
    reference runs it (quirk Q4): BatchNorm2d normalises with the sub-batch statistics and updates its running
    buffers (momentum 0.1), Dropout(0.2) follows linear5 and fusion.0/3/6/9 (utils.py:168,245-254).
    `masks`: per step a list of five uint8 keep-masks ([n,512] x3, [n,256], [n,128]) applied as x * mask / (1-p);
    None = dropout off.  As written the loss broadcasts (N,1) against (N,) to N x N (quirk Q5).
    Returns (losses, scores of the first step, gradients of the first step {key: array}, final state_dict)."""
    import torch
    import torch.nn.functional as F
    import warnings
    P, buf = {}, {}
    for k in AVM_PARAM_KEYS:
        if k.startswith("audbl") and not audio_included:
            continue
        P[k + ".weight"] = torch.tensor(np.asarray(sd[k + ".weight"], np.float32), requires_grad=True)
        P[k + ".bias"] = torch.tensor(np.asarray(sd[k + ".bias"], np.float32), requires_grad=True)
    for i in (1, 2, 3):
        buf[f"visbl.bnorm{i}.running_mean"] = torch.tensor(np.asarray(sd[f"visbl.bnorm{i}.running_mean"], np.float32))
        buf[f"visbl.bnorm{i}.running_var"] = torch.tensor(np.asarray(sd[f"visbl.bnorm{i}.running_var"], np.float32))
    opt = torch.optim.Adam(list(P.values()), lr=lr, betas=betas, eps=eps)
    crit = torch.nn.MSELoss()
    vis = torch.as_tensor(np.asarray(visual, np.float32))
    aud = torch.as_tensor(np.asarray(audio, np.float32)) if audio_included else None
    lab = torch.as_tensor(np.asarray(labels, np.float32))
    losses, scores0, grads0 = [], None, None

    def drop(h, s, j):
        if masks is None:
            return h
        return h * torch.as_tensor(np.asarray(masks[s][j], np.float32)) / (1.0 - dropout_p)

    for s in range(steps):
        opt.zero_grad()
        x = vis
        for i, (stride, pad) in zip((1, 2, 3), ((3, 3), (1, 1), (1, 1))):           # utils.py:174-187
            x = F.conv2d(x, P[f"visbl.conv{i}.weight"], P[f"visbl.conv{i}.bias"], stride=stride, padding=pad)
            x = F.max_pool2d(F.relu(x), kernel_size=3, stride=1)
            x = F.batch_norm(x, buf[f"visbl.bnorm{i}.running_mean"], buf[f"visbl.bnorm{i}.running_var"],
                             P[f"visbl.bnorm{i}.weight"], P[f"visbl.bnorm{i}.bias"], True, 0.1, 1e-5)
        v = drop(F.relu(F.linear(x.flatten(1), P["visbl.linear5.weight"], P["visbl.linear5.bias"])), s, 0)
        if audio_included:
            a = F.relu(F.conv1d(aud, P["audbl.conv1.weight"], P["audbl.conv1.bias"], stride=2, padding=1))
            a = F.relu(F.conv1d(a, P["audbl.conv2.weight"], P["audbl.conv2.bias"], stride=2, padding=1))
            a = F.relu(F.linear(a.flatten(1), P["audbl.linear3.weight"], P["audbl.linear3.bias"]))
            h = torch.cat((a, v), dim=-1)
        else:
            h = v
        for j, idx in enumerate((0, 3, 6, 9)):
            h = drop(F.relu(F.linear(h, P[f"fusion.{idx}.weight"], P[f"fusion.{idx}.bias"])), s, j + 1)
        pred = 4 * torch.sigmoid(F.linear(h, P["fusion.12.weight"], P["fusion.12.bias"])) + 1
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            loss = crit(pred, lab) if not fixed_loss else crit(pred[:, 0], lab)
        loss.backward()
        if s == 0:
            scores0 = pred.detach().numpy()[:, 0].copy()
            grads0 = {k: p.grad.detach().numpy().copy() for k, p in P.items()}
        opt.step()
        losses.append(float(loss.item()))
    out = {k: p.detach().numpy().copy() for k, p in P.items()}
    out.update({k: b.numpy().copy() for k, b in buf.items()})
    return np.array(losses), scores0, grads0, out
