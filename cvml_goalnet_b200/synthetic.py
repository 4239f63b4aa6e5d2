"""Deterministic synthetic workloads for the per-frame scoring path.

These are workload *definitions*, shared by the parity tests, `bench.py` and
`__graft_entry__.smoke()`; they depend only on NumPy so that the same seed gives
the same bytes in the build container and on the GPU box.

Seed base 12344321 is the reference's own (`main.py:53`).
"""
from __future__ import annotations

import numpy as np

SEED_BASE = 12344321

# (name, shape-builder) of every tensor in `AVM(audio_included).state_dict()`
# (reference utils.py:145-258; lazies materialised at S x S frames, B audio bins).


def conv_out(n: int, k: int, s: int, p: int) -> int:
    return (n + 2 * p - k) // s + 1


def visbl_dims(S: int = 40):
    """Spatial sizes through VisBl (utils.py:151-164): conv1 k3 s3 p3, pools 3/1, convs k3 s1 p1."""
    c1 = conv_out(S, 3, 3, 3)
    p1 = c1 - 2
    c2 = p1
    p2 = c2 - 2
    c3 = p2
    p3 = c3 - 2
    return dict(c1=c1, p1=p1, c2=c2, p2=p2, c3=c3, p3=p3, flat=512 * p3 * p3)


def audbl_dims(B: int = 30):
    """Lengths through AudBl (utils.py:203-211): two Conv1d k3 s2 p1."""
    l1 = conv_out(B, 3, 2, 1)
    l2 = conv_out(l1, 3, 2, 1)
    return dict(l1=l1, l2=l2, flat=128 * l2)


def state_dict_shapes(audio_included: bool = True, S: int = 40, B: int = 30):
    vd, ad = visbl_dims(S), audbl_dims(B)
    shapes = {
        "visbl.conv1.weight": (64, 3, 3, 3), "visbl.conv1.bias": (64,),
        "visbl.conv2.weight": (256, 64, 3, 3), "visbl.conv2.bias": (256,),
        "visbl.conv3.weight": (512, 256, 3, 3), "visbl.conv3.bias": (512,),
        "visbl.linear5.weight": (512, vd["flat"]), "visbl.linear5.bias": (512,),
    }
    for i, c in ((1, 64), (2, 256), (3, 512)):
        for k in ("weight", "bias", "running_mean", "running_var"):
            shapes[f"visbl.bnorm{i}.{k}"] = (c,)
        shapes[f"visbl.bnorm{i}.num_batches_tracked"] = ()
    if audio_included:
        shapes.update({
            "audbl.conv1.weight": (64, 30, 3), "audbl.conv1.bias": (64,),
            "audbl.conv2.weight": (128, 64, 3), "audbl.conv2.bias": (128,),
            "audbl.linear3.weight": (128, ad["flat"]), "audbl.linear3.bias": (128,),
        })
    fin = 512 + (128 if audio_included else 0)
    for idx, (o, i) in zip((0, 3, 6, 9, 12), ((512, fin), (512, 512), (256, 512), (128, 256), (1, 128))):
        shapes[f"fusion.{idx}.weight"] = (o, i)
        shapes[f"fusion.{idx}.bias"] = (o,)
    return shapes


def make_weights(seed: int = 0, audio_included: bool = True, S: int = 40, B: int = 30,
                 score_gain: float = 1.0) -> dict:
    """Seeded AVM weights as a dict of float32 NumPy arrays keyed like the state_dict.

    Conv/linear tensors follow PyTorch's default init law, U(-1/sqrt(fan_in), 1/sqrt(fan_in));
    BatchNorm gets gamma~U(.5,1.5), beta~N(0,.1^2), running_mean~N(0,.1^2), running_var~U(.5,1.5)
    so that it is not an identity.  `score_gain` != 1 replaces the last layer by
    (w - mean(w)) * gain with zero bias: a default-init AVM emits ~3.05 for every frame, the gain
    variant spreads scores across a rounding boundary so that parity tests are not vacuous.
    """
    rng = np.random.default_rng(SEED_BASE + 7919 * seed)
    out = {}
    for name, shape in state_dict_shapes(audio_included, S, B).items():
        if name.endswith("num_batches_tracked"):
            out[name] = np.zeros((), dtype=np.int64)
        elif ".bnorm" in name:
            kind = name.rsplit(".", 1)[1]
            if kind == "weight":
                out[name] = rng.uniform(0.5, 1.5, shape).astype(np.float32)
            elif kind == "running_var":
                out[name] = rng.uniform(0.5, 1.5, shape).astype(np.float32)
            else:
                out[name] = (0.1 * rng.standard_normal(shape)).astype(np.float32)
        else:
            wshape = state_dict_shapes(audio_included, S, B)[name.rsplit(".", 1)[0] + ".weight"]
            fan_in = int(np.prod(wshape[1:]))
            bound = 1.0 / np.sqrt(fan_in)
            out[name] = rng.uniform(-bound, bound, shape).astype(np.float32)
    if score_gain != 1.0:
        w12 = out["fusion.12.weight"]
        out["fusion.12.weight"] = ((w12 - w12.mean()) * np.float32(score_gain)).astype(np.float32)
        out["fusion.12.bias"] = np.zeros_like(out["fusion.12.bias"])
    return out


# Last-layer gain of the "spread" weights the end-to-end tests, smoke() and bench.py score with: large enough that
# per-frame scores cover a rounding boundary (so the keyshot knapsack is not one long tie), small enough that the
# bf16 / TF32 arithmetic stays within the 1e-3 per-element score tolerance (the synthetic frames are statistically
# alike, so a frame-to-frame score spread can only come from amplifying small feature differences -- and the
# rounding noise with them).
SPREAD_GAIN = 100.0


def center_scores(weights: dict, scores, target: float = 3.5) -> dict:
    """Copy of `weights` whose last-layer bias is shifted so that the MEDIAN of `scores` (what the model produced with
    `weights`) becomes `target`.  With target 3.5 half of the frames round to 3 and half to 4 (utils.py:611), which
    makes the selection problem non-degenerate whatever the audio/frame seed."""
    s = np.asarray(scores, dtype=np.float64).reshape(-1)
    z = np.log((s - 1.0) / (5.0 - s))                       # logit of (s - 1) / 4   (utils.py:256,270 inverted)
    zt = np.log((target - 1.0) / (5.0 - target))
    out = dict(weights)
    out["fusion.12.bias"] = (np.asarray(weights["fusion.12.bias"], np.float64) + (zt - np.median(z))).astype(np.float32)
    return out


def make_frames(n: int, H: int = 224, W: int = 224, seed: int = 0) -> np.ndarray:
    """uint8 BGR HWC frames.  Per frame a random range [lo,hi] (lo in 0..63, hi in 192..255, so min/max
    differ per frame, section 8d) filled with 70 % smooth per-frame pattern (a random 2-D sinusoid per
    channel, so frames differ in content and the scores spread) + 30 % uniform noise."""
    rng = np.random.default_rng(SEED_BASE + 104729 * seed + 1)
    out = np.empty((n, H, W, 3), dtype=np.uint8)
    yy = (np.arange(H, dtype=np.float32) / H)[:, None, None]
    xx = (np.arange(W, dtype=np.float32) / W)[None, :, None]
    for i in range(n):
        lo = int(rng.integers(0, 64))
        hi = int(rng.integers(192, 256))
        fy, fx = rng.uniform(0.5, 4.0, size=2).astype(np.float32)
        ph = rng.uniform(0, 2 * np.pi, size=3).astype(np.float32)[None, None, :]
        pat = 0.5 + 0.5 * np.sin(np.float32(2 * np.pi) * (fy * yy + fx * xx) + ph)
        u = rng.random((H, W, 3), dtype=np.float32)
        v = lo + (hi - lo) * (np.float32(0.7) * pat + np.float32(0.3) * u)
        out[i] = np.clip(np.rint(v), 0, 255).astype(np.uint8)
    return out


def make_audio(n_slots: int, samples_per_slot: int = 8000, sr: int = 16000, seed: int = 0) -> np.ndarray:
    """float32 mono: 3 random sinusoids (100-4000 Hz, amp .1-.5) + N(0,.05^2) noise (section 8d)."""
    rng = np.random.default_rng(SEED_BASE + 1299709 * seed + 2)
    n = n_slots * samples_per_slot
    t = np.arange(n, dtype=np.float64) / sr
    y = np.zeros(n, dtype=np.float64)
    for _ in range(3):
        f = rng.uniform(100.0, 4000.0)
        a = rng.uniform(0.1, 0.5)
        ph = rng.uniform(0, 2 * np.pi)
        y += a * np.sin(2 * np.pi * f * t + ph)
    # slow amplitude modulation so that per-slot features differ
    y *= 0.6 + 0.4 * np.sin(2 * np.pi * 0.05 * t + rng.uniform(0, 2 * np.pi))
    y += 0.05 * rng.standard_normal(n)
    return y.astype(np.float32)


def make_change_points(n_raw: int, seed: int = 0, lo: int = 30, hi: int = 300) -> np.ndarray:
    """Inclusive [a,b] clip intervals covering 0..n_raw-1, lengths U{lo..hi}; last b = n_raw-1."""
    rng = np.random.default_rng(SEED_BASE + 15485863 * seed + 3)
    cps = []
    a = 0
    while a < n_raw:
        b = min(a + int(rng.integers(lo, hi + 1)) - 1, n_raw - 1)
        cps.append((a, b))
        a = b + 1
    return np.asarray(cps, dtype=np.int32).reshape(-1, 2)


def make_scores(n: int, seed: int = 0) -> np.ndarray:
    """Synthetic per-frame importance scores in [1,5], smooth + noise, float32 (N,1)."""
    rng = np.random.default_rng(SEED_BASE + 32452843 * seed + 4)
    x = np.linspace(0, 1, n)
    s = 3.0 + 1.6 * np.sin(2 * np.pi * (3 * x + rng.uniform())) + 0.5 * rng.standard_normal(n)
    return np.clip(s, 1.0, 5.0).astype(np.float32)[:, None]


def make_annotator_importances(n_raw: int, n_users: int = 20, seed: int = 0, shot: int = 60) -> np.ndarray:
    """(n_users, n_raw) integer 1..5 importances, piecewise-constant over `shot`-frame shots."""
    rng = np.random.default_rng(SEED_BASE + 49979687 * seed + 5)
    n_shots = (n_raw + shot - 1) // shot
    v = rng.integers(1, 6, size=(n_users, n_shots))
    return np.repeat(v, shot, axis=1)[:, :n_raw].astype(np.float32)
