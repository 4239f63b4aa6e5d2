"""Parity tests proper: the CUDA path (through the C-ABI / ctypes) against the CPU oracle and the golden
vectors produced by the reference's own code.  Integer work is bit-exact; floating point within the
tolerance written in each test."""
import numpy as np
import pytest
import torch

from cvml_goalnet_b200 import synthetic as syn
from oracle import goalnet_oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def U(cuda_device):
    from cvml_goalnet_b200 import utils
    return utils


@pytest.fixture(scope="module")
def weights():
    return syn.make_weights(seed=0)


# ------------------------------------------------------------------ (a1) preprocess
PRE_TOL = 5e-7   # abs, on values in [0,1]: closed form vs cv2 is <= 1.2e-7, plus fp32 FMA reassociation


def test_preprocess_golden(U, golden):
    g = golden("preprocess")
    for name in ("down", "down2", "up", "flat"):
        out = U.extract_condensed_frame_tensor_array(g[f"{name}_frames"], as_numpy=True)
        assert out.shape == g[f"{name}_out"].shape and out.dtype == np.float32
        assert np.abs(out - g[f"{name}_out"]).max() <= PRE_TOL, name
    if "file_out" in g.files:
        dec = g["file_decoded_frames"][:: int(g["file_skip_frames"])]
        assert np.abs(U.extract_condensed_frame_tensor_array(dec, as_numpy=True) - g["file_out"]).max() <= PRE_TOL


@pytest.mark.parametrize("n,H,W", [(1, 224, 224), (37, 224, 224), (3, 45, 45), (2, 270, 480), (0, 32, 32),
                                   (5, 17, 23), (2, 1080, 1920)])
def test_preprocess_vs_oracle(U, n, H, W):
    fr = syn.make_frames(n, H, W, seed=H + W)
    out = U.extract_condensed_frame_tensor_array(fr, as_numpy=True)
    ref = O.extract_condensed_frame_tensor_array(fr, 40)
    assert out.shape == (n, 3, 40, 40)
    if n:
        assert np.abs(out - ref).max() <= PRE_TOL


@pytest.mark.parametrize("n,H,W,S", [(3, 1080, 1920, 224), (9, 224, 224, 224), (4, 270, 480, 224), (2, 64, 64, 2),
                                     (3, 720, 1280, 40), (300, 96, 128, 40)])
def test_preprocess_other_sizes_and_cluster_split(U, n, H, W, S, monkeypatch):
    """R224 (BASELINE config 5: 1080p -> 224 runs as a thread-block cluster per frame), up-scaling, tiny S;
    the tap-capturing kernel and the L2 kernel must agree bit for bit (same arithmetic, different staging)."""
    fr = syn.make_frames(n, H, W, seed=H + W + S)
    out = U.extract_condensed_frame_tensor_array(fr, S, as_numpy=True)
    ref = O.extract_condensed_frame_tensor_array(fr[: min(n, 4)], S)
    assert out.shape == (n, 3, S, S)
    assert np.abs(out[: min(n, 4)] - ref).max() <= PRE_TOL
    monkeypatch.setenv("GN_PREPROCESS_L2", "1")
    out_l2 = U.extract_condensed_frame_tensor_array(fr, S, as_numpy=True)
    assert np.array_equal(out, out_l2)


def test_preprocess_unaligned_device_view(U):
    """frames whose device address is not 16-byte aligned (head/tail scalar path)."""
    fr = syn.make_frames(3, 33, 35, seed=9)
    buf = torch.zeros(fr.size + 64, dtype=torch.uint8, device="cuda")
    for off in (1, 7, 16):
        view = buf[off:off + fr.size].view(fr.shape)
        view.copy_(torch.from_numpy(fr))
        out = U.extract_condensed_frame_tensor_array(view).cpu().numpy()
        assert np.abs(out - O.extract_condensed_frame_tensor_array(fr, 40)).max() <= PRE_TOL


# ------------------------------------------------------------------ (a2) audio
# MFCC values span roughly [-600, 200]; the oracle evaluates the STFT in float64 and the rest in float32,
# the kernel is float32 throughout: 2e-3 abs + 1e-4 rel.
MFCC_ATOL, MFCC_RTOL = 2e-3, 1e-4


def test_audio_golden(U, golden):
    g = golden("audio")
    for name in ("int", "ragged", "b16"):
        sr, n_frames, B = (int(x) for x in g[f"{name}_args"])
        out = U.extract_audio_features_array(g[f"{name}_y"], sr, n_frames, B, as_numpy=True)
        ref = g[f"{name}_features"]
        assert out.shape == ref.shape
        np.testing.assert_allclose(out, ref, rtol=MFCC_RTOL, atol=MFCC_ATOL)


@pytest.mark.parametrize("n,spf,sr,B", [(64, 8000, 16000, 30), (5, 11025, 22050, 30), (3, 2100, 16000, 30),
                                        (0, 8000, 16000, 30),
                                        (2, 40000, 16000, 30),      # 79 STFT columns: the DCT goes in three column chunks
                                        (2, 40000, 16000, 60)])     # B q > 4096: W is read from global memory
def test_audio_vs_oracle(U, n, spf, sr, B):
    y = syn.make_audio(max(n, 1), spf, sr, seed=n)[: max(n, 1) * spf]
    if n == 0:
        assert U.extract_audio_features_array(y, sr, 0, B).shape == (0, 30, B)
        return
    out = U.extract_audio_features_array(y, sr, n, B, as_numpy=True)
    ref = O.extract_audio_features_array(y, sr, n, B)
    np.testing.assert_allclose(out, ref, rtol=MFCC_RTOL, atol=MFCC_ATOL)


def test_audio_matches_independent_mfcc_vectors(U, golden):
    """The CUDA MFCC against vectors the oracle did not produce: torchaudio.transforms.MFCC configured with librosa
    0.10's defaults (oracle/gen_mfcc_crosscheck.py).  One slot spanning the whole signal and B = q makes the cubic
    resampling (utils.py:336-343) the identity, so the output is the raw MFCC matrix of utils.py:333."""
    g = golden("mfcc_torchaudio")
    checked = 0
    for name in [k[:-5] for k in g.files if k.endswith("_mfcc")]:
        y, sr, ref = g[f"{name}_y"], int(g[f"{name}_sr"]), g[f"{name}_mfcc"]
        q = ref.shape[1]
        if q < 4:                              # interp1d(kind='cubic') needs 4 columns (reference fails too)
            continue
        out = U.extract_audio_features_array(y, sr, 1, q, as_numpy=True)
        assert out.shape == (1, 30, q)
        np.testing.assert_allclose(out[0], ref, rtol=MFCC_RTOL, atol=MFCC_ATOL, err_msg=name)
        checked += 1
    assert checked >= 5


def test_audio_silence_and_short_segment_errors(U):
    out = U.extract_audio_features_array(np.zeros(4 * 8000, np.float32), 16000, 4, 30, as_numpy=True)
    ref = O.extract_audio_features_array(np.zeros(4 * 8000, np.float32), 16000, 4, 30)
    np.testing.assert_allclose(out, ref, rtol=MFCC_RTOL, atol=MFCC_ATOL)
    with pytest.raises(AssertionError):          # q < 4 columns: interp1d(kind='cubic') fails in the reference too
        U.extract_audio_features_array(np.zeros(1000, np.float32), 16000, 1, 30)


# ------------------------------------------------------------------ (a3-a5) model
SCORE_RTOL = 1e-3          # north_star: <= 1e-3 relative on scores, PER ELEMENT: |gpu - ref| <= 1e-3 * |ref| for every frame
                           # (visual backbone bf16 x bf16 -> fp32; AudBl GEMMs and fusion MLP TF32 x TF32 -> fp32)
FEAT_TOL = 2e-2            # visual features: bf16 activations through K = 41472; relative to the max feature


def assert_scores(s, ref, what=""):
    s, ref = np.asarray(s, np.float64).reshape(-1), np.asarray(ref, np.float64).reshape(-1)
    assert s.shape == ref.shape, what
    rel = np.abs(s - ref) / np.abs(ref)
    assert rel.max() <= SCORE_RTOL, (what, float(rel.max()), int(rel.argmax()))


def spread_weights(frames, y, sr, audio_included=True):
    """Weights whose scores cross a rounding boundary on THESE inputs: last-layer gain syn.SPREAD_GAIN, bias centred so
    that the oracle's median score is 3.5 (half the frames round to 3, half to 4 -- the knapsack is no longer one long
    tie).  Returns (weights, oracle scores (n,1) float32, oracle features dict)."""
    w = syn.make_weights(seed=0, audio_included=audio_included, score_gain=syn.SPREAD_GAIN)
    vis = O.extract_condensed_frame_tensor_array(frames, 40)
    aud = O.extract_audio_features_array(y, sr, len(frames), 30).astype(np.float32)
    w = syn.center_scores(w, O.avm_forward(w, aud, vis, audio_included=audio_included).numpy())
    ref, feats = O.avm_forward(w, aud, vis, audio_included=audio_included, return_features=True)
    ref = ref.numpy()
    classes = np.unique(np.round(ref))
    assert len(classes) >= 2 and ref.max() - ref.min() > 0.3, (classes, ref.min(), ref.max())   # the spread is real
    return w, ref, feats


def assert_keyshots(U, s_gpu, ind_gpu, o_scores, skip, n_raw, cps):
    """Scores -> selection continuity, unconditionally.  Frames whose ORACLE score lies within the score tolerance of
    x.5 may legitimately round either way (utils.py:611); they are listed and excluded from the rounding check, every
    other frame must round exactly as the oracle's.  The keyshot indicator must then equal, bit for bit, what the
    oracle's integer path selects from the oracle's scores with only those ambiguous frames taken from the GPU."""
    s_gpu = np.asarray(s_gpu, np.float32).reshape(-1)
    o = np.asarray(o_scores, np.float32).reshape(-1)
    frac = np.abs(o - np.floor(o) - 0.5)
    ambiguous = frac <= SCORE_RTOL * np.abs(o)
    assert ambiguous.mean() < 0.05, "too many frames sit on a rounding boundary for the check to mean anything"
    assert np.array_equal(np.round(s_gpu[~ambiguous]), np.round(o[~ambiguous])), "a frame far from x.5 rounded differently"
    patched = np.where(ambiguous, s_gpu, o).astype(np.float32)
    _, o_ind = O.postprocess_array(patched[:, None], skip, n_raw, cps)
    np.testing.assert_array_equal(np.asarray(ind_gpu), o_ind)
    assert 0 < int(o_ind.sum()) < n_raw
    return int(ambiguous.sum())


def _model(U, w, audio_included=True):
    m = U.AVM(audio_included)
    m.load_state_dict(w)
    return m


def test_model_golden_reference_avm(U, golden):
    g = golden("model")
    for tag, audio_included, gain in (("av", True, 1.0), ("v", False, 1.0), ("av_gain", True, 100.0)):
        w = syn.make_weights(seed=0, audio_included=audio_included, score_gain=gain)
        m = _model(U, w, audio_included)
        for simt in (True, False):
            y = m(g["audio_input"], g["visual_input"], return_features=True, _simt=simt).cpu().numpy()
            ref = g[f"{tag}_scores"]
            assert y.shape == ref.shape == (6, 1)
            assert_scores(y, ref, (tag, simt))            # also the x100 last layer (TF32 head)
            vf = m.last_visual_features.cpu().numpy()
            rv = g[f"{tag}_visual_features"]
            assert np.abs(vf - rv).max() <= FEAT_TOL * max(np.abs(rv).max(), 1e-3), (tag, simt)
            for which, key in ((1, "vis1"), (2, "vis2"), (3, "vis3")):
                act = m.read_activation(which, 6).cpu().numpy()[:2]
                r = g[f"{tag}_{key}"]
                assert np.abs(act - r).max() <= FEAT_TOL * np.abs(r).max(), (tag, simt, key)


@pytest.mark.parametrize("n", [1, 7, 128, 129, 300])
def test_model_vs_oracle_and_simt(U, weights, n):
    vis = O.extract_condensed_frame_tensor_array(syn.make_frames(n, 48, 48, seed=n), 40)
    rng = np.random.default_rng(n)
    aud = (25.0 * rng.standard_normal((n, 30, 30))).astype(np.float32)
    aud[:, 0] -= 250.0
    m = _model(U, weights)
    y_tc = m(aud, vis, return_features=True).cpu().numpy()
    f_tc = m.last_visual_features.cpu().numpy()
    h0_tc = m.read_activation(4, n).cpu().numpy()
    y_simt = m(aud, vis, return_features=True, _simt=True).cpu().numpy()
    f_simt = m.last_visual_features.cpu().numpy()
    ref, feats = O.avm_forward(weights, aud, vis, return_features=True)
    ref = ref.numpy()
    assert_scores(y_tc, ref, "tc")
    assert_scores(y_simt, ref, "simt")
    rv = feats["visual_features"].numpy()
    assert np.abs(f_tc - rv).max() <= FEAT_TOL * np.abs(rv).max()
    # tensor-core path vs the CUDA-core validation path: same bf16 operands, only the summation order differs,
    # which can flip the final bf16 rounding of a feature by one ulp (2^-8 relative)
    assert np.abs(f_tc - f_simt).max() <= 8e-3 * np.abs(rv).max()
    ra = feats["audio_features"].numpy()
    assert np.abs(h0_tc[:, :128] - ra).max() <= 2e-3 * max(np.abs(ra).max(), 1e-3)   # TF32 AudBl GEMMs, TF32-rounded store


def test_block1_tensor_core_kernel_vs_cuda_core_kernel_and_oracle(U, weights):
    """VisBl block 1 (utils.py:174-177) of the 40x40 model runs as one tcgen05 TF32 GEMM per frame (gn_front_tc.cu); the
    CUDA-core kernel (fp32 operands) is the validation path.  Both against the oracle's block-1 output, and against each
    other, on a frame count that is neither a multiple of the grid nor of two; all-black and all-white frames included
    (min == max: the normalisation's 1e-7 guard, utils.py:284)."""
    n = 601
    frames = syn.make_frames(n, 48, 48, seed=5)
    frames[7] = 0
    frames[8] = 255
    vis = O.extract_condensed_frame_tensor_array(frames, 40)
    aud = np.zeros((n, 30, 30), np.float32)
    m = _model(U, weights)
    m(aud, vis)
    a_tc = m.read_activation(1, n).cpu().numpy()
    m(aud, vis, _simt=True)
    a_simt = m.read_activation(1, n).cpu().numpy()
    _, feats = O.avm_forward(weights, aud, vis, return_features=True)
    assert a_tc.shape == a_simt.shape == (n, 64, 13, 13)
    scale = np.abs(a_simt).max()
    # the stored activations are bf16: TF32 operand rounding can move a value across a bf16 rounding boundary (2^-8)
    assert np.abs(a_tc - a_simt).max() <= 2.0 ** -7 * scale
    r = feats["vis1"].numpy()
    for a in (a_tc, a_simt):                   # one bf16 rounding of the output (2^-9 relative) + the operand rounding
        assert np.abs(a - r).max() <= 2.0 ** -8 * np.abs(r).max()


@pytest.mark.parametrize("S,n", [(64, 5), (100, 3), (46, 130), (224, 2)])
def test_model_other_frame_sizes_tiled_conv_and_splitk(U, S, n):
    """Frame sizes other than the reference's 40 (R224 = BASELINE config 5; legal because every layer of the
    reference AVM is Lazy*): overlapping 16x16 conv tiles, separate border pass, K-split linear5."""
    w = syn.make_weights(seed=3, S=S)
    vis = O.extract_condensed_frame_tensor_array(syn.make_frames(n, S + 9, S + 5, seed=S), S)
    rng = np.random.default_rng(S)
    aud = (25.0 * rng.standard_normal((n, 30, 30))).astype(np.float32)
    m = U.AVM(True, S=S)
    m.load_state_dict(w)
    y = m(aud, vis, return_features=True).cpu().numpy()
    vf = m.last_visual_features.cpu().numpy()
    ref, feats = O.avm_forward(w, aud, vis, return_features=True)
    for which, key in ((1, "vis1"), (2, "vis2"), (3, "vis3")):
        act = m.read_activation(which, n).cpu().numpy()
        r = feats[key].numpy()
        assert act.shape == r.shape, key
        assert np.abs(act - r).max() <= FEAT_TOL * np.abs(r).max(), key
    rv = feats["visual_features"].numpy()
    assert np.abs(vf - rv).max() <= FEAT_TOL * np.abs(rv).max()
    assert_scores(y, ref.numpy(), S)
    del m
    torch.cuda.empty_cache()


@pytest.mark.parametrize("audio_included,n", [(True, 300), (False, 129), (True, 1)])
def test_classifier_head_cavm(U, weights, audio_included, n):
    """Linear(128,5) + Softmax head (the reference's commented variant, utils.py:257, main.py:69,97).
    Probabilities within 2e-3 abs of the fp32 oracle (bf16 operands upstream); argmax+1 equal wherever the oracle's
    top-2 margin exceeds twice that tolerance."""
    w = dict(weights)
    rng = np.random.default_rng(55)
    w["fusion.12.weight"] = (rng.standard_normal((5, 128)) * 0.15).astype(np.float32)
    w["fusion.12.bias"] = (rng.standard_normal(5) * 0.1).astype(np.float32)
    if not audio_included:
        w = {k: v for k, v in w.items() if not k.startswith("audbl.")}
        w["fusion.0.weight"] = np.ascontiguousarray(np.asarray(w["fusion.0.weight"])[:, 128:])
    vis = O.extract_condensed_frame_tensor_array(syn.make_frames(n, 48, 48, seed=n + 1), 40)
    aud = (25.0 * rng.standard_normal((n, 30, 30))).astype(np.float32)
    m = U.CAVM(audio_included)
    m.load_state_dict(w)
    probs = m(aud, vis)
    assert probs.shape == (n, 5) and probs.dtype == torch.float32 and probs.is_cuda
    p = probs.cpu().numpy()
    ref = O.avm_forward(w, aud, vis, audio_included=audio_included, classifier=True).numpy()
    assert np.abs(p.sum(1) - 1).max() <= 1e-5
    assert np.abs(p - ref).max() <= 2e-3
    pred = m.predict(aud, vis).cpu().numpy()
    assert pred.shape == (n, 1) and pred.min() >= 1 and pred.max() <= 5
    top2 = np.sort(ref, axis=1)
    sure = (top2[:, -1] - top2[:, -2]) > 4e-3
    assert np.array_equal(pred[sure, 0], ref.argmax(1)[sure] + 1.0)
    assert ref.max() < 0.999                  # not saturated: the 2e-3 comparison above is meaningful
    # predictions feed the same selection path as the regression scores
    nraw = n * 15
    cps = syn.make_change_points(nraw, seed=3) if hasattr(syn, "make_change_points") else None
    if cps is not None:
        _, ind = U.postprocess_array(pred, 15, nraw, cps)
        _, ind_ref = O.postprocess_array(pred, 15, nraw, cps)
        assert np.array_equal(np.asarray(ind), np.asarray(ind_ref))
    with pytest.raises(AssertionError):
        m(aud, vis, _simt=True)


def test_model_empty_and_shape_errors(U, weights):
    m = _model(U, weights)
    assert m(np.zeros((0, 30, 30), np.float32), np.zeros((0, 3, 40, 40), np.float32)).shape == (0, 1)
    with pytest.raises(AssertionError):
        m(np.zeros((2, 30, 30), np.float32), np.zeros((2, 3, 32, 32), np.float32))
    with pytest.raises(AssertionError):
        m(np.zeros((3, 30, 30), np.float32), np.zeros((2, 3, 40, 40), np.float32))


def test_state_dict_round_trip(U, weights):
    m = _model(U, weights)
    sd = m.state_dict()
    assert sorted(sd.keys()) == sorted(weights.keys())
    for k, v in weights.items():
        assert tuple(sd[k].shape) == tuple(np.asarray(v).shape), k


# ------------------------------------------------------------------ (a6-a9) integer path, bit-exact
def test_knapsack_golden(U, golden):
    g = golden("integer_path")
    for i in range(int(g["knap_n"])):
        v, w, c = g[f"knap_{i}_values"].tolist(), g[f"knap_{i}_weights"].tolist(), int(g[f"knap_{i}_cap"])
        assert U.knapsack(v, w, c) == g[f"knap_{i}_sel"].tolist(), i


def test_knapsack_random_vs_oracle(U):
    rng = np.random.default_rng(5)
    for t in range(25):
        n = int(rng.integers(1, 120))
        v = rng.integers(0, 400, n).tolist()
        w = (rng.integers(0, 60, n) * int(rng.choice([1, 1, 3]))).tolist()
        cap = int(rng.integers(0, max(2, sum(w))))
        scale = int(rng.choice([1, 5, 7]))
        assert U.knapsack(v, w, cap, scale) == O.knapsack_rows(v, w, cap, scale), t
    # float weights: int(w * scale) truncation (utils.py:478)
    v, w = [10, 20, 30, 40], [1.5, 2.25, 3.9, 0.1]
    assert U.knapsack(v, w, 4.4) == O.knapsack(v, w, 4.4)


def test_knapsack_wide_table_cluster_slices(U):
    """60 001 columns: 8 cluster CTAs x 7 520-column slices; weights larger than a slice reach across CTAs."""
    rng = np.random.default_rng(11)
    n = 40
    v = rng.integers(1, 1000, n).tolist()
    w = rng.integers(1000, 9000, n).tolist()     # gcd 1 -> 60001 columns
    assert U.knapsack(v, w, 60000, 1) == O.knapsack_rows(v, w, 60000, 1)
    w2 = rng.integers(1, 30000, n).tolist()      # weights spanning up to 4 slices
    assert U.knapsack(v, w2, 60000, 1) == O.knapsack_rows(v, w2, 60000, 1)


@pytest.mark.parametrize("rows_per_step", ["1", "2", "4"])
def test_knapsack_rows_per_barrier_variants(U, monkeypatch, rows_per_step):
    """The multi-row kernel (2 or 4 DP rows per cluster barrier, halo from the neighbour CTAs) against the
    one-row kernel and the oracle: match-sized table (984 clips x 24 301 columns), item counts that are not
    a multiple of the step, zero weights, a weight-0/column-0 corner, and weights that exceed the halo (the
    kernel declines and the one-row kernel behind it takes over)."""
    monkeypatch.setenv("GN_KNAP_T", rows_per_step)
    rng = np.random.default_rng(21)
    cases = []
    n = 984
    cases.append((rng.integers(30, 1500, n).tolist(), rng.integers(30, 301, n).tolist(), 24300, 1))
    for n in (1, 2, 3, 5, 33, 127):
        cases.append((rng.integers(0, 500, n).tolist(), rng.integers(0, 200, n).tolist(), 9000, 1))
    cases.append((rng.integers(1, 500, 50).tolist(), rng.integers(1, 40, 50).tolist(), 300, 5))     # narrow table
    cases.append((rng.integers(1, 500, 37).tolist(), rng.integers(3000, 20000, 37).tolist(), 50000, 1))  # > halo
    cases.append(([5, 7, 9, 11], [0, 0, 3, 0], 3, 1))
    for i, (v, w, cap, scale) in enumerate(cases):
        assert U.knapsack(v, w, cap, scale) == O.knapsack_rows(v, w, cap, scale), (rows_per_step, i)


def test_knapsack_very_wide_table_global_rows(U):
    """capacity too wide even for 8 shared-memory slices -> single-CTA kernel with rows in the workspace."""
    rng = np.random.default_rng(12)
    n = 10
    v = rng.integers(1, 1000, n).tolist()
    w = rng.integers(20000, 90000, n).tolist()
    assert U.knapsack(v, w, 230000, 1) == O.knapsack_rows(v, w, 230000, 1)


def test_knapsack_batched_problems(U):
    """gn_knapsack batch > 1 (one cluster per problem): the 20 annotator knapsacks of utils.py:102-118 at once."""
    import torch
    rng = np.random.default_rng(13)
    batch, n, cap = 20, 60, 5400
    vals = rng.integers(0, 900, (batch, n)).astype(np.int32)
    wts = np.tile(rng.integers(30, 300, n).astype(np.int32), (batch, 1))
    from cvml_goalnet_b200.utils import _knapsack_device
    dev = torch.device("cuda", 0)
    sel, best = _knapsack_device(torch.from_numpy(vals).to(dev), torch.from_numpy(wts).to(dev), n, cap, 0, batch, dev)
    sel = sel.cpu().numpy()
    for b in range(batch):
        ref = O.knapsack_rows(vals[b].tolist(), wts[b].tolist(), cap, 1)
        assert np.nonzero(sel[b])[0].tolist() == ref, b


def test_clip_information_golden(U, golden):
    g = golden("integer_path")
    v, l, _ = U.get_clip_information(g["clip_cps"].tolist(), g["clip_imp"].tolist())
    assert v == g["clip_values"].tolist() and l == g["clip_lengths"].tolist()


def test_postprocess_golden(U, golden):
    g = golden("postprocess")
    for name in ("a", "b", "c"):
        skip, n_raw = (int(x) for x in g[f"{name}_args"])
        _, ind = U.postprocess_array(g[f"{name}_scores"], skip, n_raw, g[f"{name}_cps"])
        assert ind.dtype == np.uint8
        np.testing.assert_array_equal(ind, g[f"{name}_indicator"])


@pytest.mark.parametrize("N,skip,n_raw", [(240, 15, 3600), (240, 15, 3590), (240, 15, 3700), (10800, 15, 162000),
                                          (500, 1, 500), (1, 30, 17)])
def test_postprocess_vs_oracle(U, N, skip, n_raw):
    scores = syn.make_scores(N, seed=N)
    scores[::7] = np.round(scores[::7]) + 0.5          # exercise round-half-to-even
    cps = syn.make_change_points(n_raw, seed=N + 1)
    _, ind = U.postprocess_array(scores, skip, n_raw, cps)
    _, ref = O.postprocess_array(scores, skip, n_raw, cps)
    np.testing.assert_array_equal(ind, ref)
    full = np.arange(n_raw)[:, None] * np.ones((1, 2), np.int64)
    vid, _ = U.postprocess_array(scores, skip, n_raw, cps, full_frames=full)
    sel, _ = O.postprocess_array(scores, skip, n_raw, cps)
    if sel:
        exp = np.concatenate([full[cps[j][0]:cps[j][1]] for j in sel])
        np.testing.assert_array_equal(vid, exp)


def test_fscore_golden_and_oracle(U, golden):
    g = golden("integer_path")
    assert list(U.get_fscore(g["fs_G"], g["fs_S"])) == g["fs_avg_max"].tolist()            # bit-exact float64
    assert list(U.get_fscore(g["fs_G"], np.zeros_like(g["fs_S"]))) == g["fs_empty_avg_max"].tolist()
    rng = np.random.default_rng(3)
    for n_raw in (162000, 3601, 16, 5):
        G = (rng.random((20, n_raw)) < 0.3).astype(np.uint8)
        S = (rng.random(n_raw) < 0.15).astype(np.uint8)
        assert U.get_fscore(G, S) == O.get_fscore(G, S)


def test_gt_summaries_like_get_dataloaders(U):
    """utils.py:102-118: one postprocess per annotator builds the ground-truth summaries."""
    n_raw, skip = 3600, 15
    ann = syn.make_annotator_importances(n_raw, 20, seed=2)
    cps = syn.make_change_points(n_raw, seed=2)
    for u in (0, 7, 19):
        _, ind = U.postprocess_array(ann[u][:, None], skip, n_raw, cps)
        _, ref = O.postprocess_array(ann[u][:, None], skip, n_raw, cps)
        np.testing.assert_array_equal(ind, ref)


@pytest.mark.parametrize("n_raw,users", [(3600, 20), (27000, 20), (162000, 5), (450, 3)])
def test_gt_summaries_batched_equals_per_annotator(U, n_raw, users):
    """the U per-annotator knapsacks of utils.py:102-118 as one batched DP launch."""
    skip = 15
    ann = np.stack(syn.make_annotator_importances(n_raw, users, seed=n_raw))
    cps = syn.make_change_points(n_raw, seed=n_raw + 1)
    out = U.gt_summaries_array(ann, skip, n_raw, cps)
    assert out.shape == (users, n_raw) and out.dtype == np.uint8
    for u in range(users):
        _, ref = O.postprocess_array(ann[u][:, None], skip, n_raw, cps)
        np.testing.assert_array_equal(out[u], ref, err_msg=f"annotator {u}")


# ------------------------------------------------------------------ whole path
def test_end_to_end_config1(U):
    """BASELINE config 1: 2-min clip, 224x224 source, 2 fps, 16 kHz audio -> scores, keyshots, F-score, on weights whose
    scores spread over a rounding boundary."""
    n, skip, sr = 240, 15, 16000
    n_raw = n * skip
    frames = syn.make_frames(n, 224, 224, seed=1)
    y = syn.make_audio(n, 8000, sr, seed=1)
    cps = syn.make_change_points(n_raw, seed=1)
    w, o_scores, _ = spread_weights(frames, y, sr)
    m = _model(U, w)
    vis = U.extract_condensed_frame_tensor_array(frames)
    aud = U.extract_audio_features_array(y, sr, n, 30)
    scores = m(aud, vis)
    _, ind = U.postprocess_array(scores, skip, n_raw, cps)
    s = scores.cpu().numpy()
    assert_scores(s, o_scores)
    assert_keyshots(U, s, ind, o_scores, skip, n_raw, cps)
    # the oracle's own scores through the CUDA integer path: bit-exact against the oracle's integer path
    o_sel, o_ind = O.postprocess_array(o_scores, skip, n_raw, cps)
    _, ind_from_oracle_scores = U.postprocess_array(o_scores, skip, n_raw, cps)
    np.testing.assert_array_equal(ind_from_oracle_scores, o_ind)
    assert 0 < len(o_sel) < len(cps)
    gd = np.stack([O.postprocess_array(a[:, None], skip, n_raw, cps)[1]
                   for a in syn.make_annotator_importances(n_raw, 20, seed=1)])
    assert U.get_fscore(gd, o_ind) == O.get_fscore(gd, o_ind)


def test_match_scorer_chunked_host_path_equals_device_path(U):
    """MatchScorer.score_match (pinned host inputs, chunked H2D on two copy streams, per-chunk MFCC / preprocess /
    forward) gives the same bits as the one-shot device-resident path (a frame's score does not depend on the batch it
    is scored in), and both match the oracle."""
    import torch
    from cvml_goalnet_b200.main import MatchScorer
    n, skip, sr, spf = 50, 15, 16000, 8013          # non-integer-friendly slot length: ragged slot boundaries
    n_raw = n * skip
    frames = syn.make_frames(n, 64, 80, seed=4)
    y = syn.make_audio(n, spf, sr, seed=4)[: n * spf - 7]
    cps = syn.make_change_points(n_raw, seed=4, lo=10, hi=60)
    gd = np.stack([O.postprocess_array(a[:, None], skip, n_raw, cps)[1]
                   for a in syn.make_annotator_importances(n_raw, 20, seed=4)])
    w, o_scores, _ = spread_weights(frames, y, sr)
    m = _model(U, w)
    sc = MatchScorer(m, n, 64, 80, y.size, sr, skip, n_raw, h2d_chunks=3)
    s_host, ind_host, fs_host = sc.score_match(frames, y, cps, gd)
    s_host, ind_host = s_host.copy(), ind_host.copy()
    dev = torch.device("cuda", 0)
    s_dev, ind_dev, fs_dev = sc.score_device(torch.from_numpy(frames).to(dev), torch.from_numpy(y).to(dev), cps,
                                             torch.from_numpy(gd).to(dev))
    np.testing.assert_array_equal(s_host, s_dev.cpu().numpy())
    np.testing.assert_array_equal(ind_host, ind_dev.cpu().numpy())
    assert fs_host == fs_dev
    # a second step through the same (reused) buffers gives the same bits
    s2, ind2, fs2 = sc.score_match(frames, y, cps, gd)
    np.testing.assert_array_equal(s2, s_host)
    np.testing.assert_array_equal(ind2, ind_host)
    assert_scores(s_host, o_scores)
    assert_keyshots(U, s_host, ind_host, o_scores, skip, n_raw, cps)
    assert fs_host == O.get_fscore(gd, ind_host)


def test_scores_do_not_depend_on_the_batch(U):
    """Batch invariance: every frame's score is bit-identical whether it is scored alone, in a small batch, or inside a
    large one at any row position (linear5 always splits K into the same ranges; every other layer is per frame or per
    row).  This is what lets chunked, sharded and whole-video runs reproduce each other's bits."""
    n = 300
    frames = syn.make_frames(n, 48, 48, seed=77)
    vis = U.extract_condensed_frame_tensor_array(frames)
    rng = np.random.default_rng(7)
    aud = torch.from_numpy((25.0 * rng.standard_normal((n, 30, 30))).astype(np.float32)).cuda()
    m = _model(U, syn.make_weights(seed=0, score_gain=syn.SPREAD_GAIN))
    full = m(aud, vis).cpu().numpy()
    for a, b in ((0, 1), (7, 8), (0, 129), (129, 300), (50, 178), (299, 300)):
        part = m(aud[a:b].contiguous(), vis[a:b].contiguous()).cpu().numpy()
        np.testing.assert_array_equal(part, full[a:b], err_msg=f"rows {a}:{b}")
    big_v = vis.repeat(12, 1, 1, 1)
    big_a = aud.repeat(12, 1, 1)
    big = m(big_a, big_v).cpu().numpy().reshape(12, n)
    for r in range(12):
        np.testing.assert_array_equal(big[r], full[:, 0], err_msg=f"repetition {r}")


@pytest.mark.parametrize("reps", [25, 19])
def test_full_size_config2_match(U, reps):
    """BASELINE config 2 at full size (reps = 25: 10 800 frames of 224x224, 86.4 M audio samples, 162 000 raw frames)
    through the same MatchScorer.score_device call bench.py times, on weights whose scores spread over a rounding
    boundary; reps = 19 (8 208 frames) is a second wave/tile geometry.  The oracle cannot score 10 800 frames in
    seconds, so the match repeats a 432-frame base (frames and their 8 000-sample audio slots): every frame's score
    depends only on that frame and its slot (utils.py:260-272, eval mode), hence
      * the VisBl FEATURES (utils.py:189-192: flatten -> linear5 -> ReLU) of ALL rows equal the oracle's features of the
        base within FEAT_TOL -- the visual branch moves the score by less than the score tolerance, so it is checked
        where it can be seen;
      * all scores equal the oracle's scores of the base per element, and every repetition gives the same BITS;
      * selection and F-score counts are bit-exact against the oracle's integer path at full size."""
    from cvml_goalnet_b200.main import MatchScorer
    base, skip, sr, spf = 432, 15, 16000, 8000
    n = base * reps
    n_raw = n * skip
    frames = syn.make_frames(base, 224, 224, seed=9)
    y = syn.make_audio(base, spf, sr, seed=9)
    assert y.size == base * spf
    cps = syn.make_change_points(n_raw, seed=9)
    w, o_base, o_feats = spread_weights(frames, y, sr)
    dev = torch.device("cuda", 0)
    frames_dev = torch.from_numpy(frames).to(dev).repeat(reps, 1, 1, 1)
    y_dev = torch.from_numpy(y).to(dev).repeat(reps)
    m = _model(U, w)
    sc = MatchScorer(m, n, 224, 224, y_dev.numel(), sr, skip, n_raw)
    scores, ind, _ = sc.score_device(frames_dev, y_dev, cps, None, return_features=True)
    s = scores.cpu().numpy()
    ind = ind.cpu().numpy()
    vf = sc.features.cpu().numpy()
    assert s.shape == (n, 1) and ind.shape == (n_raw,) and ind.dtype == np.uint8 and vf.shape == (n, 512)

    rv = o_feats["visual_features"].numpy()
    assert np.abs(rv).max() > 0.05
    err = np.abs(vf.reshape(reps, base, 512) - rv[None]).max()
    assert err <= FEAT_TOL * np.abs(rv).max(), err                      # every row: all tiles, waves and K ranges
    o_full = np.tile(o_base.reshape(base, 1), (reps, 1))
    assert_scores(s, o_full)
    # position independence, exactly: the same frame gives the same bits in every repetition
    sr_ = s.reshape(reps, base)
    assert np.array_equal(sr_, np.broadcast_to(sr_[0], sr_.shape))
    assert np.array_equal(vf.reshape(reps, base, 512), np.broadcast_to(vf[:base], (reps, base, 512)))

    # scores -> selection at full size: rounding agrees off the x.5 boundaries, keyshots bit-exact
    assert_keyshots(U, s, ind, o_full, skip, n_raw, cps)
    assert 0 < int(ind.sum()) <= int(0.15 * n_raw) + len(cps)     # budget (+1 frame per clip: get_clip_information is end-exclusive)
    gd = np.stack([O.postprocess_array(a[:, None], skip, n_raw, cps)[1]
                   for a in syn.make_annotator_importances(n_raw, 20, seed=9)])
    assert U.get_fscore(gd, ind) == O.get_fscore(gd, ind)


def test_out_of_bounds_canaries(U, weights):
    """compute-sanitizer is closed on this GPU pool, so out-of-bounds writes are hunted with guard bands: every output
    and workspace buffer of the path is carved out of a larger allocation filled with a canary byte, and the bands on
    both sides must be intact after the kernels ran (ragged sizes on purpose)."""
    import ctypes as C
    from cvml_goalnet_b200 import _lib
    dev = torch.device("cuda", 0)
    lib, h = _lib.load(), _lib.handle(0)
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    G = 4096

    def guarded(nbytes, align=1024):
        buf = torch.full((nbytes + 2 * G + align,), 0xA5, dtype=torch.uint8, device=dev)
        off = G + (-(buf.data_ptr() + G)) % align
        return buf, off, buf[off:off + nbytes]

    def intact(buf, off, nbytes, what):
        torch.cuda.synchronize()
        assert bool((buf[:off] == 0xA5).all()) and bool((buf[off + nbytes:] == 0xA5).all()), f"guard band hit: {what}"

    # preprocess (tap kernel, cluster split, L2 kernel via an odd width)
    for (n, H, W, S) in ((7, 224, 224, 40), (3, 1080, 1920, 224), (5, 45, 47, 40)):
        fr = torch.from_numpy(syn.make_frames(n, H, W, seed=3)).to(dev)
        nb = n * 3 * S * S * 4
        buf, off, view = guarded(nb)
        U.extract_condensed_frame_tensor_array(fr, S, out=view.view(torch.float32).view(n, 3, S, S))
        intact(buf, off, nb, f"preprocess {H}x{W}->{S}")
    # MFCC with ragged slots
    n, spf = 37, 8013
    y = torch.from_numpy(syn.make_audio(n, spf, 16000, seed=2)[: n * spf - 11]).to(dev)
    nb = n * 30 * 30 * 4
    buf, off, view = guarded(nb)
    U.extract_audio_features_array(y, 16000, n, 30, out=view.view(torch.float32).view(n, 30, 30))
    intact(buf, off, nb, "mfcc")
    # forward: workspace, scores and features
    m = _model(U, weights)
    for n in (1, 129, 300):
        vis = torch.rand((n, 3, 40, 40), device=dev)
        aud = 25.0 * torch.randn((n, 30, 30), device=dev)
        mptr = m._model(dev)
        need = int(lib.gn_avm_workspace_bytes(mptr, n))
        wb, woff, ws = guarded(need)
        sb, soff, sv = guarded(n * 4, 16)
        fb, foff, fv = guarded(n * 512 * 4, 16)
        _lib.check(lib.gn_avm_forward(mptr, C.c_void_p(vis.data_ptr()), C.c_void_p(aud.data_ptr()), n,
                                      C.c_void_p(sv.data_ptr()), C.c_void_p(fv.data_ptr()), C.c_void_p(ws.data_ptr()),
                                      need, stream), "gn_avm_forward")
        intact(wb, woff, need, f"forward workspace n={n}")
        intact(sb, soff, n * 4, f"scores n={n}")
        intact(fb, foff, n * 512 * 4, f"features n={n}")
    # knapsack: cluster kernels, the global-row fallback forced by a long item list, batches
    rng = np.random.default_rng(4)
    for (n_items, cap, batch) in ((984, 24300, 1), (60, 5400, 20), (30000, 1000, 1), (10, 230000, 1), (37, 50000, 3)):
        vals = torch.from_numpy(rng.integers(0, 500, (batch, n_items)).astype(np.int32)).to(dev)
        wts = torch.from_numpy(rng.integers(1, max(2, cap // 8), (batch, n_items)).astype(np.int32)).to(dev)
        need = int(lib.gn_knapsack_workspace_bytes(n_items, cap)) * batch
        wb, woff, ws = guarded(need, 256)
        sb, soff, sv = guarded(batch * n_items, 16)
        bb, boff, bv = guarded(batch * 4, 16)
        _lib.check(lib.gn_knapsack(h, C.c_void_p(vals.data_ptr()), C.c_void_p(wts.data_ptr()), n_items, cap, 0, batch,
                                   C.c_void_p(sv.data_ptr()), C.c_void_p(bv.data_ptr()), C.c_void_p(ws.data_ptr()), need,
                                   stream), "gn_knapsack")
        intact(wb, woff, need, f"knapsack workspace {n_items}x{cap}x{batch}")
        intact(sb, soff, batch * n_items, "knapsack selected")
        intact(bb, boff, batch * 4, "knapsack best")
        sel = sv.view(batch, n_items).cpu().numpy()
        v0, w0 = vals[0].cpu().numpy().tolist(), wts[0].cpu().numpy().tolist()
        assert np.nonzero(sel[0])[0].tolist() == O.knapsack_rows(v0, w0, cap, 1), (n_items, cap)


def test_handles_on_two_device_indices(U, weights):
    """Function attributes (dynamic shared memory limit, cluster size) are per DEVICE: the library records them per
    handle.  With two GPUs the second handle must launch every > 48 KB kernel too; with one GPU the same handle logic is
    exercised by a fresh handle created after the first one has set its attributes."""
    import ctypes as C
    from cvml_goalnet_b200 import _lib
    lib = _lib.load()
    ndev = torch.cuda.device_count()
    idx = 1 if ndev > 1 else 0
    dev = torch.device("cuda", idx)
    _lib.handle(0)
    p = C.c_void_p()
    with torch.cuda.device(dev):
        _lib.check(lib.gn_create(C.byref(p), idx), "gn_create")
        saved = _lib._handles.get(idx)
        _lib._handles[idx] = p
        try:
            fr = syn.make_frames(3, 224, 224, seed=5)
            out = U.extract_condensed_frame_tensor_array(fr, device=dev, as_numpy=True)
            assert np.abs(out - O.extract_condensed_frame_tensor_array(fr, 40)).max() <= PRE_TOL
            y = syn.make_audio(4, 8000, 16000, seed=5)
            a = U.extract_audio_features_array(y, 16000, 4, 30, device=dev, as_numpy=True)
            np.testing.assert_allclose(a, O.extract_audio_features_array(y, 16000, 4, 30), rtol=MFCC_RTOL, atol=MFCC_ATOL)
            v, w = [10, 40, 30, 50], [5, 4, 6, 3]
            assert U.knapsack(v, w, 10, device=dev) == O.knapsack(v, w, 10)
            m = _model(U, weights)
            vis = torch.from_numpy(out).to(dev)
            s = m(torch.from_numpy(a).to(dev)[:3], vis).cpu().numpy()
            ref = O.avm_forward(weights, O.extract_audio_features_array(y, 16000, 4, 30).astype(np.float32)[:3],
                                O.extract_condensed_frame_tensor_array(fr, 40)).numpy()
            assert_scores(s, ref)
            del m
        finally:
            if saved is not None:
                _lib._handles[idx] = saved
            else:
                del _lib._handles[idx]
            torch.cuda.synchronize(dev)
            lib.gn_destroy(p)
