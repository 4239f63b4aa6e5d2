"""Generate tests/golden/*.npz by running the UNMODIFIED reference code (build container only).

    PYTHONBREAKPOINT=0 python -m oracle.gen_golden

TEST INFRASTRUCTURE ONLY.  Every vector below is produced by a function imported from
`/root/reference/utils.py` (see oracle/ref_import.py for the four stubbed I/O modules); nothing
from the oracle restatement feeds the outputs except `mfcc`, which stands in for the absent
`librosa.feature.mfcc` when the reference's own `extract_audio_features` is executed (that stage is
"parity unpinned", see oracle/goalnet_oracle.py header).  Inputs come from the seeded generators in
`cvml_goalnet_b200/synthetic.py` so that tests can rebuild the big ones (weights) from the seed.
"""
from __future__ import annotations

import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ["PYTHONBREAKPOINT"] = "0"

from oracle.ref_import import import_reference_utils  # noqa: E402
from oracle import goalnet_oracle as O  # noqa: E402
from cvml_goalnet_b200 import synthetic as syn  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def gen_preprocess(R):
    """utils.py:274-292 through a real (lossless FFV1) video file, plus the bare arithmetic."""
    import cv2
    out = {}
    cases = {"down": (5, 72, 96), "down2": (3, 135, 240), "up": (3, 24, 32), "flat": (2, 48, 48)}
    for name, (n, H, W) in cases.items():
        fr = syn.make_frames(n, H, W, seed=hash(name) % 1000 if False else len(name))
        if name == "flat":
            fr[0][:] = 17           # max == min -> the +1e-7 guard (utils.py:284)
        out[f"{name}_frames"] = fr
        # literal arithmetic of utils.py:284-285,292 on in-memory frames
        lit = []
        for image in fr:
            image = ((image - image.min()) / (image.max() - image.min() + 1e-7)).astype(np.float32)
            image = cv2.resize(image, (40, 40))
            lit.append(image)
        out[f"{name}_out"] = np.transpose(np.array(lit), axes=(0, 3, 1, 2))
    # the whole reference function on a lossless file, skip_frames = 2
    fr = syn.make_frames(7, 64, 80, seed=99)
    with tempfile.TemporaryDirectory() as td:
        fp = os.path.join(td, "clip.avi")
        vw = cv2.VideoWriter(fp, cv2.VideoWriter_fourcc(*"FFV1"), 30, (80, 64))
        ok = vw.isOpened()
        if ok:
            for f in fr:
                vw.write(f)
            vw.release()
            cap = cv2.VideoCapture(fp)
            dec = []
            while True:
                s, im = cap.read()
                if not s:
                    break
                dec.append(im)
            cap.release()
            if len(dec) == 7:
                res, full_n = R.extract_condensed_frame_tensor(fp, skip_frames=2)
                out["file_decoded_frames"] = np.array(dec)
                out["file_out"] = res
                out["file_full_n_frames"] = np.int64(full_n)
                out["file_skip_frames"] = np.int64(2)
            else:
                ok = False
    print("preprocess: video-file case", "written" if ok else "SKIPPED (no FFV1 writer)")
    np.savez_compressed(os.path.join(GOLD, "preprocess.npz"), **out)


def gen_model(R):
    """AVM.forward (utils.py:260-272), eval mode, seeded synthetic weights, with and without audio."""
    import torch
    out = {}
    rng = np.random.default_rng(syn.SEED_BASE + 11)
    N = 6
    vis = O.extract_condensed_frame_tensor_array(syn.make_frames(N, 56, 56, seed=5), 40, literal=True)
    aud = (30.0 * rng.standard_normal((N, 30, 30))).astype(np.float32)
    aud[:, 0, :] -= 300.0  # MFCC c0 is large and negative for real audio
    out["visual_input"] = vis
    out["audio_input"] = aud
    for tag, audio_included, gain in (("av", True, 1.0), ("v", False, 1.0), ("av_gain", True, 100.0)):
        w = syn.make_weights(seed=0, audio_included=audio_included, score_gain=gain)
        m = R.AVM(audio_included=audio_included)
        m.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in w.items()})
        m.eval()
        feats = {}
        m.visbl.register_forward_hook(lambda mod, i, o: feats.__setitem__("visual_features", o.detach().numpy().copy()))
        m.visbl.bnorm1.register_forward_hook(lambda mod, i, o: feats.__setitem__("vis1", o.detach().numpy().copy()))
        m.visbl.bnorm2.register_forward_hook(lambda mod, i, o: feats.__setitem__("vis2", o.detach().numpy().copy()))
        m.visbl.bnorm3.register_forward_hook(lambda mod, i, o: feats.__setitem__("vis3", o.detach().numpy().copy()))
        if audio_included:
            m.audbl.register_forward_hook(lambda mod, i, o: feats.__setitem__("audio_features", o.detach().numpy().copy()))
        with torch.no_grad():
            y = m(torch.from_numpy(aud), torch.from_numpy(vis))
        out[f"{tag}_scores"] = y.numpy()
        for k, v in feats.items():
            if k in ("vis1", "vis2", "vis3"):
                v = v[:2]  # keep fixtures small
            out[f"{tag}_{k}"] = v
        # state_dict key/shape contract
        out[f"{tag}_keys"] = np.array(sorted(m.state_dict().keys()))
    np.savez_compressed(os.path.join(GOLD, "model.npz"), **out)


def gen_integer_path(R):
    """expand_array, get_clip_information, knapsack, get_fscore (utils.py:396-410,445-510,552-580)."""
    out = {}
    rng = np.random.default_rng(syn.SEED_BASE + 23)
    # expand_array: truncate, pad-with-last, identity
    for name, (n, rate, length) in {"trunc": (7, 4, 26), "pad": (7, 4, 33), "ident": (9, 4, 9), "exact": (5, 3, 15)}.items():
        arr = rng.integers(1, 6, size=n).tolist()
        out[f"expand_{name}_in"] = np.array(arr)
        out[f"expand_{name}_args"] = np.array([rate, length])
        out[f"expand_{name}_out"] = np.array(R.expand_array(list(arr), rate, length))
    # knapsack cases: random, ties (all equal values), zero-length clips, nothing fits, empty
    cases = []
    for t in range(12):
        n = int(rng.integers(1, 40))
        values = rng.integers(0, 60, size=n).tolist()
        weights = rng.integers(1, 30, size=n).tolist()
        cap = int(rng.integers(1, max(2, sum(weights) // 2)))
        cases.append((values, weights, cap))
    cases.append(([3] * 10, [4] * 10, 13))                   # ties
    cases.append(([5, 0, 7, 9], [0, 3, 0, 4], 6))            # zero-weight clips (quirk Q11)
    cases.append(([5, 6], [50, 60], 10))                     # nothing fits
    cases.append(([], [], 10))                               # empty
    cases.append(([4, 4, 4], [1, 1, 1], 0))                  # zero capacity
    cases.append((rng.integers(30, 1500, 60).tolist(), rng.integers(30, 301, 60).tolist(), 1500))
    for i, (v, w, c) in enumerate(cases):
        out[f"knap_{i}_values"] = np.array(v, dtype=np.int64)
        out[f"knap_{i}_weights"] = np.array(w, dtype=np.int64)
        out[f"knap_{i}_cap"] = np.int64(c)
        out[f"knap_{i}_sel"] = np.array(R.knapsack(list(v), list(w), c), dtype=np.int64)
    out["knap_n"] = np.int64(len(cases))
    # get_clip_information
    imp = rng.integers(1, 6, size=500).tolist()
    cps = syn.make_change_points(500, seed=3, lo=5, hi=60)
    v, l, _ = R.get_clip_information(cps.tolist(), imp)
    out["clip_imp"] = np.array(imp)
    out["clip_cps"] = cps
    out["clip_values"] = np.array(v, dtype=np.int64)
    out["clip_lengths"] = np.array(l, dtype=np.int64)
    # get_fscore: the unmodified function on int64 arrays == NumPy-1 semantics (quirk Q6)
    G = (rng.random((20, 700)) < 0.2).astype(np.int64)
    G[3] = 0
    S = (rng.random(700) < 0.15).astype(np.int64)
    fa, fm = R.get_fscore(G, S)
    out["fs_G"] = G.astype(np.uint8)
    out["fs_S"] = S.astype(np.uint8)
    out["fs_avg_max"] = np.array([fa, fm], dtype=np.float64)
    fa0, fm0 = R.get_fscore(G, np.zeros(700, dtype=np.int64))
    out["fs_empty_avg_max"] = np.array([fa0, fm0], dtype=np.float64)
    np.savez_compressed(os.path.join(GOLD, "integer_path.npz"), **out)


def gen_postprocess(R):
    """utils.py:606-643 executed literally with an in-memory stand-in for the two HDF5 files."""
    import torch
    out = {}
    cases = {"a": (120, 15, 1800, 20, 200), "b": (77, 30, 2290, 20, 200), "c": (50, 1, 50, 2, 8)}
    for name, (N, skip, n_raw, lo, hi) in cases.items():
        scores = syn.make_scores(N, seed=len(out))
        cps = syn.make_change_points(n_raw, seed=len(out) + 1, lo=lo, hi=hi)

        class FakeFile(dict):
            def __enter__(self):
                return self

            def __exit__(self, *a):
                return False

        R.h5py.File = lambda *a, **k: FakeFile({"video_1": {"change_points": cps}})
        R.get_video_data_from_h5 = lambda fp: [["video_1", int(cps[-1][1])]]
        R.get_video_data_from_mat = lambda fp: [["vid", int(cps[-1][1]) + 1]]
        _, ind = R.postprocess(video_id="vid", h5_file_path="x.h5", mat_file_path="x.mat",
                               batch_importances=torch.from_numpy(scores), skip_frames=skip,
                               full_n_frames=n_raw, full_frames=None)
        out[f"{name}_scores"] = scores
        out[f"{name}_cps"] = cps
        out[f"{name}_args"] = np.array([skip, n_raw])
        out[f"{name}_indicator"] = ind
    np.savez_compressed(os.path.join(GOLD, "postprocess.npz"), **out)


def gen_audio(R):
    """utils.py:313-349 executed literally; librosa.load / librosa.feature.mfcc are the stub + oracle mfcc."""
    out = {}
    sr = 16000
    for name, (n_frames, spf, B) in {"int": (4, 8000, 30), "ragged": (5, 7777, 30), "b16": (3, 8000, 16)}.items():
        y = syn.make_audio(1, n_frames * spf + (3 if name == "ragged" else 0), sr, seed=len(out))
        R.librosa.load = lambda fp, _y=y: (_y, sr)
        R.librosa.feature.mfcc = lambda y, sr, n_mfcc: O.mfcc(y, sr, n_mfcc)
        feats = R.extract_audio_features("x.wav", n_frames, B)
        out[f"{name}_y"] = y
        out[f"{name}_args"] = np.array([sr, n_frames, B])
        out[f"{name}_features"] = feats
    # mfcc restatement snapshot (the definition of the unpinned stage), one segment
    y = syn.make_audio(1, 8000, sr, seed=42)
    out["mfcc_y"] = y
    out["mfcc_out"] = O.mfcc(y, sr)
    np.savez_compressed(os.path.join(GOLD, "audio.npz"), **out)


def gen_train(R):
    """The reference's train step (main.py:68-70,187-193) applied to AVM.fusion, run with the REFERENCE model object:
    eval-mode forward (Dropout identity, so that the vectors are reproducible), nn.MSELoss as written (pred (N,1) vs
    labels (N,): N x N broadcast, quirk Q5), optim.Adam(lr=0.001).  Features are what `fusion` receives."""
    import torch
    import warnings
    out = {}
    rng = np.random.default_rng(syn.SEED_BASE + 41)
    for tag, audio_included, n in (("av", True, 96), ("v", False, 40)):
        w = syn.make_weights(seed=0, audio_included=audio_included)
        m = R.AVM(audio_included=audio_included)
        m.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in w.items()})
        m.eval()
        fin = 640 if audio_included else 512
        x = np.abs(rng.standard_normal((n, fin))).astype(np.float32)      # fusion inputs are post-ReLU features
        lab = rng.integers(1, 6, n).astype(np.float32)
        opt = torch.optim.Adam(params=m.fusion.parameters(), lr=0.001)
        crit = torch.nn.MSELoss()
        losses = []
        for step in range(3):
            opt.zero_grad()
            pred = 4 * m.fusion(torch.from_numpy(x)) + 1                       # utils.py:269-270
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                loss = crit(pred, torch.from_numpy(lab))                       # main.py:191
            loss.backward()
            if step == 0:
                g = np.concatenate([p.grad.numpy().ravel() for p in m.fusion.parameters()])
                out[f"{tag}_grads_s97"] = g[::97].copy()                        # strided sample keeps the fixture small
                out[f"{tag}_grads_l2"] = np.array([np.sqrt(np.sum(np.square(p.grad.numpy().astype(np.float64))))
                                                   for p in m.fusion.parameters()])
                out[f"{tag}_pred"] = pred.detach().numpy()[:, 0].copy()
            opt.step()
            losses.append(float(loss.item()))
        out[f"{tag}_x"] = x
        out[f"{tag}_labels"] = lab
        out[f"{tag}_losses"] = np.array(losses)
        out[f"{tag}_params3_s97"] = np.concatenate([p.detach().numpy().ravel() for p in m.fusion.parameters()])[::97].copy()
    np.savez_compressed(os.path.join(GOLD, "train.npz"), **out)


def gen_train_avm(R):
    """The reference's train sub-step (main.py:187-193) on the WHOLE reference AVM object in train mode (BatchNorm
    batch statistics + running-buffer updates, as the reference runs it), nn.MSELoss as written, optim.Adam(lr=0.001)
    over model.parameters() (main.py:70), sub-batch of 10 frames (main.py:47).  Dropout probabilities are set to 0 so
    that the vectors are reproducible (mask semantics are covered by the oracle <-> CUDA tests)."""
    import torch
    import warnings
    out = {}
    rng = np.random.default_rng(syn.SEED_BASE + 43)
    for tag, audio_included, n in (("av", True, 10), ("v", False, 7)):
        w = syn.make_weights(seed=1, audio_included=audio_included)
        m = R.AVM(audio_included=audio_included)
        vis = rng.random((n, 3, 40, 40)).astype(np.float32)
        aud = (25.0 * rng.standard_normal((n, 30, 30))).astype(np.float32)
        lab = rng.integers(1, 6, n).astype(np.float32)
        with torch.no_grad():
            m(torch.from_numpy(aud), torch.from_numpy(vis))                    # materialise the Lazy* layers
        m.load_state_dict({k: torch.from_numpy(np.asarray(v)) for k, v in w.items()})
        m.train()
        for mod in m.modules():
            if isinstance(mod, torch.nn.Dropout):
                mod.p = 0.0
        opt = torch.optim.Adam(params=m.parameters(), lr=0.001)
        crit = torch.nn.MSELoss()
        losses = []
        for step in range(2):
            opt.zero_grad()
            pred = m(torch.from_numpy(aud), torch.from_numpy(vis))
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                loss = crit(pred, torch.from_numpy(lab))
            loss.backward()
            if step == 0:
                out[f"{tag}_pred"] = pred.detach().numpy()[:, 0].copy()
                for k, p_ in m.named_parameters():
                    g = p_.grad.numpy()
                    out[f"{tag}_gl2_{k}"] = np.array(np.sqrt(np.sum(np.square(g.astype(np.float64)))))
                    out[f"{tag}_gs_{k}"] = g.ravel()[::997].copy()
            opt.step()
            losses.append(float(loss.item()))
        out[f"{tag}_vis"], out[f"{tag}_aud"], out[f"{tag}_labels"] = vis, aud, lab
        out[f"{tag}_losses"] = np.array(losses)
        for k, v in m.state_dict().items():
            if "num_batches" in k:
                continue
            out[f"{tag}_p2_{k}"] = v.detach().numpy().ravel()[::997].copy()
    np.savez_compressed(os.path.join(GOLD, "train_avm.npz"), **out)


def main():
    os.makedirs(GOLD, exist_ok=True)
    R = import_reference_utils()
    gen_preprocess(R)
    gen_model(R)
    gen_integer_path(R)
    gen_postprocess(R)
    gen_audio(R)
    gen_train(R)
    gen_train_avm(R)
    for f in sorted(os.listdir(GOLD)):
        print(f, os.path.getsize(os.path.join(GOLD, f)))


if __name__ == "__main__":
    main()
