import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from cvml_goalnet_b200 import synthetic as syn, utils as U
from cvml_goalnet_b200.train import FusionTrainer
from oracle import goalnet_oracle as O
m = U.AVM(True); w = syn.make_weights(seed=0); m.load_state_dict(w)
sizes = [512 * 640, 512, 512 * 512, 512, 256 * 512, 256, 128 * 256, 128, 128, 1]
for n, fixed in [(300, True), (320, True), (300, False), (4096, True), (4096, False)]:
    rng = np.random.default_rng(100 + n)
    x = np.abs(rng.standard_normal((n, 640))).astype(np.float32)
    lab = rng.integers(1, 6, n).astype(np.float32)
    tr = FusionTrainer(m, max_rows=4096, fixed_loss=fixed, dropout_p=0.0, process_group=False)
    _, ref, pred, _ = O.fusion_train_steps(w, x, lab, steps=1, fixed_loss=fixed)
    tr.forward_backward(x, lab)
    g = tr.grads.cpu().numpy()
    off = 0; errs = []
    for s in sizes:
        a, b = g[off:off + s].astype(np.float64), ref[off:off + s].astype(np.float64)
        errs.append(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-12)); off += s
    print(n, fixed, " ".join(f"{e:.3f}" for e in errs), "pred err", np.abs(tr.scores[:n].cpu().numpy() - pred).max())
