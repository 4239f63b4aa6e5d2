"""GPU probe: per-layer error of the AVM forward (tensor-core or CUDA-core validation path) vs the oracle.
usage: python tools/probe_model.py [tc|simt] [n]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from cvml_goalnet_b200 import synthetic as syn, utils as U
from oracle import goalnet_oracle as O

path = sys.argv[1] if len(sys.argv) > 1 else "tc"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 5
w = syn.make_weights(seed=0)
vis = O.extract_condensed_frame_tensor_array(syn.make_frames(n, 48, 48, seed=3), 40)
rng = np.random.default_rng(0)
aud = (25.0 * rng.standard_normal((n, 30, 30))).astype(np.float32); aud[:, 0] -= 250.0
ref, feats = O.avm_forward(w, aud, vis, return_features=True)
m = U.AVM(True); m.load_state_dict(w)
print("running", path, "n =", n, flush=True)
y = m(aud, vis, return_features=True, _simt=(path == "simt"))
torch.cuda.synchronize()
print("forward done", flush=True)
def rel(a, b): return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-9))
for which, key in ((1, "vis1"), (2, "vis2"), (3, "vis3")):
    a = m.read_activation(which, n).cpu().numpy()
    print(key, "rel err", rel(a, feats[key].numpy()), "finite", bool(np.isfinite(a).all()), flush=True)
h0 = m.read_activation(4, n).cpu().numpy()
print("audio feats rel", rel(h0[:, :128], feats["audio_features"].numpy()))
print("visual feats rel", rel(h0[:, 128:], feats["visual_features"].numpy()))
print("vf(out) rel", rel(m.last_visual_features.cpu().numpy(), feats["visual_features"].numpy()))
print("scores", y.flatten()[:5].tolist(), "ref", ref.flatten()[:5].tolist())
print("scores rel", rel(y.cpu().numpy(), ref.numpy()))
