"""Score spread vs arithmetic error of the CUDA forward on config-2-shaped inputs (run on the GPU box).
For each last-layer gain: range of the oracle's scores, rounding classes, and the per-element relative error of the
CUDA scores against the fp32 CPU oracle.  Prints one JSON line per gain; used to choose the weights the parity tests
and the benchmark run with."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cvml_goalnet_b200 import synthetic as syn, utils as U  # noqa: E402
from oracle import goalnet_oracle as O  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 216
torch.set_num_threads(os.cpu_count() or 1)
frames = syn.make_frames(n, 224, 224, seed=9)
y = syn.make_audio(n, 8000, 16000, seed=9)
vis_o = O.extract_condensed_frame_tensor_array(frames, 40)
aud_o = O.extract_audio_features_array(y, 16000, n, 30).astype(np.float32)
vis = U.extract_condensed_frame_tensor_array(frames)
aud = U.extract_audio_features_array(y, 16000, n, 30)
for gain in (1.0, 30.0, 100.0, 300.0):
    w = syn.make_weights(seed=0, score_gain=gain)
    m = U.AVM(True)
    m.load_state_dict(w)
    s = m(aud, vis).cpu().numpy()
    ref = O.avm_forward(w, aud_o, vis_o).numpy()
    rel = np.abs(s - ref) / np.abs(ref)
    print(json.dumps({"gain": gain, "n": n, "ref_min": float(ref.min()), "ref_max": float(ref.max()),
                      "classes": np.unique(np.round(ref)).tolist(), "max_rel": float(rel.max()),
                      "mean_rel": float(rel.mean()), "max_abs": float(np.abs(s - ref).max())}), flush=True)
