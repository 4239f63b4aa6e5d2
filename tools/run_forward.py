"""Run the AVM forward (and optionally the front kernels) a few times on config-2-sized synthetic inputs: the command ncu
wraps when one kernel of the forward is profiled (`ncu -k regex:... -s N -c 1 python tools/run_forward.py`)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cvml_goalnet_b200 import synthetic as syn, utils as U  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10800
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
front = len(sys.argv) > 3 and sys.argv[3] == "front"
dev = torch.device("cuda", 0)
m = U.AVM(True)
m.load_state_dict(syn.make_weights(seed=0, score_gain=syn.SPREAD_GAIN))
g = torch.Generator(device=dev)
g.manual_seed(1)
vis = torch.rand((n, 3, 40, 40), device=dev, generator=g)
aud = 25.0 * torch.randn((n, 30, 30), device=dev, generator=g)
out = torch.empty((n, 1), device=dev)
if front:
    frames = torch.randint(0, 256, (n, 224, 224, 3), device=dev, dtype=torch.uint8, generator=g)
    y = torch.randn(n * 8000, device=dev, generator=g) * 0.1
for _ in range(reps):
    if front:
        U.extract_condensed_frame_tensor_array(frames, 40, out=vis)
        U.extract_audio_features_array(y, 16000, n, 30, out=aud)
    m(aud, vis, out=out)
torch.cuda.synchronize()
print("ok", float(out.mean()))
