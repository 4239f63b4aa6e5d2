"""Run the config-2 selection (clip statistics, knapsack, indicator) a few times on synthetic scores: the command ncu wraps
when the knapsack kernels are profiled (`ncu -k regex:knapsack ... python tools/run_select.py`)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cvml_goalnet_b200 import synthetic as syn, utils as U  # noqa: E402

n, skip = 10800, 15
n_raw = n * skip
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
dev = torch.device("cuda", 0)
cps = syn.make_change_points(n_raw, seed=1)
g = torch.Generator(device=dev)
g.manual_seed(3)
scores = 3.0 + torch.rand((n, 1), device=dev, generator=g)
bufs = U.SelectBuffers(int(cps.shape[0]), n_raw, 1, dev)
for _ in range(reps):
    _, ind = U.postprocess_array(scores, skip, n_raw, cps, device=dev, return_device=True, bufs=bufs)
torch.cuda.synchronize()
print("ok", int(cps.shape[0]), int(ind.sum()))
