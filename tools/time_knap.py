"""Device time of gn_knapsack alone for a few table shapes (CUDA events, 20 repetitions).

    python tools/time_knap.py            # rows x columns sweep; honours GN_KNAP_T / GN_KNAP_CL
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from cvml_goalnet_b200.utils import _knapsack_device

dev = torch.device("cuda", 0)
rng = np.random.default_rng(3)
for n, cap, wlo, whi in [(984, 24300, 30, 301), (492, 24300, 30, 301), (984, 12150, 30, 301), (984, 24300, 300, 3000),
                         (246, 24300, 30, 301), (984, 24300, 3, 31)]:
    v = torch.from_numpy(rng.integers(1, 1500, n).astype(np.int32)).to(dev)
    w = torch.from_numpy(rng.integers(wlo, whi, n).astype(np.int32)).to(dev)
    for _ in range(3):
        sel, best = _knapsack_device(v, w, n, cap, 0, 1, dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(20):
        sel, best = _knapsack_device(v, w, n, cap, 0, 1, dev)
    e1.record()
    torch.cuda.synchronize()
    print(f"T={os.environ.get('GN_KNAP_T', 'default')} n={n} cap={cap} w=[{wlo},{whi}) selected={int(sel.sum())} "
          f"{e0.elapsed_time(e1) / 20 * 1e3:.1f} us")
