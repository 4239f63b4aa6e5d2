import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from cvml_goalnet_b200 import utils as U
from oracle import goalnet_oracle as O
rng = np.random.default_rng(11)
n = int(sys.argv[1]); cap = int(sys.argv[2]); wlo = int(sys.argv[3]); whi = int(sys.argv[4])
v = rng.integers(1, 1000, n).tolist()
w = rng.integers(wlo, whi, n).tolist()
try:
    got = U.knapsack(v, w, cap, 1)
    torch.cuda.synchronize()
    print(sys.argv[1:], os.environ.get("GN_KNAP_CL"), got == O.knapsack_rows(v, w, cap, 1))
except Exception as e:
    print(sys.argv[1:], os.environ.get("GN_KNAP_CL"), "EXC", str(e)[:120])
