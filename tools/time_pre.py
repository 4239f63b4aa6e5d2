"""Time gn_preprocess_u8_bgr for a (H, W, S) on the device: python tools/time_pre.py H W S n"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cvml_goalnet_b200 import utils as U

H, W, S, n = (int(x) for x in sys.argv[1:5])
dev = torch.device("cuda", 0)
fr = torch.randint(0, 256, (n, H, W, 3), dtype=torch.uint8, device=dev)
out = torch.empty((n, 3, S, S), dtype=torch.float32, device=dev)
for _ in range(3):
    U.extract_condensed_frame_tensor_array(fr, S, out=out)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    U.extract_condensed_frame_tensor_array(fr, S, out=out)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
print(f"H={H} W={W} S={S} n={n} L2={os.environ.get('GN_PREPROCESS_L2')}: {ms:.3f} ms  {n * (H * W * 3 + S * S * 12) / ms / 1e6:.0f} GB/s")
