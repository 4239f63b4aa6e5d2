"""goalnet_b200: B200-native (sm_100a) per-frame multimodal scoring path of CVML-GoalNet.

`cvml_goalnet_b200.utils` mirrors the reference's `utils.py` names; `cvml_goalnet_b200.main` mirrors
`main.py`'s inference entry points; `cvml_goalnet_b200.synthetic` defines the benchmark workloads.
The arithmetic lives in csrc/ (hand-written CUDA behind the C-ABI of include/goalnet_b200.h).
"""
__version__ = "0.1.0"
