"""Fusion-network training step on the B200 (BASELINE config 4).

Host mirror of the reference's train step for `AVM.fusion` (main.py:68-70,187-193):

    criterion = nn.MSELoss(); optimizer = optim.Adam(params = model.parameters(), lr = 0.001)
    optimizer.zero_grad(); pred = model(...); loss = criterion(pred, labels); loss.backward(); optimizer.step()

`FusionTrainer.step(features, labels)` does exactly that for the five Linear layers of `model.fusion` on
precomputed fusion inputs (the (N, 640) rows `torch.cat((audio_features, visual_features))` of utils.py:266).
All arithmetic is in libgoalnet_b200.so (tcgen05 GEMMs for forward / dgrad / wgrad, fused Adam); this module
only owns buffers and, for data-parallel training, the NCCL all-reduce of the flat gradient vector.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from .utils import AVM, _device, _ptr, _stream, _to_dev

FUSION_LAYERS = ("fusion.0", "fusion.3", "fusion.6", "fusion.9", "fusion.12")
HIDDEN_WIDTHS = (512, 512, 256, 128)


class FusionTrainer:
    """Adam training of `model.fusion` (reference defaults: lr 1e-3, betas (0.9, 0.999), eps 1e-8).

    fixed_loss=False reproduces the reference's loss as written: MSELoss(pred (N,1), labels (N,)) broadcasts to
    N x N (quirk Q5).  dropout_p is the reference's 0.2 (utils.py:245-254); 0 disables it.
    process_group: torch.distributed group for data-parallel training (None = the default group when
    torch.distributed is initialised, False = never reduce); gradients are averaged over its ranks.
    """

    def __init__(self, model: AVM, max_rows: int, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 fixed_loss: bool = False, dropout_p: float = 0.2, device=None, process_group=None, seed: int = 12344321):
        self.dev = _device(device)
        self.model = model
        self.fin = 640 if model.audio_included else 512
        self.fixed_loss = bool(fixed_loss)
        self.dropout_p = float(dropout_p)
        self.group = process_group
        self.seed = int(seed)
        self.steps = 0
        self.max_rows = int(max_rows)
        lib = _lib.load()
        sd = model.state_dict()
        keep = []
        st = _lib.gn_fusion_state()
        for k, name in enumerate(FUSION_LAYERS):
            w = sd[name + ".weight"].detach().to(self.dev, torch.float32).contiguous()
            b = sd[name + ".bias"].detach().to(self.dev, torch.float32).contiguous()
            keep += [w, b]
            st.weight[k] = w.data_ptr()
            st.bias[k] = b.data_ptr()
        p = C.c_void_p()
        with torch.cuda.device(self.dev):
            _lib.check(lib.gn_fusion_trainer_create(_lib.handle(self.dev.index), C.byref(p), C.byref(st), self.fin,
                                                    self.max_rows, _stream(self.dev)), "gn_fusion_trainer_create")
            torch.cuda.current_stream(self.dev).synchronize()      # `keep` may be freed after this
        self._t = p
        _lib.check(lib.gn_fusion_set_hyper(self._t, lr, betas[0], betas[1], eps), "gn_fusion_set_hyper")
        self.n_params = int(lib.gn_fusion_param_count(self._t))
        self.grads = torch.zeros(self.n_params, dtype=torch.float32, device=self.dev)
        self.scores = torch.empty(self.max_rows, dtype=torch.float32, device=self.dev)
        self.loss = torch.zeros(1, dtype=torch.float64, device=self.dev)

    def __del__(self):
        try:
            if getattr(self, "_t", None):
                _lib.load().gn_fusion_trainer_destroy(self._t)
                self._t = None
        except Exception:
            pass

    # ------------------------------------------------------------------------------------------------
    def _mask_ptrs(self, masks):
        if masks is None:
            return None, None
        assert len(masks) == 4, "E: four dropout masks expected (after fusion.1, .4, .7, .10)"
        keep = [_to_dev(m, torch.uint8, self.dev) for m in masks]
        arr = (C.c_void_p * 4)(*[m.data_ptr() for m in keep])
        return arr, keep

    def forward_backward(self, features, labels, masks=None):
        """Forward + loss + backward; leaves the flat gradient in `self.grads` (this rank's, un-reduced)."""
        x = _to_dev(features, torch.float32, self.dev)
        lab = _to_dev(labels, torch.float32, self.dev).reshape(-1)
        assert x.dim() == 2 and x.shape[1] == self.fin, "E: features must have shape (N, %d)" % self.fin
        assert lab.numel() == x.shape[0], "E: Inconsistency in data loader definition"
        n = int(x.shape[0])
        arr, keep = self._mask_ptrs(masks)
        marg = C.cast(arr, C.POINTER(C.c_void_p)) if arr is not None else None
        seed = (self.seed + 0x9E3779B97F4A7C15 * (self.steps + 1)) & 0xFFFFFFFFFFFFFFFF
        with torch.cuda.device(self.dev):
            _lib.check(_lib.load().gn_fusion_forward_backward(
                self._t, _ptr(x), _ptr(lab), n, int(self.fixed_loss), self.dropout_p, marg, seed, _ptr(self.grads),
                _ptr(self.scores), _ptr(self.loss), _stream(self.dev)), "gn_fusion_forward_backward")
        return n

    def optimizer_step(self):
        """All-reduce (data parallel) and Adam."""
        scale = 1.0
        import torch.distributed as dist
        if self.group is not False and dist.is_available() and dist.is_initialized():
            group = self.group if self.group is not None else dist.group.WORLD
            world = dist.get_world_size(group)
            if world > 1:
                dist.all_reduce(self.grads, op=dist.ReduceOp.SUM, group=group)
                scale = 1.0 / world
        with torch.cuda.device(self.dev):
            _lib.check(_lib.load().gn_fusion_adam_step(self._t, _ptr(self.grads), scale, _stream(self.dev)),
                       "gn_fusion_adam_step")
        self.steps += 1

    def step(self, features, labels, masks=None):
        """One train step (main.py:187-193).  Returns (loss tensor float64[1] on the device, predictions (N,1))."""
        n = self.forward_backward(features, labels, masks)
        self.optimizer_step()
        return self.loss, self.scores[:n].reshape(n, 1)

    # ------------------------------------------------------------------------------------------------
    def flat_params(self) -> torch.Tensor:
        out = torch.empty(self.n_params, dtype=torch.float32, device=self.dev)
        with torch.cuda.device(self.dev):
            _lib.check(_lib.load().gn_fusion_get_params(self._t, _ptr(out), _stream(self.dev)), "gn_fusion_get_params")
        return out

    def sync_model(self):
        """Write the trained parameters back into `model` (state_dict keys fusion.{0,3,6,9,12}.*)."""
        flat = self.flat_params().cpu()
        sd = self.model.state_dict()
        off = 0
        for name in FUSION_LAYERS:
            for suffix in (".weight", ".bias"):
                t = sd[name + suffix]
                t.copy_(flat[off:off + t.numel()].reshape(t.shape))
                off += t.numel()
        self.model.invalidate()
        return self.model


def fusion_flat(state_dict) -> np.ndarray:
    """Parameters of AVM.fusion as one float32 vector in `model.fusion.parameters()` order."""
    parts = []
    for name in FUSION_LAYERS:
        for suffix in (".weight", ".bias"):
            v = state_dict[name + suffix]
            parts.append((v.detach().cpu().numpy() if torch.is_tensor(v) else np.asarray(v)).astype(np.float32).ravel())
    return np.concatenate(parts)
