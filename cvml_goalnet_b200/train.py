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


class _GraphedStep:
    """A whole train step (forward, loss, backward, gradient all-reduce, Adam) captured ONCE in a CUDA graph and replayed:
    the step is a chain of 30-90 small dependent launches, so issuing it as one graph launch removes the per-launch host
    cost that bounded it.  What changes from step to step lives on the device (`gn_step_state`: Adam step count, bias
    corrections, dropout seed bump), so a replay is a new step, not a repetition.  Inputs are copied into static buffers
    (same addresses every replay); the first `warmup` calls run eagerly (they also set kernel attributes and let NCCL
    build its communicator outside the capture)."""

    def __init__(self, dev, warmup: int = 2):
        self.dev, self.warmup, self.calls, self.graphs = dev, warmup, 0, {}

    def run(self, key, eager_fn):
        """eager_fn() must launch the step on the current stream reading only static buffers."""
        self.calls += 1
        if self.calls <= self.warmup:
            eager_fn()
            return
        ent = self.graphs.get(key)
        if ent is None:
            torch.cuda.synchronize(self.dev)
            g = torch.cuda.CUDAGraph()
            l0 = int(_lib.load().gn_launch_count(_lib.handle(self.dev.index)))
            with torch.cuda.graph(g):
                eager_fn()
            captured = int(_lib.load().gn_launch_count(_lib.handle(self.dev.index))) - l0
            # the capture only RECORDED those launches: take them back out, every replay adds them
            _lib.add_graph_launches(self.dev.index, -captured)
            ent = self.graphs[key] = (g, captured)
        ent[0].replay()
        _lib.add_graph_launches(self.dev.index, ent[1])


class FusionTrainer:
    """Adam training of `model.fusion` (reference defaults: lr 1e-3, betas (0.9, 0.999), eps 1e-8).

    fixed_loss=False reproduces the reference's loss as written: MSELoss(pred (N,1), labels (N,)) broadcasts to
    N x N (quirk Q5).  dropout_p is the reference's 0.2 (utils.py:245-254); 0 disables it.
    With a `CAVM` model (fusion.12 = Linear(128,5) + Softmax, the reference's commented classifier variant) the loss is
    nn.CrossEntropyLoss: fixed_loss=False applies it to the model's softmax OUTPUT with target labels-1, exactly what
    un-commenting main.py:69,96 computes; fixed_loss=True applies it to the logits.  `scores` is then (N,5).
    process_group: torch.distributed group for data-parallel training (None = the default group when
    torch.distributed is initialised, False = never reduce); gradients are averaged over its ranks.
    """

    def __init__(self, model: AVM, max_rows: int, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 fixed_loss: bool = False, dropout_p: float = 0.2, device=None, process_group=None, seed: int = 12344321,
                 use_graph: bool = False):
        self.dev = _device(device)
        self.model = model
        self.use_graph = bool(use_graph)
        self._graph = _GraphedStep(self.dev)
        self.fin = 640 if model.audio_included else 512
        self.n_cls = int(getattr(model, "OUT_COLS", 1))
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
            _lib.check(lib.gn_fusion_trainer_create_head(_lib.handle(self.dev.index), C.byref(p), C.byref(st), self.fin,
                                                         self.max_rows, self.n_cls, _stream(self.dev)),
                       "gn_fusion_trainer_create_head")
            torch.cuda.current_stream(self.dev).synchronize()      # `keep` may be freed after this
        self._t = p
        _lib.check(lib.gn_fusion_set_hyper(self._t, lr, betas[0], betas[1], eps), "gn_fusion_set_hyper")
        self.n_params = int(lib.gn_fusion_param_count(self._t))
        self.grads = torch.zeros(self.n_params, dtype=torch.float32, device=self.dev)
        self.scores = torch.empty(self.max_rows * self.n_cls, dtype=torch.float32, device=self.dev)
        self.loss = torch.zeros(1, dtype=torch.float64, device=self.dev)
        if self.use_graph:
            self._gx = torch.zeros((self.max_rows, self.fin), dtype=torch.float32, device=self.dev)
            self._glab = torch.zeros(self.max_rows, dtype=torch.float32, device=self.dev)

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
                self._t, _ptr(x), _ptr(lab), n, int(self.fixed_loss) + (2 if self.n_cls > 1 else 0), self.dropout_p, marg,
                seed, _ptr(self.grads),
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
        if self.use_graph and masks is None:
            x = _to_dev(features, torch.float32, self.dev)
            lab = _to_dev(labels, torch.float32, self.dev).reshape(-1)
            n = int(x.shape[0])
            assert x.dim() == 2 and x.shape[1] == self.fin and lab.numel() == n and n <= self.max_rows, \
                "E: Inconsistency in data loader definition"
            self._gx[:n].copy_(x, non_blocking=True)
            self._glab[:n].copy_(lab, non_blocking=True)

            def body():
                self.forward_backward(self._gx[:n], self._glab[:n])
                self.optimizer_step()
            self._graph.run(n, body)
            return self.loss, self.scores[:n * self.n_cls].reshape(n, self.n_cls)
        n = self.forward_backward(features, labels, masks)
        self.optimizer_step()
        return self.loss, self.scores[:n * self.n_cls].reshape(n, self.n_cls)

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


DROPOUT_WIDTHS = (512, 512, 512, 256, 128)     # linear5 (utils.py:168), fusion.0/3/6/9 (utils.py:245-254)


class AVMTrainer:
    """The reference's train sub-step (main.py:187-193) for the WHOLE model in train mode.

    `step(audios, frames, labels)` = optimizer.zero_grad(); pred = model(audios, frames); loss = MSELoss()(pred,
    labels); loss.backward(); optimizer.step() with optim.Adam(model.parameters(), lr) -- BatchNorm uses the
    sub-batch statistics and updates its running buffers, Dropout(0.2) is active (quirk Q4), the loss broadcasts to
    N x N unless fixed_loss (quirk Q5).  Arithmetic: fp32 CUDA kernels in libgoalnet_b200.so (gn_train_avm.cu).
    `sync_model()` writes the trained parameters and running statistics back into `model`, whose eval-mode forward
    (the tensor-core scoring path) then uses them.  Data parallel: gradients are averaged over `process_group`.
    """

    def __init__(self, model: AVM, max_frames: int = 10, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 fixed_loss: bool = False, dropout_p: float = 0.2, device=None, process_group=None, seed: int = 12344321,
                 use_graph: bool = False):
        self.dev = _device(device)
        self.model = model
        self.use_graph = bool(use_graph)
        self._graph = _GraphedStep(self.dev)
        self.fixed_loss, self.dropout_p = bool(fixed_loss), float(dropout_p)
        self.group, self.seed, self.steps, self.max_frames = process_group, int(seed), 0, int(max_frames)
        lib = _lib.load()
        st, keep = self._state_struct(model.state_dict(), clone=False)
        p = C.c_void_p()
        with torch.cuda.device(self.dev):
            _lib.check(lib.gn_avm_trainer_create(_lib.handle(self.dev.index), C.byref(p), C.byref(st),
                                                 int(model.audio_included), model.S, model.B, self.max_frames,
                                                 _stream(self.dev)), "gn_avm_trainer_create")
            torch.cuda.current_stream(self.dev).synchronize()
        del keep
        self._t = p
        _lib.check(lib.gn_avm_trainer_set_hyper(self._t, lr, betas[0], betas[1], eps), "gn_avm_trainer_set_hyper")
        self.n_params = int(lib.gn_avm_trainer_param_count(self._t))
        g = C.c_void_p()
        _lib.check(lib.gn_avm_trainer_grads(self._t, C.byref(g)), "gn_avm_trainer_grads")
        self._grads_ptr = g.value
        self.scores = torch.empty(self.max_frames, dtype=torch.float32, device=self.dev)
        self.loss = torch.zeros(1, dtype=torch.float64, device=self.dev)
        if self.use_graph:
            self._gvis = torch.zeros((self.max_frames, 3, model.S, model.S), dtype=torch.float32, device=self.dev)
            self._gaud = torch.zeros((self.max_frames, 30, model.B), dtype=torch.float32, device=self.dev)
            self._glab = torch.zeros(self.max_frames, dtype=torch.float32, device=self.dev)

    def __del__(self):
        try:
            if getattr(self, "_t", None):
                _lib.load().gn_avm_trainer_destroy(self._t)
                self._t = None
        except Exception:
            pass

    def _state_struct(self, sd, clone):
        st, keep = _lib.gn_avm_state(), {}
        for field, key in _lib.STATE_KEYS.items():
            if key not in sd:
                setattr(st, field, None)
                continue
            t = sd[key].detach().to(device=self.dev, dtype=torch.float32).contiguous()
            if clone:
                t = t.clone()
            keep[key] = t
            setattr(st, field, t.data_ptr())
        return st, keep

    def grads(self) -> torch.Tensor:
        """Flat float32 gradient of the last forward_backward (a view of the library's buffer)."""
        n = self.n_params

        class _Holder:     # __cuda_array_interface__ view of library-owned memory
            pass
        h = _Holder()
        h.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (self._grads_ptr, False), "version": 2}
        return torch.as_tensor(h, device=self.dev)

    def forward_backward(self, audios, frames, labels, masks=None):
        vis = _to_dev(frames, torch.float32, self.dev)
        n = int(vis.shape[0])
        assert vis.dim() == 4 and tuple(vis.shape[1:]) == (3, self.model.S, self.model.S), "E: Invalid shape for visual input"
        assert 1 <= n <= self.max_frames, "E: sub-batch larger than max_frames"
        aud = None
        if self.model.audio_included:
            aud = _to_dev(audios, torch.float32, self.dev)
            assert tuple(aud.shape) == (n, 30, self.model.B), "E: Inconsistency in data loader definition"
        lab = _to_dev(labels, torch.float32, self.dev).reshape(-1)
        assert lab.numel() == n, "E: Inconsistency in data loader definition"
        marg, keep = None, None
        if masks is not None:
            assert len(masks) == 5, "E: five dropout masks expected (linear5, fusion.0/3/6/9)"
            keep = [_to_dev(m, torch.uint8, self.dev) for m in masks]
            for m, wdt in zip(keep, DROPOUT_WIDTHS):
                assert tuple(m.shape) == (n, wdt), "E: dropout mask shape"
            marg = C.cast((C.c_void_p * 5)(*[m.data_ptr() for m in keep]), C.POINTER(C.c_void_p))
        seed = (self.seed + 0x9E3779B97F4A7C15 * (self.steps + 1)) & 0xFFFFFFFFFFFFFFFF
        with torch.cuda.device(self.dev):
            _lib.check(_lib.load().gn_avm_trainer_forward_backward(
                self._t, _ptr(vis), _ptr(aud), _ptr(lab), n, int(self.fixed_loss), self.dropout_p, marg, seed,
                _ptr(self.scores), _ptr(self.loss), _stream(self.dev)), "gn_avm_trainer_forward_backward")
        return n

    def optimizer_step(self):
        scale = 1.0
        import torch.distributed as dist
        if self.group is not False and dist.is_available() and dist.is_initialized():
            group = self.group if self.group is not None else dist.group.WORLD
            world = dist.get_world_size(group)
            if world > 1:
                dist.all_reduce(self.grads(), op=dist.ReduceOp.SUM, group=group)
                scale = 1.0 / world
        with torch.cuda.device(self.dev):
            _lib.check(_lib.load().gn_avm_trainer_adam_step(self._t, scale, _stream(self.dev)), "gn_avm_trainer_adam_step")
        self.steps += 1

    def step(self, audios, frames, labels, masks=None):
        """One train sub-step.  Returns (loss float64[1] on the device, train-mode predictions (n,1))."""
        if self.use_graph and masks is None:
            vis = _to_dev(frames, torch.float32, self.dev)
            n = int(vis.shape[0])
            assert 1 <= n <= self.max_frames, "E: sub-batch larger than max_frames"
            self._gvis[:n].copy_(vis, non_blocking=True)
            if self.model.audio_included:
                self._gaud[:n].copy_(_to_dev(audios, torch.float32, self.dev), non_blocking=True)
            self._glab[:n].copy_(_to_dev(labels, torch.float32, self.dev).reshape(-1), non_blocking=True)

            def body():
                self.forward_backward(self._gaud[:n] if self.model.audio_included else None, self._gvis[:n], self._glab[:n])
                self.optimizer_step()
            self._graph.run(n, body)
            return self.loss, self.scores[:n].reshape(n, 1)
        n = self.forward_backward(audios, frames, labels, masks)
        self.optimizer_step()
        return self.loss, self.scores[:n].reshape(n, 1)

    def state_dict(self) -> dict:
        """Trained parameters + BatchNorm running buffers keyed like the reference's state_dict (CUDA tensors)."""
        st, keep = self._state_struct(self.model.state_dict(), clone=True)
        with torch.cuda.device(self.dev):
            _lib.check(_lib.load().gn_avm_trainer_export(self._t, C.byref(st), _stream(self.dev)), "gn_avm_trainer_export")
        return keep

    def sync_model(self):
        """Write the trained parameters and BatchNorm running buffers back into `model`.  Data parallel: gradients are
        all-reduced every step, so the parameters agree on every rank, but each rank's running statistics come from its
        own sub-batches -- they are averaged over the group here so that every rank ends with the same eval model."""
        import torch.distributed as dist
        trained = self.state_dict()
        if self.group is not False and dist.is_available() and dist.is_initialized():
            group = self.group if self.group is not None else dist.group.WORLD
            world = dist.get_world_size(group)
            if world > 1:
                for k in sorted(trained):
                    if k.endswith("running_mean") or k.endswith("running_var"):
                        dist.all_reduce(trained[k], op=dist.ReduceOp.SUM, group=group)
                        trained[k] /= world
        sd = self.model.state_dict()
        for k, v in trained.items():
            sd[k].copy_(v.to(sd[k].device))
        self.model.invalidate()

    PARAM_ORDER = ("visbl.conv1", "visbl.bnorm1", "visbl.conv2", "visbl.bnorm2", "visbl.conv3", "visbl.bnorm3",
                   "visbl.linear5", "audbl.conv1", "audbl.conv2", "audbl.linear3",
                   "fusion.0", "fusion.3", "fusion.6", "fusion.9", "fusion.12")

    def grad_dict(self) -> dict:
        """Gradients of the last forward_backward split per parameter (model.parameters() order; every tensor starts
        at a multiple of 4 elements in the flat vector)."""
        flat, sd, out, off = self.grads(), self.model.state_dict(), {}, 0
        for name in self.PARAM_ORDER:
            if name.startswith("audbl") and not self.model.audio_included:
                continue
            for suffix in (".weight", ".bias"):
                shp = sd[name + suffix].shape
                cnt = int(np.prod(shp))
                out[name + suffix] = flat[off:off + cnt].reshape(tuple(shp)).clone()
                off += (cnt + 3) // 4 * 4
        return out
