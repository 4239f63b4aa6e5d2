"""gn_knapsack_static: the allocation-free variant of gn_knapsack (scratch in a static device array inside the library;
no counterpart in the reference).  Called through the C-ABI and compared,
bit for bit (integer work), with gn_knapsack on the same inputs, on two calls in a row that share the scratch.  The file
sorts last on purpose: a failure here must not hide the other tests."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GN_E_STATIC_LIMIT = -6
GN_KNAPSACK_STATIC_MAX_BYTES = 8 << 20


def _problems():
    rng = np.random.default_rng(2026)
    # (n_items, capacity, batch): config 2's shape first (multi-row cluster kernel), then a smaller batched problem that
    # finds the first call's bits, rows and status words in the shared scratch, then the global-row fallback
    for n_items, cap, batch in ((380, 24300, 1), (60, 5400, 20), (10, 230000, 1)):
        vals = rng.integers(0, 900, (batch, n_items)).astype(np.int32)
        wts = rng.integers(1, max(2, cap // 6), (batch, n_items)).astype(np.int32)
        yield n_items, cap, batch, vals, wts


def test_knapsack_static_matches_knapsack_on_consecutive_calls(cuda_device):
    import torch
    from cvml_goalnet_b200 import _lib
    from cvml_goalnet_b200.utils import _h, _stream

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    lib, h, stream = _lib.load(), _h(dev), _stream(dev)

    def ptr(t):
        return C.c_void_p(t.data_ptr())

    results = []
    for n_items, cap, batch, vals, wts in _problems():
        v, w = torch.from_numpy(vals).to(dev), torch.from_numpy(wts).to(dev)
        need = int(lib.gn_knapsack_workspace_bytes(n_items, cap)) * batch
        assert 0 < need <= GN_KNAPSACK_STATIC_MAX_BYTES
        ws = torch.empty(need, dtype=torch.uint8, device=dev)
        sel_ref = torch.zeros((batch, n_items), dtype=torch.uint8, device=dev)
        best_ref = torch.zeros(batch, dtype=torch.int32, device=dev)
        _lib.check(lib.gn_knapsack(h, ptr(v), ptr(w), n_items, cap, 0, batch, ptr(sel_ref), ptr(best_ref), ptr(ws), need,
                                   stream), "gn_knapsack")
        results.append((n_items, cap, batch, v, w, sel_ref, best_ref))
    torch.cuda.synchronize()

    # outputs are allocated up front: between here and the end of the loop nothing allocates
    outs = [(torch.full_like(s, 7), torch.full_like(b, -1)) for (_, _, _, _, _, s, b) in results]
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info(dev)
    reserved_before = torch.cuda.memory_reserved(dev)
    for rep in range(2):                                   # the whole sequence twice: every call after the first finds
        for (n_items, cap, batch, v, w, sel_ref, best_ref), (sel, best) in zip(results, outs):   # stale scratch
            sel.fill_(7)
            best.fill_(-1)
            _lib.check(lib.gn_knapsack_static(h, ptr(v), ptr(w), n_items, cap, 0, batch, ptr(sel), ptr(best), stream),
                       "gn_knapsack_static")
            torch.cuda.synchronize()
            assert torch.equal(best, best_ref), (rep, n_items, cap, batch)
            assert torch.equal(sel, sel_ref), (rep, n_items, cap, batch)
    free_after, _ = torch.cuda.mem_get_info(dev)
    assert torch.cuda.memory_reserved(dev) == reserved_before
    assert free_after >= free_before, "gn_knapsack_static must not allocate device memory"


def test_knapsack_static_rejects_problems_beyond_its_scratch(cuda_device):
    import torch
    from cvml_goalnet_b200 import _lib
    from cvml_goalnet_b200.utils import _h, _stream

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    lib, h, stream = _lib.load(), _h(dev), _stream(dev)
    n_items, cap = 4000, 100000
    assert int(lib.gn_knapsack_workspace_bytes(n_items, cap)) > GN_KNAPSACK_STATIC_MAX_BYTES
    v = torch.ones((1, n_items), dtype=torch.int32, device=dev)
    sel = torch.zeros((1, n_items), dtype=torch.uint8, device=dev)
    best = torch.zeros(1, dtype=torch.int32, device=dev)
    rc = lib.gn_knapsack_static(h, C.c_void_p(v.data_ptr()), C.c_void_p(v.data_ptr()), n_items, cap, 0, 1,
                                C.c_void_p(sel.data_ptr()), C.c_void_p(best.data_ptr()), stream)
    assert rc == GN_E_STATIC_LIMIT
    assert b"static scratch limit" in lib.gn_status_string(rc)
    # a batch that fits one by one but not together
    n_items, cap, batch = 380, 24300, 8
    assert int(lib.gn_knapsack_workspace_bytes(n_items, cap)) * batch > GN_KNAPSACK_STATIC_MAX_BYTES
    rc = lib.gn_knapsack_static(h, C.c_void_p(v.data_ptr()), C.c_void_p(v.data_ptr()), n_items, cap, 0, batch,
                                C.c_void_p(sel.data_ptr()), C.c_void_p(best.data_ptr()), stream)
    assert rc == GN_E_STATIC_LIMIT
    torch.cuda.synchronize()
