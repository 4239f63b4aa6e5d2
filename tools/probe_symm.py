"""Probe (2+ GPUs, run under torchrun): does torch's symmetric memory rendezvous work on this box, i.e. can a kernel on
rank r store straight into rank p's HBM through a mapped peer pointer (the mechanism for writing per-frame scores into
every peer from the fusion epilogue)?  Prints one JSON line per rank."""
import json
import os
import sys
import time

import torch
import torch.distributed as dist

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
out = {"rank": rank, "world": world}
try:
    import torch.distributed._symmetric_memory as symm
    t = symm.empty((1 << 20,), dtype=torch.float32, device=dev)
    t.fill_(float(rank))
    hdl = symm.rendezvous(t, dist.group.WORLD.group_name)
    out["buffer_ptrs"] = [hex(int(p)) for p in hdl.buffer_ptrs]
    out["signal_pad_ptrs"] = len(hdl.signal_pad_ptrs)
    hdl.barrier()
    peer = (rank + 1) % world
    pt = hdl.get_buffer(peer, (1 << 20,), torch.float32)
    pt[rank * 16:(rank + 1) * 16] = 100.0 + rank          # a store into the peer's memory from this rank's kernel
    hdl.barrier()
    torch.cuda.synchronize()
    src = (rank - 1) % world
    out["peer_store_seen"] = bool((t[src * 16:(src + 1) * 16] == 100.0 + src).all().item())
    # barrier latency
    for _ in range(5):
        hdl.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50):
        hdl.barrier()
    e1.record()
    torch.cuda.synchronize()
    out["symm_barrier_us"] = e0.elapsed_time(e1) * 1e3 / 50
    g = torch.empty(world * 86400, device=dev)
    s = torch.ones(86400, device=dev)
    for _ in range(5):
        dist.all_gather_into_tensor(g, s)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(50):
        dist.all_gather_into_tensor(g, s)
    e1.record()
    torch.cuda.synchronize()
    out["nccl_allgather_345KB_us"] = e0.elapsed_time(e1) * 1e3 / 50
    out["ok"] = True
except Exception as e:  # noqa: BLE001
    out["ok"] = False
    out["error"] = repr(e)[:500]
print(json.dumps(out), flush=True)
dist.destroy_process_group()
