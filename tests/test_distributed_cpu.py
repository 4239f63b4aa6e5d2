"""world_size-2 gloo tests (CPU) of the multi-GPU host logic: shard bounds, the padded score all-gather for a
frame-sharded match, and the per-match row gather of BASELINE config 3."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from cvml_goalnet_b200 import distributed as D


def test_shard_bounds_cover_and_order():
    for n in (0, 1, 7, 240, 10800, 10801):
        for world in (1, 2, 3, 4, 8):
            b = D.shard_bounds(n, world)
            assert len(b) == world and b[0][0] == 0 and b[-1][1] == n
            for (a0, b0), (a1, b1) in zip(b, b[1:]):
                assert b0 == a1 and a0 <= b0
            per = -(-n // world) if n else 0
            assert all(hi - lo <= per for lo, hi in b)


def test_match_ownership_partitions():
    for n in (0, 1, 5, 64):
        for world in (1, 2, 8):
            seen = sorted(i for r in range(world) for i in D.matches_of_rank(n, r, world))
            assert seen == list(range(n))
            assert all(D.owner_of_match(i, world) == i % world for i in range(n))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n_total, n_matches, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # frame-sharded match: score of global frame i is f(i)
        full = (torch.arange(n_total, dtype=torch.float32) * 0.37 % 4.0 + 1.0).reshape(-1, 1)
        a, b = D.shard_bounds(n_total, world)[rank]
        got = D.gather_frame_scores(full[a:b].clone(), n_total)
        ok1 = torch.equal(got, full)
        # whole matches per rank: row of match m is g(m)
        width = 11
        rows_full = (torch.arange(n_matches * width, dtype=torch.float32).reshape(n_matches, width) * 1.5)
        mine = D.matches_of_rank(n_matches, rank, world)
        got2 = D.gather_match_rows(rows_full[mine].clone(), n_matches)
        ok2 = torch.equal(got2, rows_full)
        ind_full = (torch.arange(n_matches * width).reshape(n_matches, width) % 2).to(torch.uint8)
        got3 = D.gather_match_rows(ind_full[mine].clone(), n_matches)
        ok3 = torch.equal(got3, ind_full)
        np.save(os.path.join(out_dir, f"ok_{rank}.npy"), np.array([ok1, ok2, ok3]))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_total,n_matches", [(240, 5), (7, 2), (10801, 64), (1, 1)])
def test_gloo_world2_gathers(tmp_path, n_total, n_matches):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), n_total, n_matches, str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        assert np.load(tmp_path / f"ok_{r}.npy").all(), r
