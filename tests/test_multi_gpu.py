"""2-GPU NCCL tests (skipped with fewer than two devices): a frame-sharded match and a batch of matches give the
same bits as the single-GPU path."""
import os
import socket

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    import torch.distributed as dist
    from cvml_goalnet_b200 import distributed as D, synthetic as syn, utils as U
    from cvml_goalnet_b200.main import MatchScorer
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        n, skip, sr, spf = 101, 15, 16000, 8000
        n_raw = n * skip
        frames = syn.make_frames(n, 48, 64, seed=3)
        y = syn.make_audio(n, spf, sr, seed=3)
        cps = syn.make_change_points(n_raw, seed=3, lo=10, hi=80)
        model = U.AVM(True)
        model.load_state_dict(syn.make_weights(seed=0))
        # single-GPU result on this rank
        sc = MatchScorer(model, n, 48, 64, y.size, sr, skip, n_raw, device=dev, h2d_chunks=2)
        s_ref, ind_ref, _ = sc.score_match(frames, y, cps, None)
        s_ref, ind_ref = s_ref.copy(), ind_ref.copy()
        # one match, frames sharded
        a, b = D.shard_bounds(n, world)[rank]
        scores, ind = D.score_match_frame_sharded(model, frames[a:b], y, sr, n, skip, n_raw, cps)
        ok1 = np.array_equal(scores.cpu().numpy(), s_ref) and np.array_equal(ind.cpu().numpy(), ind_ref)
        # three matches, whole matches per rank
        matches = []
        for mi in range(3):
            matches.append((syn.make_frames(n, 48, 64, seed=10 + mi), syn.make_audio(n, spf, sr, seed=10 + mi),
                            syn.make_change_points(n_raw, seed=10 + mi, lo=10, hi=80)))
        S, I = D.score_matches_sharded(sc, matches)
        ok2 = True
        for mi, (fr, yy, cp) in enumerate(matches):
            s1, i1, _ = sc.score_match(fr, yy, cp, None)
            ok2 = ok2 and np.array_equal(S[mi].cpu().numpy(), s1.reshape(-1)) and np.array_equal(I[mi].cpu().numpy(), i1)
        np.save(os.path.join(out_dir, f"ok_{rank}.npy"), np.array([ok1, ok2]))
    finally:
        dist.destroy_process_group()


def _batch_worker(rank, world, port, out_dir):
    """BASELINE config 3 in small: 5 matches (not a multiple of the rank count: padded ownership), frames sharded,
    both score-exchange variants, device and pinned-host entry points -- all against the single-GPU MatchScorer."""
    import torch.distributed as dist
    from cvml_goalnet_b200 import distributed as D, synthetic as syn, utils as U
    from cvml_goalnet_b200.main import MatchScorer, ShardedBatchScorer
    from oracle import goalnet_oracle as O
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        M, n, skip, sr, spf, Hh, Ww, users = 5, 100, 15, 16000, 8000, 48, 64, 20
        n_raw, samples = n * skip, n * spf
        a, b = D.shard_bounds(n, world)[rank]
        fr = [syn.make_frames(n, Hh, Ww, seed=20 + m) for m in range(M)]
        au = [syn.make_audio(n, spf, sr, seed=20 + m) for m in range(M)]
        cps = [syn.make_change_points(n_raw, seed=20 + m, lo=10, hi=80) for m in range(M)]
        gd = [np.stack([O.postprocess_array(x[:, None], skip, n_raw, cps[m])[1]
                        for x in syn.make_annotator_importances(n_raw, users, seed=20 + m)]) for m in range(M)]
        w = syn.make_weights(seed=0, score_gain=syn.SPREAD_GAIN)
        model = U.AVM(True)
        model.load_state_dict(w)
        single = MatchScorer(model, n, Hh, Ww, samples, sr, skip, n_raw, device=dev, h2d_chunks=2)
        ref = []
        for m in range(M):
            s1, i1, f1 = single.score_device(torch.from_numpy(fr[m]).to(dev), torch.from_numpy(au[m]).to(dev), cps[m],
                                             torch.from_numpy(gd[m]).to(dev))
            ref.append((s1.reshape(-1).clone(), i1.clone(), f1))
        frames_loc = torch.from_numpy(np.stack([f[a:b] for f in fr])).to(dev)
        audio_loc = torch.from_numpy(np.stack([x[a * spf:b * spf] for x in au])).to(dev)
        gd_dev = [torch.from_numpy(g).to(dev) for g in gd]
        oks = []
        for exchange in ("peer", "nccl"):
            bs = ShardedBatchScorer(model, M, n, Hh, Ww, samples, sr, skip, n_raw, max(len(c) for c in cps), users,
                                    device=dev, exchange=exchange, chunk_frames=130)      # ragged chunks across matches
            bs.set_change_points(cps)
            for entry in ("device", "host", "device"):
                if entry == "device":
                    S = bs.score_device(frames_loc, audio_loc, gd_dev)
                else:
                    S = bs.score_host(frames_loc.cpu().pin_memory(), audio_loc.cpu().pin_memory(), gd_dev)
                torch.cuda.synchronize(dev)
                ok = True
                for m in range(M):
                    ok = ok and bool(torch.equal(S[m], ref[m][0])) and bool(torch.equal(bs.indicator(m), ref[m][1]))
                    ov, gs, ss = bs.counts(m)
                    ok = ok and U.fscore_from_counts(ov, gs, ss) == ref[m][2]
                oks.append(ok)
            del bs
        np.save(os.path.join(out_dir, f"batch_{rank}.npy"), np.array(oks))
    finally:
        dist.destroy_process_group()


def test_two_gpu_config3_frame_sharded_batch(tmp_path):
    """ShardedBatchScorer (frames sharded, scores stored into the peer from the fusion epilogue or NCCL-gathered,
    selection distributed, indicators gathered) reproduces the single-GPU bits for every match on every rank."""
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices")
    import torch.multiprocessing as mp
    mp.spawn(_batch_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    for r in range(2):
        assert np.load(tmp_path / f"batch_{r}.npy").all(), (r, np.load(tmp_path / f"batch_{r}.npy"))


def test_two_gpu_sharded_paths_match_single_gpu(tmp_path):
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices")
    import torch.multiprocessing as mp
    mp.spawn(_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    for r in range(2):
        assert np.load(tmp_path / f"ok_{r}.npy").all(), r


def _train_worker(rank, world, port, out_dir):
    import torch.distributed as dist
    from cvml_goalnet_b200 import synthetic as syn, utils as U
    from cvml_goalnet_b200.train import FusionTrainer
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        n = 256                                     # rows per rank
        rng = np.random.default_rng(55)
        x = np.abs(rng.standard_normal((world * n, 640))).astype(np.float32)
        lab = rng.integers(1, 6, world * n).astype(np.float32)
        model = U.AVM(True)
        model.load_state_dict(syn.make_weights(seed=0))
        # data parallel: each rank its own shard, NCCL all-reduce of the flat gradient, identical Adam step
        dp = FusionTrainer(model, max_rows=n, fixed_loss=True, dropout_p=0.0, device=dev)
        for _ in range(3):
            dp.step(x[rank * n:(rank + 1) * n], lab[rank * n:(rank + 1) * n])
        p_dp = dp.flat_params()
        # every rank holds the same parameters afterwards
        gathered = [torch.empty_like(p_dp) for _ in range(world)]
        dist.all_gather(gathered, p_dp)
        same = all(torch.equal(gathered[0], g) for g in gathered)
        # with the element-wise loss and equal shards, the averaged gradient is the full-batch gradient
        full = FusionTrainer(model, max_rows=world * n, fixed_loss=True, dropout_p=0.0, device=dev, process_group=False)
        full.forward_backward(x, lab)
        g_full = full.grads.clone()
        dp2 = FusionTrainer(model, max_rows=n, fixed_loss=True, dropout_p=0.0, device=dev)
        dp2.forward_backward(x[rank * n:(rank + 1) * n], lab[rank * n:(rank + 1) * n])
        g = dp2.grads.clone()
        dist.all_reduce(g)
        g /= world
        rel = float((g - g_full).norm() / g_full.norm())
        np.save(os.path.join(out_dir, f"train_{rank}.npy"), np.array([float(same), rel]))
    finally:
        dist.destroy_process_group()


def test_two_gpu_data_parallel_fusion_training(tmp_path):
    """BASELINE config 4 at world size 2: flat-gradient all-reduce + identical Adam step on every rank."""
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs two CUDA devices")
    import torch.multiprocessing as mp
    mp.spawn(_train_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    for r in range(2):
        same, rel = np.load(tmp_path / f"train_{r}.npy")
        assert same == 1.0, r
        assert rel <= 2e-2, rel          # bf16 rounding of per-shard terms vs the full batch
