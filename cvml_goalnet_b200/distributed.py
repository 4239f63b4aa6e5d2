"""Multi-GPU plumbing for the scoring path: one process per GPU, `torch.distributed` (NCCL on the GPUs,
gloo in the CPU tests) for the one exchange step the path has.

The reference has no distributed code.  In eval mode every frame is scored
independently (reference utils.py:260-272), so the path shards with no data-path collective until the
selection step:

* one match on P ranks  -> contiguous frame ranges per rank, then an all-gather of the fp32 scores
  (43 KB for a 90-minute match) so that the knapsack sees the whole score vector;
* many matches on P ranks (BASELINE config 3) -> whole matches per rank (round-robin), scores and
  indicators all-gathered at the end so that every rank holds every summary.

Everything here is device-agnostic tensor plumbing (it runs under gloo on CPU tensors in the tests);
the CUDA work is done by `MatchScorer` / `utils` on each rank's own device.
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

import torch
import torch.distributed as dist


def shard_bounds(n: int, world: int) -> List[Tuple[int, int]]:
    """Contiguous frame ranges [a, b) per rank, ceil(n / world) frames each (the last ranks may be short or
    empty).  Rank r gets frames [r*ceil(n/P), ...) as in SURVEY.md section 8(e)."""
    assert n >= 0 and world >= 1, "E: invalid shard request"
    per = -(-n // world) if n else 0
    return [(min(r * per, n), min((r + 1) * per, n)) for r in range(world)]


def matches_of_rank(n_matches: int, rank: int, world: int) -> List[int]:
    """Round-robin ownership of whole matches (selection is sequential per match, so matches are the unit)."""
    return list(range(rank, n_matches, world))


def owner_of_match(i: int, world: int) -> int:
    return i % world


def _world(group=None) -> Tuple[int, int]:
    if not dist.is_available() or not dist.is_initialized():
        return 0, 1
    return dist.get_rank(group), dist.get_world_size(group)


def gather_frame_scores(local_scores: torch.Tensor, n_total: int, group=None) -> torch.Tensor:
    """All-gather the per-frame scores of one match whose frames were sharded with `shard_bounds`.

    local_scores: (k, 1) or (k,) float32 scores of this rank's frame range; returns (n_total, 1) on the same
    device on every rank.  Ranges are padded to ceil(n/P) rows so that one `all_gather_into_tensor`
    (NCCL: a single ring/NVLS collective) moves everything."""
    rank, world = _world(group)
    ls = local_scores.reshape(-1)
    bounds = shard_bounds(n_total, world)
    a, b = bounds[rank]
    assert ls.numel() == b - a, "E: Inconsistency in data loader definition"
    if world == 1:
        return ls.reshape(-1, 1)
    per = -(-n_total // world) if n_total else 0
    send = torch.zeros(per, dtype=ls.dtype, device=ls.device)
    send[: b - a] = ls
    recv = torch.empty(world * per, dtype=ls.dtype, device=ls.device)
    dist.all_gather_into_tensor(recv, send, group=group)
    # with ceil(n/P)-sized contiguous ranges the padding is all at the tail: rank r's rows sit at [r*per, ...)
    return recv[:n_total].reshape(-1, 1)


def gather_match_rows(local_rows: torch.Tensor, n_matches: int, group=None) -> torch.Tensor:
    """All-gather per-match rows (scores (m, N) or indicators (m, N_raw)) computed under `matches_of_rank`
    ownership; returns (n_matches, width) in match order on every rank."""
    rank, world = _world(group)
    mine = matches_of_rank(n_matches, rank, world)
    assert local_rows.shape[0] == len(mine), "E: Inconsistency in data loader definition"
    if world == 1:
        return local_rows
    per = -(-n_matches // world)
    width = local_rows.shape[1]
    send = torch.zeros((per, width), dtype=local_rows.dtype, device=local_rows.device)
    send[: len(mine)] = local_rows
    recv = torch.empty((world * per, width), dtype=local_rows.dtype, device=local_rows.device)
    dist.all_gather_into_tensor(recv, send, group=group)
    out = torch.empty((n_matches, width), dtype=local_rows.dtype, device=local_rows.device)
    for r in range(world):
        idx = matches_of_rank(n_matches, r, world)
        if idx:
            out[idx] = recv[r * per: r * per + len(idx)]
    return out


def score_match_frame_sharded(model, frames_local, audio, sr: int, n_total: int, skip_frames: int, full_n_frames: int,
                              change_points, bin_length: int = 30, group=None, select_on_all_ranks: bool = True):
    """One match on P GPUs: this rank preprocesses and scores its contiguous frame range, the scores are
    all-gathered, and the (sequential) knapsack selection runs on every rank (or on rank 0 only).

    frames_local : uint8 (k, H, W, 3) -- the frames of this rank's `shard_bounds` range
    audio        : float32 (samples,) -- the whole track (346 MB for 90 minutes; each rank reads only the
                   slots of its range)
    Returns (scores (n_total, 1) cuda, indicator (N_raw,) uint8 cuda | None)."""
    from . import utils as U
    rank, world = _world(group)
    a, b = shard_bounds(n_total, world)[rank]
    vis = U.extract_condensed_frame_tensor_array(frames_local, model.S)
    assert vis.shape[0] == b - a, "E: Inconsistency in data loader definition"
    aud = None
    if model.audio_included:
        aud_full = torch.empty((n_total, 30, bin_length), dtype=torch.float32, device=vis.device)
        U.extract_audio_features_array(audio, sr, n_total, bin_length, device=vis.device, out=aud_full, slots=(a, b))
        aud = aud_full[a:b]
    local = model(aud, vis)
    scores = gather_frame_scores(local, n_total, group)
    ind = None
    if select_on_all_ranks or rank == 0:
        _, ind = U.postprocess_array(scores, skip_frames, full_n_frames, change_points, device=vis.device,
                                     return_device=True)
    return scores, ind


def score_matches_sharded(scorer, matches: Sequence, group=None):
    """BASELINE config 3: a batch of matches, whole matches per rank, summaries all-gathered.

    scorer  : a `MatchScorer` for the common match shape on this rank's device
    matches : sequence of (frames, audio, change_points) for ALL matches (a rank only touches its own);
              entries of other ranks may be None
    Returns (scores (n_matches, N) float32, indicators (n_matches, N_raw) uint8) on this rank's device."""
    rank, world = _world(group)
    mine = matches_of_rank(len(matches), rank, world)
    dev = scorer.dev
    s_rows = torch.empty((len(mine), scorer.n), dtype=torch.float32, device=dev)
    i_rows = torch.empty((len(mine), scorer.n_raw), dtype=torch.uint8, device=dev)
    for k, mi in enumerate(mine):
        frames, audio, cps = matches[mi]
        on_device = torch.is_tensor(frames) and frames.is_cuda
        if on_device:
            s, ind, _ = scorer.score_device(frames, audio, cps, None)
            s_rows[k] = s.reshape(-1)
            i_rows[k] = ind
        else:
            s, ind, _ = scorer.score_match(frames, audio, cps, None)
            s_rows[k] = torch.from_numpy(s).reshape(-1).to(dev)
            i_rows[k] = torch.from_numpy(ind).to(dev)
    return gather_match_rows(s_rows, len(matches), group), gather_match_rows(i_rows, len(matches), group)
