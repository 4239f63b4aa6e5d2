"""ShardedBatchScorer on ONE GPU (world size 1: no exchange): the config-3 pipeline -- flat (match, frame) chunks that
straddle match boundaries, scattered scores, padded batched knapsacks, per-match indicators and F-score counts -- against
the single-match MatchScorer, bit for bit.  The 2-GPU version lives in tests/test_multi_gpu.py."""
import numpy as np
import pytest
import torch

from cvml_goalnet_b200 import synthetic as syn
from oracle import goalnet_oracle as O

pytestmark = pytest.mark.gpu


def test_batch_scorer_single_rank_equals_match_scorer(cuda_device):
    from cvml_goalnet_b200 import utils as U
    from cvml_goalnet_b200.main import MatchScorer, ShardedBatchScorer
    dev = cuda_device
    M, n, skip, sr, spf, H, W, users = 4, 90, 15, 16000, 8013, 40, 56, 20
    n_raw = n * skip
    fr = [syn.make_frames(n, H, W, seed=40 + m) for m in range(M)]
    au = [syn.make_audio(n, spf, sr, seed=40 + m)[: n * spf - 5] for m in range(M)]     # ragged slot bounds
    samples = au[0].size
    cps = [syn.make_change_points(n_raw, seed=40 + m, lo=10, hi=30 + 40 * m) for m in range(M)]   # very different clip counts
    gd = [np.stack([O.postprocess_array(x[:, None], skip, n_raw, cps[m])[1]
                    for x in syn.make_annotator_importances(n_raw, users, seed=40 + m)]) for m in range(M)]
    model = U.AVM(True)
    model.load_state_dict(syn.make_weights(seed=0, score_gain=syn.SPREAD_GAIN))
    single = MatchScorer(model, n, H, W, samples, sr, skip, n_raw, device=dev, h2d_chunks=3)
    bs = ShardedBatchScorer(model, M, n, H, W, samples, sr, skip, n_raw, max(len(c) for c in cps), users, device=dev,
                            chunk_frames=77)
    bs.set_change_points(cps)
    frames = torch.from_numpy(np.stack(fr)).to(dev)
    audio = torch.from_numpy(np.stack(au)).to(dev)
    gd_dev = [torch.from_numpy(g).to(dev) for g in gd]
    for entry in ("device", "host"):
        if entry == "device":
            S = bs.score_device(frames, audio, gd_dev)
        else:
            S = bs.score_host(frames.cpu().pin_memory(), audio.cpu().pin_memory(), gd_dev)
        torch.cuda.synchronize(dev)
        for m in range(M):
            s1, i1, f1 = single.score_device(frames[m], audio[m], cps[m], gd_dev[m])
            assert torch.equal(S[m], s1.reshape(-1)), (entry, m)
            assert torch.equal(bs.indicator(m), i1), (entry, m)
            assert U.fscore_from_counts(*bs.counts(m)) == f1, (entry, m)
            # and the oracle's integer path on those scores
            _, o_ind = O.postprocess_array(s1.cpu().numpy(), skip, n_raw, cps[m])
            np.testing.assert_array_equal(i1.cpu().numpy(), o_ind)
