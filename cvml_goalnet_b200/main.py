"""Inference entry points mirroring the reference's `main.py` (infer, main.py:300-348) plus the
array-based scorer the benchmark and the multi-GPU code drive.

`MatchScorer.score_match` is the public end-to-end call: frames (uint8 BGR) + mono audio of one video
-> per-frame scores -> keyshot indicator (-> F-scores), i.e. main.py:315-345 without the file I/O.
Host inputs are staged through pinned memory in chunks on a copy stream so that the PCIe transfer of
chunk i+1 overlaps the preprocessing kernel of chunk i.
"""
from __future__ import annotations

import os

import numpy as np
import torch

from . import utils as U


class MatchScorer:
    """Preallocated, stream-ordered pipeline for videos of a fixed shape.

    n_frames kept frames of HxW source pixels, `samples` audio samples at `sr` Hz, raw-rate length
    `full_n_frames`, `skip_frames` raw frames per kept frame (utils.py:283).
    """

    def __init__(self, model: U.AVM, n_frames: int, H: int, W: int, samples: int, sr: int, skip_frames: int,
                 full_n_frames: int, bin_length: int = 30, device=None, h2d_chunks: int = 16):
        self.dev = U._device(device)
        self.model = model
        self.n, self.H, self.W = int(n_frames), int(H), int(W)
        self.samples, self.sr = int(samples), int(sr)
        self.skip, self.n_raw, self.B = int(skip_frames), int(full_n_frames), int(bin_length)
        d = self.dev
        self.frames_dev = torch.empty((self.n, H, W, 3), dtype=torch.uint8, device=d)
        self.audio_dev = torch.empty(self.samples, dtype=torch.float32, device=d)
        self.vis = torch.empty((self.n, 3, model.S, model.S), dtype=torch.float32, device=d)
        self.aud = torch.empty((self.n, 30, self.B), dtype=torch.float32, device=d)
        self.scores = torch.empty((self.n, 1), dtype=torch.float32, device=d)
        self.copy_stream = torch.cuda.Stream(device=d)
        self.h2d_chunks = max(1, min(int(h2d_chunks), self.n))
        self.scores_host = torch.empty((self.n, 1), dtype=torch.float32).pin_memory()
        self.ind_host = torch.empty(self.n_raw, dtype=torch.uint8).pin_memory()
        self.h2d_bytes = 0
        self.d2h_bytes = 0
        # audio samples each H2D chunk must bring along (slot boundaries of utils.py:322-330); consecutive
        # chunks may share a boundary sample, so each range starts where the previous one ended at the latest
        bounds = np.linspace(0, self.n, self.h2d_chunks + 1).astype(np.int64)
        self._audio_ranges, prev_hi = [], 0
        for i in range(self.h2d_chunks):
            lo, hi = U.audio_slot_sample_range(self.samples, self.n, int(bounds[i]), int(bounds[i + 1]))
            lo = min(lo, prev_hi)
            self._audio_ranges.append((lo, max(hi, lo)))
            prev_hi = max(hi, prev_hi)

    # ---- device-resident path (inputs already in HBM) ------------------------------------------------
    def score_device(self, frames_dev, audio_dev, change_points_dev_or_host, gd_dev=None):
        """frames_dev uint8 (N,H,W,3) and audio_dev float32 (samples,) on the device.  Returns
        (scores (N,1) cuda, indicator (N_raw,) uint8 cuda, (f_avg, f_max) | None)."""
        U.extract_condensed_frame_tensor_array(frames_dev, self.model.S, out=self.vis)
        if self.model.audio_included:
            U.extract_audio_features_array(audio_dev, self.sr, self.n, self.B, out=self.aud)
        self.model(self.aud, self.vis, out=self.scores)
        _, ind = U.postprocess_array(self.scores, self.skip, self.n_raw, change_points_dev_or_host, device=self.dev,
                                     return_device=True)
        fs = U.get_fscore(gd_dev, ind, device=self.dev) if gd_dev is not None else None
        return self.scores, ind, fs

    # ---- host path: what a user of the reference calls with decoded arrays ----------------------------
    def score_match(self, frames_host, audio_host, change_points, gd=None):
        """frames_host: uint8 (N,H,W,3) (NumPy or pinned torch CPU tensor); audio_host: float32 (samples,).
        Copies in (chunked, overlapped with preprocessing), scores, selects, copies scores + indicator out.
        Returns (scores np (N,1), indicator np (N_raw,), (f_avg, f_max) | None)."""
        d = self.dev
        fh = frames_host if torch.is_tensor(frames_host) else torch.from_numpy(np.ascontiguousarray(frames_host))
        ah = audio_host if torch.is_tensor(audio_host) else torch.from_numpy(np.ascontiguousarray(audio_host))
        assert tuple(fh.shape) == (self.n, self.H, self.W, 3), "E: Inconsistency in data loader definition"
        assert ah.numel() == self.samples, "E: Inconsistency in data loader definition"
        main = torch.cuda.current_stream(d)
        cs = self.copy_stream
        cs.wait_stream(main)                       # previous step's readers of the staging buffers are done
        ah = ah.reshape(-1)
        bounds = np.linspace(0, self.n, self.h2d_chunks + 1).astype(np.int64)
        events = []
        with torch.cuda.stream(cs):
            for i in range(self.h2d_chunks):
                a, b = int(bounds[i]), int(bounds[i + 1])
                if self.model.audio_included:
                    lo, hi = self._audio_ranges[i]
                    self.audio_dev[lo:hi].copy_(ah[lo:hi], non_blocking=True)
                self.frames_dev[a:b].copy_(fh[a:b], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(cs)
                events.append((a, b, ev))
        # chunk i is preprocessed and scored while chunk i+1 is still crossing PCIe
        for a, b, ev in events:
            main.wait_event(ev)
            if self.model.audio_included:
                U.extract_audio_features_array(self.audio_dev, self.sr, self.n, self.B, out=self.aud, slots=(a, b))
            U.extract_condensed_frame_tensor_array(self.frames_dev[a:b], self.model.S, out=self.vis[a:b])
            self.model(self.aud[a:b], self.vis[a:b], out=self.scores[a:b])
        _, ind = U.postprocess_array(self.scores, self.skip, self.n_raw, change_points, device=d, return_device=True)
        self.scores_host.copy_(self.scores, non_blocking=True)
        self.ind_host.copy_(ind, non_blocking=True)
        fs = U.get_fscore(gd, ind, device=d) if gd is not None else None     # synchronises (reads the counts)
        main.synchronize()
        self.h2d_bytes = fh.numel() + ah.numel() * 4
        self.d2h_bytes = self.scores_host.numel() * 4 + self.ind_host.numel()
        return self.scores_host.numpy(), self.ind_host.numpy(), fs


def infer_arrays(frames_u8, y, sr, skip_frames, full_n_frames, change_points, model: U.AVM, full_frames=None,
                 bin_length: int = 30):
    """main.py:315-345 on decoded arrays: returns (summarized_video | None, indicator, scores)."""
    vis = U.extract_condensed_frame_tensor_array(frames_u8, model.S)
    aud = U.extract_audio_features_array(y, sr, len(frames_u8), bin_length) if model.audio_included else None
    scores = model(aud, vis)
    video, ind = U.postprocess_array(scores, skip_frames, full_n_frames, change_points, full_frames=full_frames)
    return video, ind, scores.cpu().numpy()


def infer(video_fp: str, audio_included: bool, weights_fp: str | None = None,
          h5_file_path: str = "ydata-tvsum50-v1_1/ground_truth/eccv16_dataset_tvsum_google_pool5.h5",
          mat_file_path: str = "ydata-tvsum50-v1_1/ground_truth/ydata-tvsum50.mat"):
    """main.py:300-348 with the reference's paths and skip_frames = 60.  Needs the TVSum files, a trained
    checkpoint, h5py and librosa -- none of which ship with the reference or this image -- so it is kept
    for interface parity; `infer_arrays` is the tested path."""
    if weights_fp is None:
        weights_fp = "./models/opt_frame_importance_model.pt" if audio_included else \
            "./models/opt_frame_importance_model_no_audio.pt"
    skip_frames = 60                                                           # main.py:311
    video_frames, full_n = U.extract_condensed_frame_tensor(video_fp, skip_frames=skip_frames)
    full_frames = U.get_frame_tensor(video_fp)
    audio_fp = ".".join(video_fp.split(".")[:-1]) + ".wav"
    model = U.AVM(audio_included)
    model.load_state_dict(torch.load(weights_fp, map_location="cpu"))
    aud = U.extract_audio_features(audio_fp, len(video_frames), 30) if audio_included else None   # quirk Q2
    scores = model(aud, video_frames)
    video_id = os.path.basename(video_fp).split(".")[0]
    return U.postprocess(video_id, h5_file_path, mat_file_path, scores, skip_frames, full_n, full_frames)
