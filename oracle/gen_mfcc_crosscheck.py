"""TEST INFRASTRUCTURE (not product code).  Independent cross-check vectors for the one stage the reference
cannot pin here: `librosa.feature.mfcc(y=…, sr=sr, n_mfcc=30)` (utils.py:333; librosa is absent from the image).

`torchaudio.transforms.MFCC` is a separate third-party implementation of the same published definition
(torchaudio's own test-suite checks it against librosa).  Configured with librosa 0.10's defaults
(n_fft 2048, hop 512, periodic Hann, centred zero padding, power 2, 128 Slaney mels with Slaney norm,
AmplitudeToDB power/top_db 80, DCT-II ortho) it gives vectors the oracle's restatement did not produce itself.
Run here once (`python oracle/gen_mfcc_crosscheck.py`); the output tests/golden/mfcc_torchaudio.npz is committed.
"""
import os
import sys

import numpy as np
import torch
import torchaudio

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cvml_goalnet_b200 import synthetic as syn  # noqa: E402


def torchaudio_mfcc(y: np.ndarray, sr: int, n_mfcc: int = 30) -> np.ndarray:
    t = torchaudio.transforms.MFCC(
        sample_rate=sr, n_mfcc=n_mfcc, dct_type=2, norm="ortho", log_mels=False,
        melkwargs=dict(n_fft=2048, hop_length=512, win_length=2048, n_mels=128, f_min=0.0, f_max=sr / 2,
                       center=True, pad_mode="constant", power=2.0, norm="slaney", mel_scale="slaney"))
    return t(torch.from_numpy(np.ascontiguousarray(y, dtype=np.float32))).numpy()


CASES = {"sr16k_8000": (16000, 8000), "sr16k_4000": (16000, 4000), "sr22k_11025": (22050, 11025),
         "sr44k_20000": (44100, 20000), "sr16k_short_700": (16000, 700)}


def main():
    out = {"torchaudio_version": np.array(torchaudio.__version__)}
    for i, (name, (sr, n)) in enumerate(CASES.items()):
        y = syn.make_audio(1, n, sr, seed=100 + i)
        out[f"{name}_y"] = y
        out[f"{name}_sr"] = np.array(sr)
        out[f"{name}_mfcc"] = torchaudio_mfcc(y, sr)
    # a near-silent segment exercises the amin / top_db clamps
    y = (1e-6 * syn.make_audio(1, 8000, 16000, seed=7)).astype(np.float32)
    out["quiet_y"], out["quiet_sr"], out["quiet_mfcc"] = y, np.array(16000), torchaudio_mfcc(y, 16000)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "mfcc_torchaudio.npz"), **out)
    print("wrote tests/golden/mfcc_torchaudio.npz:", [k for k in out if k.endswith("_mfcc")])


if __name__ == "__main__":
    main()
