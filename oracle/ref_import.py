"""Import the UNMODIFIED reference `utils.py`: from /root/reference in the build container, or from
`oracle/_ref/utils.py` -- the copy `__graft_entry__.build()` drops there (git-ignored, never part of the
history; it travels to the GPU box with the working tree the way the built .so does) so that
`bench.py --impl reference` / `cpu_baseline` can time the reference's OWN `AVM`, `knapsack`,
`get_clip_information`, `expand_array` and `get_fscore` on the box's host cores.

TEST INFRASTRUCTURE ONLY.  Used by `oracle/gen_golden.py`, by the `not gpu` tests that pin the oracle
restatement against the reference's own code, and by the CPU-baseline legs of `bench.py`.  Nothing in
the product package or the `-m gpu` tests may import this module.

The reference imports four third-party modules that are absent from this image
(`moviepy.editor`, `h5py`, `matplotlib.pyplot`, `librosa`; reference
utils.py:7,9,10,13).  They are only touched by file-I/O functions that are out
of scope, so they are stubbed with MagicMock and the
pure-arithmetic functions (`AVM`, `knapsack`, `get_clip_information`,
`expand_array`, `get_fscore`, the body of `extract_condensed_frame_tensor`) run
unmodified.
"""
import os
import sys
from unittest.mock import MagicMock

REFERENCE_DIR = "/root/reference"
LOCAL_COPY_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


def reference_dir():
    """Directory holding the reference's utils.py: the reference tree itself, else the travelled copy, else None."""
    for d in (REFERENCE_DIR, LOCAL_COPY_DIR):
        if os.path.isfile(os.path.join(d, "utils.py")):
            return d
    return None


def reference_available() -> bool:
    return reference_dir() is not None


def stage_reference_copy() -> str:
    """build(): copy the reference's utils.py byte for byte to oracle/_ref/ (only where /root/reference exists).
    Returns the destination path, or '' when there is nothing to copy."""
    import shutil
    src = os.path.join(REFERENCE_DIR, "utils.py")
    if not os.path.isfile(src):
        return ""
    os.makedirs(LOCAL_COPY_DIR, exist_ok=True)
    dst = os.path.join(LOCAL_COPY_DIR, "utils.py")
    shutil.copyfile(src, dst)
    return dst


def import_reference_utils():
    """Return the reference `utils` module (cached), or raise FileNotFoundError."""
    if not reference_available():
        raise FileNotFoundError("reference tree not present (expected only in the build container)")
    if "goalnet_reference_utils" in sys.modules:
        return sys.modules["goalnet_reference_utils"]
    os.environ.setdefault("PYTHONBREAKPOINT", "0")  # quirk Q1: stray breakpoint() at utils.py:321
    for name in ("moviepy", "moviepy.editor", "h5py", "matplotlib", "matplotlib.pyplot",
                 "matplotlib.ticker", "librosa"):
        sys.modules.setdefault(name, MagicMock())
    import importlib.util
    spec = importlib.util.spec_from_file_location("goalnet_reference_utils",
                                                  os.path.join(reference_dir(), "utils.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules["goalnet_reference_utils"] = mod
    spec.loader.exec_module(mod)
    return mod
