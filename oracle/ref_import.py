"""Import the UNMODIFIED reference `utils.py` from /root/reference (this container only).

TEST INFRASTRUCTURE ONLY.  Used by `oracle/gen_golden.py` and by the `not gpu`
tests that pin the oracle restatement against the reference's own code.  Nothing
in the product package, `bench.py` or the `-m gpu` tests may import this module:
`/root/reference` does not exist on the GPU box.

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


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_DIR, "utils.py"))


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
                                                  os.path.join(REFERENCE_DIR, "utils.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules["goalnet_reference_utils"] = mod
    spec.loader.exec_module(mod)
    return mod
