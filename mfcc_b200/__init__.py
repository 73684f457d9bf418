"""B200-native MFCC extractor behind the stage API of lambdaconcept/mfcc.

``MFCC`` and the stage operators need the in-tree CUDA library
(mfcc_b200/libmfcc_b200.so, built by ``__graft_entry__.build()``) and a GPU;
there is no CPU fallback.  ``MFCCConfig`` and ``tables`` are plain Python.
"""
from .config import MFCCConfig, MODE_INT, MODE_FP32  # noqa: F401
from . import tables  # noqa: F401


def __getattr__(name):
    # torch-dependent parts are imported on first use so that table/config code stays
    # importable in tools that have no torch
    if name == "MFCC":
        from .mfcc import MFCC
        return MFCC
    if name == "LiveSession":
        from .live import LiveSession
        return LiveSession
    if name == "LibrosaMFCC":
        from .librosa_mode import LibrosaMFCC
        return LibrosaMFCC
    if name in ("Preemph", "Frame", "WindowHamming", "FftStream", "PowerSpectrum", "FilterBank",
                "Log2Fix", "DCTStream", "Discard"):
        from . import stages
        return getattr(stages, name)
    raise AttributeError(name)
