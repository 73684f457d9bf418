"""CPU checks of the C-ABI boundary: the library builds for sm_100a, loads, and
exports every symbol include/mfcc_b200.h declares.  No compute calls (no GPU here)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def libpath():
    from mfcc_b200 import build
    return build.build()


def test_header_symbols_exported(libpath):
    hdr = open(os.path.join(ROOT, "include", "mfcc_b200.h")).read()
    declared = set(re.findall(r"\b(mfcc_b200_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 10
    lib = ctypes.CDLL(libpath)
    for name in declared:
        assert hasattr(lib, name), name
    from mfcc_b200 import _lib
    assert declared == set(_lib.EXPORTS)


def test_cfg_struct_matches_header(libpath):
    from mfcc_b200 import _lib
    # 16 int32, 8 pointers, 2 floats
    assert ctypes.sizeof(_lib.Cfg) == 16 * 4 + 8 * 8 + 2 * 4
    assert ctypes.sizeof(_lib.Taps) == 6 * 8


def test_version_and_error_string(libpath):
    from mfcc_b200 import _lib
    L = _lib.lib()
    assert L.mfcc_b200_version() == 1
    assert L.mfcc_b200_launch_count() == 0
    # argument validation happens before any CUDA call
    h = ctypes.c_void_p()
    assert L.mfcc_b200_create(None, ctypes.byref(h)) == _lib.EINVAL
    assert b"null" in L.mfcc_b200_last_error()
    c = _lib.Cfg()
    assert L.mfcc_b200_create(ctypes.byref(c), ctypes.byref(h)) == _lib.EINVAL   # struct_size
    assert L.mfcc_b200_run(None, None, 0, 0, 0, 0, None, 0, None, None) == _lib.EINVAL


def test_sass_uses_tma(libpath):
    """The fused kernel's sample tiles arrive by TMA bulk copy: UBLKCP in the SASS."""
    import shutil
    import subprocess
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    sass = subprocess.run([cuobjdump, "-sass", libpath], capture_output=True, text=True).stdout
    assert "UBLKCP" in sass and "sm_100a" in sass


def test_no_cpu_fallback_in_product():
    """The product never imports the oracle."""
    pkg = os.path.join(ROOT, "mfcc_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dp, f)).read()
                assert "oracle" not in src.replace("no CPU fallback", ""), f


def test_stage_constructors_validate():
    import torch
    from mfcc_b200 import stages as S
    with pytest.raises(ValueError):
        S.Frame(windowlen=600, nfft=512)                 # frame.py:51
    with pytest.raises(ValueError):
        S.WindowHamming(nfft=4)                          # window.py:13
    with pytest.raises(ValueError):
        S.FftStream(nfft=500)                            # fft.py:351-353
    d = S.Discard(first=0, count=16)
    assert d(torch.arange(64).reshape(2, 32)).shape == (2, 16)
    from mfcc_b200 import MFCCConfig
    with pytest.raises(ValueError):
        MFCCConfig(nfft=500)
    with pytest.raises(ValueError):
        MFCCConfig(nfilters=32, nceptrums=33)
