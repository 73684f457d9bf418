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


def test_every_entry_point_rejects_bad_arguments_without_cuda(libpath):
    """0 on success, a negative code and a message otherwise, never an abort
    (software/ft601.h:43-50 convention): invalid sizes and null buffers are refused before
    any CUDA call, so this runs on a box without a GPU.  A non-null dummy stands in for
    device pointers that must not be dereferenced on these paths."""
    from mfcc_b200 import _lib
    L = _lib.lib()
    E = _lib.EINVAL
    buf = (ctypes.c_int64 * 64)()
    p = ctypes.cast(buf, ctypes.c_void_p)
    n64 = ctypes.c_int64()
    h = ctypes.c_void_p()
    calls = [
        ("run: null handle", lambda: L.mfcc_b200_run(None, p, 8, 8, 1, 0, p, 0, None, None)),
        ("run_host: null handle", lambda: L.mfcc_b200_run_host(None, p, 8, 8, 1, 0, p, 0)),
        ("stage: null handle", lambda: L.mfcc_b200_stage(None, 0, p, p, 8, 0, 0, None)),
        ("state_bytes: null handle", lambda: L.mfcc_b200_state_bytes(None, 4, ctypes.byref(n64))),
        ("step: null handle", lambda: L.mfcc_b200_step(None, p, p, 4, p, None)),
        ("step_host: null handle", lambda: L.mfcc_b200_step_host(None, 4, p, p, 0)),
        ("frame: windowlen > nfft", lambda: L.mfcc_b200_frame(p, p, 100, 1, 64, 8, 32, None)),
        ("frame: null", lambda: L.mfcc_b200_frame(None, None, 100, 1, 32, 8, 32, None)),
        ("window: nfft", lambda: L.mfcc_b200_window(p, p, 1, 12, 8, p, None)),
        ("window: precision", lambda: L.mfcc_b200_window(p, p, 1, 64, 15, p, None)),
        ("window: null curve", lambda: L.mfcc_b200_window(p, p, 1, 64, 8, None, None)),
        ("fft_stream: 16 points", lambda: L.mfcc_b200_fft_stream(p, p, 1, 16, p, None)),
        ("fft_stream: 8192 points", lambda: L.mfcc_b200_fft_stream(p, p, 1, 8192, p, None)),
        ("fft_stream: not a power of two", lambda: L.mfcc_b200_fft_stream(p, p, 1, 96, p, None)),
        ("fft_stream: null twiddles", lambda: L.mfcc_b200_fft_stream(p, p, 1, 64, None, None)),
        ("fft_stream: rows", lambda: L.mfcc_b200_fft_stream(p, p, -1, 64, p, None)),
        ("dct_stream: 4 inputs", lambda: L.mfcc_b200_dct_stream(p, p, 1, 4, p, None)),
        ("dct_stream: 2048 inputs", lambda: L.mfcc_b200_dct_stream(p, p, 1, 2048, p, None)),
        ("power: width", lambda: L.mfcc_b200_power(p, p, 8, 33, None)),
        ("power: null", lambda: L.mfcc_b200_power(None, p, 8, 24, None)),
        ("log2fix: widths", lambda: L.mfcc_b200_log2fix(p, p, 8, 16, 3, None)),
        ("log2fix: null", lambda: L.mfcc_b200_log2fix(None, p, 8, 16, 15, None)),
        ("filterbank: null steps", lambda: L.mfcc_b200_filterbank(p, p, 1, 256, 16, 30, 64, 30, 16, None, 0, None)),
        ("lifter: null", lambda: L.mfcc_b200_lifter(None, p, 4, 16, p, None)),
        ("standardize: null", lambda: L.mfcc_b200_standardize(None, 0, p, 4, 16, 1, None)),
        ("to_int16: null", lambda: L.mfcc_b200_to_int16(None, p, 4, None)),
        ("uart_frame: null", lambda: L.mfcc_b200_uart_frame(None, p, 4, 16, 0xa55a, None)),
        ("vad_step: null", lambda: L.mfcc_b200_vad_step(None, p, p, 4, 16, 93, 1, p, p, None)),
        ("librosa_create: null cfg", lambda: L.mfcc_b200_librosa_create(None, ctypes.byref(h))),
        ("librosa_run: null handle", lambda: L.mfcc_b200_librosa_run(None, p, 8, 8, 1, 1, p, p, 0, None, None)),
        ("librosa_workspace: null handle", lambda: L.mfcc_b200_librosa_workspace(None, 1, 1, ctypes.byref(n64))),
    ]
    for what, call in calls:
        rc = call()
        assert rc in (E, _lib.EUNSUPPORTED), (what, rc)
        assert len(L.mfcc_b200_last_error()) > 0, what
    # empty work is success, not an error
    assert L.mfcc_b200_power(None, None, 0, 24, None) == 0
    assert L.mfcc_b200_fft_stream(None, None, 0, 64, None, None) == 0
    assert L.mfcc_b200_window(None, None, 0, 64, 8, None, None) == 0
    L.mfcc_b200_destroy(None)                                   # a no-op, like free(NULL)
    L.mfcc_b200_librosa_destroy(None)
    assert L.mfcc_b200_launch_count() == 0                      # nothing above reached a launch


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
