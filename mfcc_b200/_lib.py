"""ctypes binding of libmfcc_b200.so (include/mfcc_b200.h).

There is no CPU fallback: if the library is missing or a call fails, this module
raises.  The library is built in-tree by ``__graft_entry__.build()`` /
``mfcc_b200.build.build()``.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmfcc_b200.so")

OK, EINVAL, EUNSUPPORTED, ECUDA = 0, -1, -2, -3

(STAGE_PREEMPH, STAGE_FRAME, STAGE_WINDOW, STAGE_FFT, STAGE_POWER, STAGE_MEL, STAGE_LOG2,
 STAGE_DCT) = range(8)


class Cfg(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "struct_size", "mode", "nfft", "hop", "nfilters", "nceptrums", "window_shift",
        "pow_shift", "mel_wbits", "mel_lo_bit", "mel_out_bits", "log_precision",
        "log_out_bits")] + [("reserved", C.c_int32 * 3)] + [
        (n, C.c_void_p) for n in ("window_curve", "tw_fft", "tw_dct", "mel_weight", "mel_close",
                                  "f_window", "f_mel", "f_dct")] + [
        ("f_preemph", C.c_float), ("f_mel_floor", C.c_float)]


class Taps(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("d_window", "d_fft", "d_power", "d_mel", "d_log2",
                                          "d_dct")]


class LibrosaCfg(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("struct_size", "n_fft", "hop", "n_mels", "n_mfcc", "center")] + [
        ("amin", C.c_float), ("top_db", C.c_float)] + [(n, C.c_void_p) for n in ("window", "mel", "dct")]


class LibrosaTaps(C.Structure):
    _fields_ = [("d_power", C.c_void_p), ("d_logmel", C.c_void_p)]


class MfccError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("libmfcc_b200: %s (code %d)" % (msg, code))
        self.code = code


_lib = None


def lib():
    """Load the shared library; raise loudly if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "%s is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback)" % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    vp, i32, i64 = C.c_void_p, C.c_int32, C.c_int64
    L.mfcc_b200_create.restype = C.c_int
    L.mfcc_b200_create.argtypes = [C.POINTER(Cfg), C.POINTER(vp)]
    L.mfcc_b200_destroy.restype = None
    L.mfcc_b200_destroy.argtypes = [vp]
    L.mfcc_b200_run.restype = C.c_int
    L.mfcc_b200_run.argtypes = [vp, vp, i64, i64, i32, i64, vp, i64, C.POINTER(Taps), vp]
    L.mfcc_b200_run_host.restype = C.c_int
    L.mfcc_b200_run_host.argtypes = [vp, vp, i64, i64, i32, i64, vp, i64]
    L.mfcc_b200_host_config.restype = C.c_int
    L.mfcc_b200_host_config.argtypes = [vp, i64, i32]
    L.mfcc_b200_debug_fail_chunk.restype = C.c_int
    L.mfcc_b200_debug_fail_chunk.argtypes = [vp, i32]
    L.mfcc_b200_host_alloc.restype = C.c_int
    L.mfcc_b200_host_alloc.argtypes = [C.POINTER(vp), i64, i32]
    L.mfcc_b200_host_free.restype = None
    L.mfcc_b200_host_free.argtypes = [vp]
    L.mfcc_b200_stage.restype = C.c_int
    L.mfcc_b200_stage.argtypes = [vp, i32, vp, vp, i64, i64, i32, vp]
    L.mfcc_b200_state_bytes.restype = C.c_int
    L.mfcc_b200_state_bytes.argtypes = [vp, i32, C.POINTER(i64)]
    L.mfcc_b200_step.restype = C.c_int
    L.mfcc_b200_step.argtypes = [vp, vp, vp, i32, vp, vp]
    for name in ("mfcc_b200_fft_stream", "mfcc_b200_dct_stream"):
        f = getattr(L, name)
        f.restype = C.c_int
        f.argtypes = [vp, vp, i64, i32, vp, vp]
    L.mfcc_b200_window.restype = C.c_int
    L.mfcc_b200_window.argtypes = [vp, vp, i64, i32, i32, vp, vp]
    L.mfcc_b200_step_host.restype = C.c_int
    L.mfcc_b200_step_host.argtypes = [vp, i32, vp, vp, i32]
    L.mfcc_b200_lifter.restype = C.c_int
    L.mfcc_b200_lifter.argtypes = [vp, vp, i64, i32, vp, vp]
    L.mfcc_b200_standardize.restype = C.c_int
    L.mfcc_b200_standardize.argtypes = [vp, i32, vp, i64, i32, i32, vp]
    L.mfcc_b200_frame.restype = C.c_int
    L.mfcc_b200_frame.argtypes = [vp, vp, i64, i64, i32, i32, i32, vp]
    L.mfcc_b200_power.restype = C.c_int
    L.mfcc_b200_power.argtypes = [vp, vp, i64, i32, vp]
    L.mfcc_b200_log2fix.restype = C.c_int
    L.mfcc_b200_log2fix.argtypes = [vp, vp, i64, i32, i32, vp]
    L.mfcc_b200_filterbank.restype = C.c_int
    L.mfcc_b200_filterbank.argtypes = [vp, vp, i64, i32, i32, i32, i32, i32, i32, vp, i32, vp]
    L.mfcc_b200_to_int16.restype = C.c_int
    L.mfcc_b200_to_int16.argtypes = [vp, vp, i64, vp]
    L.mfcc_b200_uart_frame.restype = C.c_int
    L.mfcc_b200_uart_frame.argtypes = [vp, vp, i64, i32, C.c_uint32, vp]
    L.mfcc_b200_vad_step.restype = C.c_int
    L.mfcc_b200_vad_step.argtypes = [vp, vp, vp, i32, i32, i32, i32, vp, vp, vp]
    L.mfcc_b200_librosa_create.restype = C.c_int
    L.mfcc_b200_librosa_create.argtypes = [C.POINTER(LibrosaCfg), C.POINTER(vp)]
    L.mfcc_b200_librosa_destroy.restype = None
    L.mfcc_b200_librosa_destroy.argtypes = [vp]
    L.mfcc_b200_librosa_frames.restype = C.c_int64
    L.mfcc_b200_librosa_frames.argtypes = [vp, i64]
    L.mfcc_b200_librosa_workspace.restype = C.c_int
    L.mfcc_b200_librosa_workspace.argtypes = [vp, i32, i64, C.POINTER(i64)]
    L.mfcc_b200_librosa_config.restype = C.c_int
    L.mfcc_b200_librosa_config.argtypes = [vp, i64]
    L.mfcc_b200_librosa_dct_path.restype = C.c_int
    L.mfcc_b200_librosa_dct_path.argtypes = [vp, i32]
    L.mfcc_b200_librosa_run.restype = C.c_int
    L.mfcc_b200_librosa_run.argtypes = [vp, vp, i64, i64, i32, i64, vp, vp, i64, C.POINTER(LibrosaTaps), vp]
    L.mfcc_b200_librosa_run_static.restype = C.c_int
    L.mfcc_b200_librosa_run_static.argtypes = [vp, vp, i64, i64, i32, i64, vp, i64, vp]
    L.mfcc_b200_launch_count.restype = C.c_int64
    L.mfcc_b200_launch_count.argtypes = []
    L.mfcc_b200_last_error.restype = C.c_char_p
    L.mfcc_b200_last_error.argtypes = []
    L.mfcc_b200_version.restype = C.c_int
    L.mfcc_b200_version.argtypes = []
    _lib = L
    return L


EXPORTS = ["mfcc_b200_create", "mfcc_b200_destroy", "mfcc_b200_run", "mfcc_b200_run_host",
           "mfcc_b200_host_config", "mfcc_b200_host_alloc", "mfcc_b200_host_free", "mfcc_b200_debug_fail_chunk",
           "mfcc_b200_stage", "mfcc_b200_frame", "mfcc_b200_window", "mfcc_b200_fft_stream",
           "mfcc_b200_dct_stream", "mfcc_b200_power", "mfcc_b200_log2fix", "mfcc_b200_filterbank", "mfcc_b200_state_bytes", "mfcc_b200_step", "mfcc_b200_step_host", "mfcc_b200_lifter",
           "mfcc_b200_standardize", "mfcc_b200_to_int16", "mfcc_b200_uart_frame", "mfcc_b200_vad_step",
           "mfcc_b200_librosa_create", "mfcc_b200_librosa_destroy", "mfcc_b200_librosa_frames",
           "mfcc_b200_librosa_workspace", "mfcc_b200_librosa_config", "mfcc_b200_librosa_dct_path", "mfcc_b200_librosa_run", "mfcc_b200_librosa_run_static",
           "mfcc_b200_launch_count", "mfcc_b200_last_error", "mfcc_b200_version"]


LIBROSA_STATIC_BYTES = 16 * 1024 * 1024       # MFCC_B200_LIBROSA_STATIC_BYTES


def check(rc):
    if rc != OK:
        msg = lib().mfcc_b200_last_error()
        raise MfccError(rc, msg.decode() if msg else "error")


def launch_count() -> int:
    return int(lib().mfcc_b200_launch_count())
