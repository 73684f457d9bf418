"""TEST INFRASTRUCTURE — ctypes binding of the C oracle (oracle/mfcc_ref.c).

Only tests/, __graft_entry__.smoke() and bench.py may import this.  Tables come
from oracle/ref_np.py (built the way the reference builds them), never from the
product package.
"""
from __future__ import annotations

import ctypes as C
import math
import os
import subprocess

import numpy as np

from . import ref_np as R

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libmfcc_ref.so")


class _Cfg(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "nfft", "hop", "nfilters", "nceptrums", "window_shift", "pow_shift", "mel_wbits",
        "mel_lo_bit", "mel_out_bits", "log_precision", "log_out_bits", "reserved")] + [
        ("window_curve", C.c_void_p), ("tw_fft", C.c_void_p), ("tw_dct", C.c_void_p),
        ("mel_step", C.c_void_p)]


class _Taps(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("window", "fft", "power", "mel", "log2", "dct")]


def build(force: bool = False) -> str:
    """Compile oracle/libmfcc_ref.so with the recipe in oracle/Makefile."""
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(
            os.path.join(_HERE, "mfcc_ref.c")):
        subprocess.check_call(["make", "-C", _HERE, "-s"] + (["-B"] if force else []))
    return _SO


FAST_FLAGS = "-O3 -march=native -fopenmp"


def _cpu_tag() -> str:
    """Hash of this machine's CPU feature flags: a -march=native binary is only valid here."""
    import hashlib
    flags = ""
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("flags"):
                flags = line
                break
    except OSError:
        pass
    return hashlib.sha1(flags.encode()).hexdigest()[:10]


def build_fast(force: bool = False) -> str:
    """Compile the timed CPU baseline (oracle/mfcc_ref_fast.c, gcc -O3 -march=native -fopenmp)
    on THIS machine; the file name carries a hash of the CPU flags so that a binary built
    elsewhere is never loaded."""
    so = os.path.join(_HERE, "libmfcc_ref_fast.%s.so" % _cpu_tag())
    src = os.path.join(_HERE, "mfcc_ref_fast.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B", "fast", "FAST_SO=" + os.path.basename(so)])
    return so


_lib = None
_fast = None


def _bind_run(L):
    L.mfcc_ref_run.restype = C.c_int
    L.mfcc_ref_run.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int32,
                               C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_int32]
    return L


def fast_lib():
    global _fast
    if _fast is None:
        _fast = _bind_run(C.CDLL(build_fast()))
    return _fast


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.mfcc_ref_run.restype = C.c_int
        _lib.mfcc_ref_run.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int32,
                                      C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_int32]
        _lib.mfcc_ref_max_threads.restype = C.c_int
        _lib.mfcc_ref_fft.restype = None
        _lib.mfcc_ref_fft.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_int64]
        _lib.mfcc_ref_log2.restype = None
        _lib.mfcc_ref_log2.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int32]
    return _lib


def _decoded_twiddles(size):
    return np.ascontiguousarray(
        np.array([R.twiddle(size, k) for k in range(size // 2)], dtype=np.int32))


class RefConfig:
    """Holds the tables (kept alive) and the C struct for one configuration."""

    def __init__(self, nfft=512, samplerate=16e3, nfilters=32, nceptrums=32):
        self.nfft, self.nfilters, self.nceptrums = nfft, nfilters, nceptrums
        self.hop = nfft // 3
        self.curve = np.ascontiguousarray(R.window_curve(nfft, 8, 1)[0].astype(np.int32))
        self.tw_fft = _decoded_twiddles(nfft)
        self.tw_dct = _decoded_twiddles(4 * nfilters)
        pts = R.get_filter_points(0, samplerate / 2, nfilters, nfft, sample_rate=samplerate)
        self.points = pts
        self.steps = np.array(R.calc_filters(pts, wsize=30), dtype=np.uint64)
        maxvalrange = int(math.log2(pts[-1] - pts[-3])) + 30 + 30
        prec = 15 - math.ceil(math.log2(16))
        self.c = _Cfg(nfft, self.hop, nfilters, nceptrums, 9, 2, 30, maxvalrange - (18 + 16), 16,
                      prec, 15, 0,
                      self.curve.ctypes.data, self.tw_fft.ctypes.data, self.tw_dct.ctypes.data,
                      self.steps.ctypes.data)

    def num_frames(self, n):
        return R.num_frames(n, self.nfft, self.hop)


def max_threads() -> int:
    """Host threads the timed runs use: the cores this process may run on.  Not
    omp_get_max_threads(): torchrun exports OMP_NUM_THREADS=1 to every rank, which would
    silently turn the all-cores CPU baseline into a single-thread one."""
    import os
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or int(lib().mfcc_ref_max_threads()))


def run(cfg: RefConfig, x: np.ndarray, n_frames=None, taps=False, threads=1, fast=False):
    """x: int16 [N] or [B, N].  Returns int16 [K, nceptrums] / [B, K, nceptrums]
    (and a dict of per-stage arrays of clip 0 when taps=True).  fast=True runs the tuned
    build (mfcc_ref_fast.c, the timed CPU baseline) instead of the checker."""
    if fast and taps:
        raise ValueError("the tuned build has no taps")
    x = np.ascontiguousarray(x, dtype=np.int16)
    single = x.ndim == 1
    x2 = x[None, :] if single else x
    B, N = x2.shape
    K = cfg.num_frames(N) if n_frames is None else int(n_frames)
    out = np.zeros((B, K, cfg.nceptrums), dtype=np.int16)
    t = None
    tp = None
    if taps:
        nb, nf = cfg.nfft // 2, cfg.nfilters
        t = dict(window=np.zeros((K, cfg.nfft), np.int16), fft=np.zeros((K, nb, 2), np.int16),
                 power=np.zeros((K, nb), np.uint32), mel=np.zeros((K, nf), np.uint16),
                 log2=np.zeros((K, nf), np.uint16), dct=np.zeros((K, nf), np.int16))
        tp = _Taps(*[t[n].ctypes.data for n in ("window", "fft", "power", "mel", "log2", "dct")])
    rc = (fast_lib() if fast else lib()).mfcc_ref_run(C.byref(cfg.c), x2.ctypes.data, N, N, B, K, out.ctypes.data,
                            K * cfg.nceptrums, C.byref(tp) if tp is not None else None, threads)
    if rc != 0:
        raise RuntimeError("mfcc_ref_run failed: %d" % rc)
    res = out[0] if single else out
    return (res, t) if taps else res


def fft(xr, xi, size):
    """Integer FFT of [n, size] int32 arrays (natural-order input); returns copies."""
    vr = np.ascontiguousarray(xr, dtype=np.int32).copy().reshape(-1, size)
    vi = np.ascontiguousarray(xi, dtype=np.int32).copy().reshape(-1, size)
    tw = _decoded_twiddles(size)
    lib().mfcc_ref_fft(vr.ctypes.data, vi.ctypes.data, size, tw.ctypes.data, vr.shape[0])
    return vr, vi


def log2(m, prec=11, out_bits=15):
    m = np.ascontiguousarray(m, dtype=np.uint32)
    o = np.zeros_like(m)
    lib().mfcc_ref_log2(m.ctypes.data, o.ctypes.data, m.size, prec, out_bits)
    return o
