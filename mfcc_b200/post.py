"""What the reference's tool chain does to the cepstra after the path (SURVEY.md 8(f)
rows f2 and f3), as GPU operators over libmfcc_b200.so:

* ``lifter``        software/lift.py:12-26, :38-39
* ``scale``         sklearn.preprocessing.scale(axis=time) + int16 cast, software/genlibrosa.py:25-28
* ``uart_frame``    MagicInserter + UART byte order, mfcc/misc/magic.py:27-39,
                    mfcc/targets/mic2mfcc.py:58-62 (parsed by software/cepstrum.c:15-68)
* ``VoiceGate``     cepstrum_refill_window / cepstrum_eval_power, software/cepstrum.c:93-183
"""
from __future__ import annotations

import ctypes as C
import math

import torch

from . import _lib

NCEPSTRUMS_MIC = 16          # software/recv.c:17
NFRAMES_WINDOW = 93          # software/recv.c:18
POWER_THRESHOLD = 100000000  # software/cepstrum.c:13
MAGIC = 0xa55a               # mfcc/misc/magic.py:11


def _sp(t):
    return C.c_void_p(torch.cuda.current_stream(t.device).cuda_stream)


def _p(t):
    return C.c_void_p(t.data_ptr())


def _need(x, dtype, what):
    if not isinstance(x, torch.Tensor) or x.dtype != dtype or x.device.type != "cuda":
        raise TypeError("%s must be a CUDA %s tensor" % (what, dtype))
    return x.contiguous()


def lifter_coeffs(ncoeff, L=22, device="cuda"):
    """``1 + (L/2) sin(pi n / L)`` in float64 (software/lift.py:21-22)."""
    n = torch.arange(ncoeff, dtype=torch.float64)
    return (1 + (L / 2.) * torch.sin(math.pi * n / L)).to(device)


def lifter(cepstra, L=22):
    """int16 ``[..., ncoeff]`` -> int16, ``(lift * cepstra).astype(np.int16)``; L <= 0
    returns the input unchanged (software/lift.py:24-26)."""
    x = _need(cepstra, torch.int16, "cepstra")
    if L <= 0:
        return x
    nc = x.shape[-1]
    out = torch.empty_like(x)
    with torch.cuda.device(x.device):
        _lib.check(_lib.lib().mfcc_b200_lifter(_p(x), _p(out), x.numel() // nc, nc,
                                               _p(lifter_coeffs(nc, L, x.device)), _sp(x)))
    return out


def scale(cepstra):
    """``[K, ncoeff]`` or ``[B, K, ncoeff]`` int16 / float32 -> int16: every coefficient
    standardised over the K frames of its clip (population std, zero std divides by 1),
    then truncated like ``astype(np.int16)``."""
    if not isinstance(cepstra, torch.Tensor) or cepstra.dtype not in (torch.int16, torch.float32) \
            or cepstra.device.type != "cuda":
        raise TypeError("cepstra must be a CUDA int16 or float32 tensor")
    if cepstra.dim() not in (2, 3):
        raise ValueError("cepstra must be [K, ncoeff] or [B, K, ncoeff]")
    x = cepstra.contiguous()
    B = 1 if x.dim() == 2 else x.shape[0]
    K, nc = x.shape[-2], x.shape[-1]
    out = torch.empty(x.shape, dtype=torch.int16, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(_lib.lib().mfcc_b200_standardize(_p(x), int(x.dtype == torch.float32), _p(out), K, nc,
                                                    B, _sp(x)))
    return out


def to_int16(x):
    """float32 CUDA tensor -> int16 like ``ndarray.astype(np.int16)`` (software/genlibrosa.py:27)."""
    if not isinstance(x, torch.Tensor) or x.dtype != torch.float32 or x.device.type != "cuda":
        raise TypeError("x must be a CUDA float32 tensor")
    x = x.contiguous()
    out = torch.empty(x.shape, dtype=torch.int16, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(_lib.lib().mfcc_b200_to_int16(_p(x), _p(out), x.numel(), _sp(x)))
    return out


def uart_frame(cols, magic=MAGIC):
    """int16 ``[n, ncoeff]`` -> uint8 ``[n, 2 * (ncoeff + 1)]``: the bytes the microphone
    build puts on the serial line for these columns."""
    x = _need(cols, torch.int16, "cols")
    if x.dim() != 2:
        raise ValueError("cols must be [n, ncoeff]")
    n, nc = x.shape
    out = torch.empty((n, 2 * (nc + 1)), dtype=torch.uint8, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(_lib.lib().mfcc_b200_uart_frame(_p(x), _p(out), n, nc, magic, _sp(x)))
    return out


class VoiceGate:
    """Sliding window of the last `nframes` columns per stream; ``step(cols)`` pushes one
    column per stream and returns (power int32 [S], active bool [S]) as
    ``cepstrum_eval_power`` computes them.  The state (ring + head) is two tensors: the
    checkpoint of the gate."""

    def __init__(self, n_streams, ncepstrums=NCEPSTRUMS_MIC, nframes=NFRAMES_WINDOW,
                 threshold=POWER_THRESHOLD, device="cuda"):
        if (nframes * ncepstrums // 3) % ncepstrums:
            raise ValueError("nframes * ncepstrums / 3 must be a multiple of ncepstrums")
        self.n_streams, self.ncepstrums, self.nframes, self.threshold = n_streams, ncepstrums, nframes, threshold
        self.device = torch.device(device)
        self.ring = torch.zeros((n_streams, nframes), dtype=torch.int16, device=self.device)
        self.head = torch.zeros((n_streams,), dtype=torch.int32, device=self.device)
        self.power = torch.zeros((n_streams,), dtype=torch.int32, device=self.device)
        self.active = torch.zeros((n_streams,), dtype=torch.uint8, device=self.device)

    def reset(self):
        self.ring.zero_()
        self.head.zero_()
        return self

    def step(self, cols):
        x = _need(cols, torch.int16, "cols")
        if x.shape != (self.n_streams, self.ncepstrums):
            raise ValueError("cols must be [%d, %d]" % (self.n_streams, self.ncepstrums))
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().mfcc_b200_vad_step(_p(self.ring), _p(self.head), _p(x), self.n_streams,
                                                     self.ncepstrums, self.nframes, self.threshold,
                                                     _p(self.power), _p(self.active), _sp(x)))
        return self.power, self.active.bool()
