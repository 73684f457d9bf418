"""One callable per stage module of the reference, same constructor names and
defaults; each maps a GPU tensor to a GPU tensor.  The parameter
combinations MFCC.elaborate generates (/root/reference/mfcc/core/mfcc.py:38-88) run the
stage kernels that share their device code with the fused kernel (``mfcc_b200_stage``);
other parameter sets of the same classes run handle-free general kernels.  Width is 16
throughout, as everywhere in the reference.

16-bit unsigned streams (mel, log) are carried in int16 tensors and 30-bit powers in
int32 tensors: same bits, torch has no general uint16/uint32 arithmetic.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib, tables
from .config import MFCCConfig
from .handle import Handle


# mfcc/misc/fft.py:251-308, :344: a stage lasts size/2 cycles, a butterfly's results land 8
# cycles after its read; tests/test_rtl_golden.py::test_small_fft_hazard_of_the_reference
_HAZARD = ("a %d-point FFT core: below 32 points the reference's scheduler has a read-before-write "
           "hazard and its output is not a transform; not reproduced")


class _Stage:
    stage_id = -1

    def _setup(self, handle, device, nfilters=32):
        if not torch.cuda.is_available():
            raise RuntimeError("stage operators need a CUDA device: there is no CPU fallback")
        self.device = torch.device(device if device is not None else
                                   "cuda:%d" % torch.cuda.current_device())
        if handle is None:
            with torch.cuda.device(self.device):
                handle = Handle(MFCCConfig(nfilters=nfilters, nceptrums=nfilters))
        self._h = handle

    def _general_device(self, device):
        """Handle-free operators (the general kernels take their tables per call)."""
        if not torch.cuda.is_available():
            raise RuntimeError("stage operators need a CUDA device: there is no CPU fallback")
        self.device = torch.device(device if device is not None else
                                   "cuda:%d" % torch.cuda.current_device())

    def _check(self, x, dtype, last_dims):
        if not isinstance(x, torch.Tensor) or x.dtype != dtype:
            raise TypeError("%s takes a %s tensor" % (type(self).__name__, dtype))
        if x.device != self.device:
            raise ValueError("input is on %s, operator on %s" % (x.device, self.device))
        if tuple(x.shape[-len(last_dims):]) != tuple(last_dims):
            raise ValueError("%s: trailing dimensions must be %s" % (type(self).__name__,
                                                                      (last_dims,)))
        return x.contiguous()

    def _run(self, x, out, n, n_frames=0, prev=0):
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().mfcc_b200_stage(
                self._h.ptr, self.stage_id, C.c_void_p(x.data_ptr()), C.c_void_p(out.data_ptr()),
                n, n_frames, prev, C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)))
        return out


class Preemph(_Stage):
    """Preemph(width=16), mfcc/core/preemph.py:9-30.  int16 [N] -> int16 [N]."""
    stage_id = _lib.STAGE_PREEMPH

    def __init__(self, width=16, _handle=None, device=None):
        if width != 16:
            raise ValueError("width must be 16")
        self.width = width
        self._setup(_handle, device)

    def __call__(self, x, prev=0):
        if x.dim() != 1:
            raise ValueError("Preemph takes one stream [N]")
        x = self._check(x, torch.int16, (x.shape[-1],))
        return self._run(x, torch.empty_like(x), x.numel(), 0, int(prev))


class Frame(_Stage):
    """Frame(width=16, windowlen=400, stepsize=160, nfft=512), mfcc/core/frame.py:49-60, any
    parameter set: int16 [N] -> int16 [K, nfft], frame k = y[stepsize*k : stepsize*k +
    windowlen] then nfft - windowlen zeros.  K = full frames unless n_frames is given (the
    stream is then zero-extended, as the reference's hosts do)."""
    stage_id = _lib.STAGE_FRAME

    def __init__(self, width=16, windowlen=400, stepsize=160, nfft=512, _handle=None, device=None):
        if windowlen > nfft:
            raise ValueError("windowlen must be <= nfft")          # frame.py:51
        if width != 16:
            raise ValueError("width must be 16")
        if windowlen < 1 or stepsize < 1:
            raise ValueError("windowlen and stepsize must be >= 1")
        self.width, self.windowlen, self.stepsize, self.nfft = width, windowlen, stepsize, nfft
        if not torch.cuda.is_available():
            raise RuntimeError("stage operators need a CUDA device: there is no CPU fallback")
        self.device = torch.device(device if device is not None else
                                   "cuda:%d" % torch.cuda.current_device())

    def __call__(self, y, n_frames=None):
        if y.dim() != 1:
            raise ValueError("Frame takes one stream [N]")
        y = self._check(y, torch.int16, (y.shape[-1],))
        n = y.numel()
        k = (0 if n < self.windowlen else (n - self.windowlen) // self.stepsize + 1) if n_frames is None \
            else int(n_frames)
        out = torch.empty((k, self.nfft), dtype=torch.int16, device=self.device)
        if k == 0:
            return out
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().mfcc_b200_frame(
                C.c_void_p(y.data_ptr() if n else 0), C.c_void_p(out.data_ptr()), n, k, self.windowlen,
                self.stepsize, self.nfft, C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)))
        return out


class WindowHamming(_Stage):
    """WindowHamming(width=16, nfft=512, precision=8), mfcc/core/window.py:10-20, any
    power-of-two nfft >= 8 and any precision whose quarter table fits (window.py:37).
    int16 [K, nfft] -> int16 [K, nfft].  MFCC's own set (16, 512, 8) runs the stage kernel
    of the fused path's tables, the others the general kernel with their own curve."""
    stage_id = _lib.STAGE_WINDOW

    def __init__(self, width=16, nfft=512, precision=8, _handle=None, device=None):
        if nfft < 8:
            raise ValueError("nfft must be >= 8")                  # window.py:13
        if nfft & (nfft - 1):
            raise ValueError("nfft must be a power of two")        # the address slices of window.py:59-66
        if width != 16:
            raise ValueError("width must be 16")
        if not 1 <= precision <= 14:
            raise ValueError("precision must be 1..14")
        self.width, self.nfft, self.precision = width, nfft, precision
        self._fast = (nfft, precision) == (512, 8)
        if self._fast:
            self._setup(_handle, device)
            return
        curve = tables.window_curve(nfft, precision)               # raises ValueError like window.py:37
        self._general_device(device)
        self._curve = torch.from_numpy(curve.astype("uint16")).to(self.device)

    def __call__(self, frames):
        x = self._check(frames, torch.int16, (self.nfft,))
        out = torch.empty_like(x)
        if self._fast:
            return self._run(x, out, x.numel() // self.nfft)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().mfcc_b200_window(
                C.c_void_p(x.data_ptr()), C.c_void_p(out.data_ptr()), x.numel() // self.nfft, self.nfft,
                self.precision, C.c_void_p(self._curve.data_ptr()),
                C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)))
        return out


class FftStream(_Stage):
    """FftStream(width=16, nfft=512), mfcc/core/fft_stream.py:6-11, any power-of-two nfft in
    32..4096.  int16 [K, nfft] -> int16 [K, nfft/2, 2] (data_r, data_i).  nfft = 512 runs the
    warp-per-frame stage kernel, other sizes the general FFT-core kernel."""
    stage_id = _lib.STAGE_FFT

    def __init__(self, width=16, nfft=512, _handle=None, device=None):
        if nfft <= 0 or nfft & (nfft - 1):
            raise ValueError("Size must be a positive power-of-two integer, not %r" % (nfft,))   # fft.py:351-353
        if width != 16:
            raise ValueError("width must be 16")
        if nfft < 32:
            raise ValueError(_HAZARD % nfft)
        if nfft > 4096:
            raise ValueError("nfft must be 32..4096")
        self.width, self.nfft = width, nfft
        self._fast = nfft == 512
        if self._fast:
            self._setup(_handle, device)
            return
        self._general_device(device)
        self._tw = torch.from_numpy(tables.twiddle_table(nfft, 16).astype("int16")).to(self.device)

    def __call__(self, frames):
        x = self._check(frames, torch.int16, (self.nfft,))
        out = torch.empty(x.shape[:-1] + (self.nfft // 2, 2), dtype=torch.int16, device=self.device)
        if self._fast:
            return self._run(x, out, x.numel() // self.nfft)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().mfcc_b200_fft_stream(
                C.c_void_p(x.data_ptr()), C.c_void_p(out.data_ptr()), x.numel() // self.nfft, self.nfft,
                C.c_void_p(self._tw.data_ptr()), C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)))
        return out


class PowerSpectrum(_Stage):
    """PowerSpectrum(width=16, width_output=24), mfcc/core/pow2.py:43-50, any output width:
    int16 [..., 2] (re, im) -> int32 [...] holding the unsigned top ``width_output`` bits of
    the 32-bit sum of squares (MFCC uses 30, mfcc.py:60-62)."""
    stage_id = _lib.STAGE_POWER

    def __init__(self, width=16, width_output=24, multiplier_cls=None, _handle=None, device=None):
        if width != 16:
            raise ValueError("width must be 16")
        if not 1 <= width_output <= 32:
            raise ValueError("width_output must be 1..32")
        self.width, self.width_output = width, width_output
        if not torch.cuda.is_available():
            raise RuntimeError("stage operators need a CUDA device: there is no CPU fallback")
        self.device = torch.device(device if device is not None else
                                   "cuda:%d" % torch.cuda.current_device())

    def __call__(self, spec):
        x = self._check(spec, torch.int16, (2,))
        out = torch.empty(x.shape[:-1], dtype=torch.int32, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().mfcc_b200_power(
                C.c_void_p(x.data_ptr()), C.c_void_p(out.data_ptr()), out.numel(), self.width_output,
                C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)))
        return out


class FilterBank(_Stage):
    """FilterBank(width=24, width_output=24, width_mul=None, gain=0, sample_rate=16000,
    nfft=512, ntap=16), mfcc/core/filterbank.py:37-63.  With the parameters MFCC uses
    (30, 16, gain 18, 16 or 32 taps: mfcc.py:69-75) int32 [K, 256] -> int16 [K, ntap]
    (16-bit unsigned values that wrap, filterbank.py:122) through the warp-per-frame stage
    kernel; with any other parameter set int32 [K, nfft/2] -> int32 [K, ntap] (the unsigned
    width_output-bit values) through the general register-level kernel."""
    stage_id = _lib.STAGE_MEL

    def __init__(self, width=24, width_output=24, width_mul=None, gain=0, sample_rate=16000,
                 nfft=512, ntap=16, multiplier_cls=None, _handle=None, device=None, _general=False):
        import math
        width_mul = width if width_mul is None else width_mul
        self.width, self.width_output, self.width_mul, self.gain = width, width_output, width_mul, gain
        self.sample_rate, self.nfft, self.ntap = sample_rate, nfft, ntap
        self._fast = (not _general and (width, width_output, width_mul, gain, nfft) == (30, 16, 30, 18, 512)
                      and ntap in (16, 32))
        if self._fast:
            if _handle is not None and (_handle.cfg.nfilters != ntap or
                                        _handle.cfg.samplerate != sample_rate):
                _handle = None
            try:
                if _handle is None and torch.cuda.is_available():
                    dev = torch.device(device if device is not None else
                                       "cuda:%d" % torch.cuda.current_device())
                    with torch.cuda.device(dev):
                        _handle = Handle(MFCCConfig(samplerate=sample_rate, nfilters=ntap, nceptrums=ntap))
                self._setup(_handle, device, ntap)
                return
            except _lib.MfccError as e:          # a sample rate whose mel segments the warp kernel's table cannot hold
                if e.code != _lib.EUNSUPPORTED:
                    raise
                self._fast = False
        if not (1 <= width <= 31 and 1 <= width_mul <= 31 and 1 <= width_output <= 32 and ntap >= 1):
            raise ValueError("FilterBank: width and width_mul must be <= 31, width_output <= 32")
        if not torch.cuda.is_available():
            raise RuntimeError("stage operators need a CUDA device: there is no CPU fallback")
        self.device = torch.device(device if device is not None else
                                   "cuda:%d" % torch.cuda.current_device())
        pts = tables.mel_points(ntap, nfft, sample_rate)
        steps = tables.mel_steps(pts, wsize=width_mul)                       # calc_filters, filterbank.py:22-34
        self.maxvalrange = int(math.log2(pts[-1] - pts[-3])) + width + width_mul       # :71
        self.lo_bit = self.maxvalrange - (gain + width_output)                         # :122
        if self.lo_bit < 0 or self.maxvalrange > 127:
            raise ValueError("FilterBank: gain + width_output exceeds the accumulator width")
        self._steps = torch.tensor([int(v) for v in steps], dtype=torch.int64, device=self.device)

    def __call__(self, power):
        if self._fast:
            x = self._check(power, torch.int32, (256,))
            out = torch.empty(x.shape[:-1] + (self.ntap,), dtype=torch.int16, device=self.device)
            return self._run(x, out, x.numel() // 256)
        nb = self.nfft // 2
        x = self._check(power, torch.int32, (nb,))
        out = torch.empty(x.shape[:-1] + (self.ntap,), dtype=torch.int32, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().mfcc_b200_filterbank(
                C.c_void_p(x.data_ptr()), C.c_void_p(out.data_ptr()), x.numel() // nb, nb, self.ntap,
                self.width_mul, self.maxvalrange, self.lo_bit, self.width_output,
                C.c_void_p(self._steps.data_ptr()), self._steps.numel(),
                C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)))
        return out


class Log2Fix(_Stage):
    """Log2Fix(width, width_output), mfcc/core/log.py:107-116.  Elementwise.  With the widths
    MFCC uses (16, 15: mfcc.py:82) int16 (16-bit unsigned) -> int16 (Q4.11) through the
    table-driven stage kernel; with any other widths int64 -> int32 through the general
    kernel (the reference's own bench uses 37 and 20, log.py:143)."""
    stage_id = _lib.STAGE_LOG2

    def __init__(self, width, width_output, multiplier_cls=None, _handle=None, device=None, _general=False):
        import math
        self.width, self.width_output = width, width_output
        self.precision = width_output - math.ceil(math.log2(width)) if width >= 2 else 0     # log.py:114
        self._fast = (width, width_output) == (16, 15) and not _general
        if self._fast:
            self._setup(_handle, device)
            return
        if width < 2 or not 1 <= width_output <= 32 or not 1 <= self.precision <= 30 or width + self.precision > 63:
            raise ValueError("Log2Fix(%d, %d): widths outside what the kernel covers" % (width, width_output))
        if not torch.cuda.is_available():
            raise RuntimeError("stage operators need a CUDA device: there is no CPU fallback")
        self.device = torch.device(device if device is not None else
                                   "cuda:%d" % torch.cuda.current_device())

    def __call__(self, mel):
        if self._fast:
            nf = self._h.cfg.nfilters
            if mel.numel() % nf:
                raise ValueError("Log2Fix: element count must be a multiple of nfilters=%d" % nf)
            x = self._check(mel, torch.int16, (mel.shape[-1],))
            return self._run(x, torch.empty_like(x), x.numel() // nf)
        x = self._check(mel, torch.int64, (mel.shape[-1],))
        out = torch.empty(x.shape, dtype=torch.int32, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().mfcc_b200_log2fix(
                C.c_void_p(x.data_ptr()), C.c_void_p(out.data_ptr()), x.numel(), self.width, self.width_output,
                C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)))
        return out


class DCTStream(_Stage):
    """DCTStream(width=16, nfft=16), mfcc/core/dct_stream.py:6-11 (nfft = number of
    inputs), any power-of-two nfft in 8..1024.  int16 [K, nfft] -> int16 [K, nfft].
    16 and 32 inputs (the sets MFCC builds, mfcc.py:84) run the warp-level stage kernel,
    other sizes the general FFT-core kernel on the 4*nfft-point transform."""
    stage_id = _lib.STAGE_DCT

    def __init__(self, width=16, nfft=16, _handle=None, device=None):
        if width != 16:
            raise ValueError("width must be 16")
        if nfft <= 0 or nfft & (nfft - 1):
            raise ValueError("Size must be a positive power-of-two integer, not %r" % (4 * nfft,))  # fft.py:351-353
        if nfft < 8:
            raise ValueError(_HAZARD % (4 * nfft))
        if nfft > 1024:
            raise ValueError("nfft must be 8..1024")
        self.width, self.nfft = width, nfft
        self._fast = nfft in (16, 32)
        if self._fast:
            if _handle is not None and _handle.cfg.nfilters != nfft:
                _handle = None
            self._setup(_handle, device, nfft)
            return
        self._general_device(device)
        self._tw = torch.from_numpy(tables.twiddle_table(4 * nfft, 16).astype("int16")).to(self.device)

    def __call__(self, logs):
        x = self._check(logs, torch.int16, (self.nfft,))
        out = torch.empty_like(x)
        if self._fast:
            return self._run(x, out, x.numel() // self.nfft)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().mfcc_b200_dct_stream(
                C.c_void_p(x.data_ptr()), C.c_void_p(out.data_ptr()), x.numel() // self.nfft, self.nfft,
                C.c_void_p(self._tw.data_ptr()), C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)))
        return out


class Discard:
    """Discard(width=16, first=0, count=16), mfcc/misc/discard.py:5-13: keeps elements
    first .. first+count-1 of each row.  A view; no kernel."""

    def __init__(self, width=16, first=0, count=16):
        self.width, self.first, self.count = width, first, count
        self.last = first + count - 1

    def __call__(self, x):
        return x[..., self.first: self.first + self.count]
