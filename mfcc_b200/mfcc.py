"""``MFCC``: the reference's top-level module as a callable GPU operator.

Mirrors ``MFCC(width=16, nfft=512, samplerate=16e3, nfilters=16, nceptrums=16)`` of
/root/reference/mfcc/core/mfcc.py:19-30: same keyword names and defaults, ``reset()``
for the ``reset`` signal, and the stage modules reachable as attributes
(``frame, window, fft_stream, powspec, filterbank, log2, dct_stream, discard``,
mfcc.py:106-114).  ``sink``/``source`` become the argument and the result of
``__call__``: int16 samples in, ``[K, nceptrums]`` int16 cepstra out, one row per
``last`` of the source stream.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from .config import MFCCConfig, MODE_INT, MODE_FP32
from .handle import Handle


def _stream_ptr(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _check_samples(x, device):
    if not isinstance(x, torch.Tensor):
        raise TypeError("samples must be a torch.Tensor")
    if x.dtype != torch.int16:
        raise TypeError("samples must be int16 (sink payload is signed 16, mfcc.py:29)")
    if x.dim() not in (1, 2):
        raise ValueError("samples must be [N] or [B, N]")
    if x.device.type != "cuda":
        raise ValueError("samples must live on the GPU; use run_host() for host buffers")
    if device is not None and x.device != device:
        raise ValueError("samples are on %s, operator is on %s" % (x.device, device))


class MFCC:
    def __init__(self, width=16, nfft=512, samplerate=16e3, nfilters=16, nceptrums=16,
                 mode="int", device=None, mel_floor=0.0):
        if mode not in ("int", "fp32"):
            raise ValueError("mode must be 'int' or 'fp32'")
        self.mode = mode
        # int: the FPGA's fixed point, bit exact.  fp32: the float prototype of
        # notebook/MFCC-INT.ipynb cells 2-10, float32 cepstra; mel_floor > 0 clamps the
        # mel energies before log2 (the notebook has none: silence gives -inf)
        self.out_dtype = torch.int16 if mode == "int" else torch.float32
        self.cfg = MFCCConfig(width=width, nfft=nfft, samplerate=samplerate, nfilters=nfilters,
                              nceptrums=nceptrums, mode=MODE_INT if mode == "int" else MODE_FP32,
                              mel_floor=float(mel_floor) if mode == "fp32" else 0.0)
        self.width, self.nfft, self.samplerate = width, nfft, samplerate
        self.nfilters, self.nceptrums = nfilters, nceptrums
        if not torch.cuda.is_available():
            raise RuntimeError("MFCC needs a CUDA device: there is no CPU fallback")
        self.device = torch.device(device if device is not None else
                                   "cuda:%d" % torch.cuda.current_device())
        with torch.cuda.device(self.device):
            self._h = Handle(self.cfg)
        from . import stages as S   # stage operators share the integer-mode handle
        if mode == "int":
            hs = self._h
        else:
            with torch.cuda.device(self.device):
                hs = Handle(MFCCConfig(width=width, nfft=nfft, samplerate=samplerate,
                                       nfilters=nfilters, nceptrums=nceptrums))
        self.preemph = S.Preemph(width, _handle=hs, device=self.device)
        self.frame = S.Frame(width, nfft, nfft // 3, nfft, _handle=hs, device=self.device)
        self.window = S.WindowHamming(width, nfft, 8, _handle=hs, device=self.device)
        self.fft_stream = S.FftStream(width, nfft, _handle=hs, device=self.device)
        self.powspec = S.PowerSpectrum(width, 30, _handle=hs, device=self.device)
        self.filterbank = S.FilterBank(30, 16, None, 18, samplerate, nfft, nfilters,
                                       _handle=hs, device=self.device)
        self.log2 = S.Log2Fix(16, 15, _handle=hs, device=self.device)
        self.dct_stream = S.DCTStream(width, nfilters, _handle=hs, device=self.device)
        self.discard = S.Discard(width, 0, nceptrums)
        self._state = None

    # ------------------------------------------------------------------ batch
    def num_frames(self, n_samples: int) -> int:
        return self.cfg.num_frames(n_samples)

    def __call__(self, samples, n_frames=None, taps=False, out=None):
        """samples: int16 ``[N]`` or ``[B, N]`` on the GPU, every row a stream after
        reset.  Returns ``[K, nceptrums]`` / ``[B, K, nceptrums]`` int16.
        n_frames > num_frames(N) zero-extends the stream (software/main.c:135-144).
        taps=True (one clip) also returns every stage output, the analogue of the
        bench collectors at mfcc.py:176-183."""
        _check_samples(samples, self.device)
        single = samples.dim() == 1
        x = samples if samples.is_contiguous() or samples.stride(-1) == 1 else samples.contiguous()
        if x.stride(-1) != 1:
            x = x.contiguous()
        B = 1 if single else x.shape[0]
        N = x.shape[-1]
        K = self.num_frames(N) if n_frames is None else int(n_frames)
        if K < 0:
            raise ValueError("n_frames must be >= 0")
        nc = self.nceptrums
        if out is None:
            out = torch.empty((B, K, nc), dtype=self.out_dtype, device=self.device)
        elif out.shape != (B, K, nc) or out.dtype != self.out_dtype or not out.is_contiguous():
            raise ValueError("out must be a contiguous %s [B, K, nceptrums] tensor" % self.out_dtype)
        tp = None
        tensors = None
        if taps:
            if self.mode != "int":
                raise ValueError("taps are an integer-mode feature")
            if B != 1:
                raise ValueError("taps need a single clip")
            nf, nb = self.nfilters, self.nfft // 2
            dev = self.device
            tensors = dict(
                window=torch.empty((K, self.nfft), dtype=torch.int16, device=dev),
                fft=torch.empty((K, nb, 2), dtype=torch.int16, device=dev),
                power=torch.empty((K, nb), dtype=torch.int32, device=dev),
                filter=torch.empty((K, nf), dtype=torch.int16, device=dev),
                log=torch.empty((K, nf), dtype=torch.int16, device=dev),
                dct=torch.empty((K, nf), dtype=torch.int16, device=dev))
            tp = _lib.Taps(*[tensors[n].data_ptr() for n in
                             ("window", "fft", "power", "filter", "log", "dct")])
        stride = N if single else x.stride(0)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().mfcc_b200_run(
                self._h.ptr, C.c_void_p(x.data_ptr()), stride, N, B, K,
                C.c_void_p(out.data_ptr()), K * nc, C.byref(tp) if tp is not None else None,
                _stream_ptr(self.device)))
        res = out[0] if single else out
        return (res, tensors) if taps else res

    def run_host(self, samples, n_frames=None, out=None):
        """Host buffers in, host buffers out (pinned buffers give full PCIe rate):
        the call a host tool in the place of software/main.c:100-177 makes."""
        if samples.dtype != torch.int16 or samples.device.type != "cpu":
            raise TypeError("run_host takes a CPU int16 tensor")
        single = samples.dim() == 1
        x = samples.contiguous()
        B = 1 if single else x.shape[0]
        N = x.shape[-1]
        K = self.num_frames(N) if n_frames is None else int(n_frames)
        nc = self.nceptrums
        if out is None:
            out = torch.empty((B, K, nc), dtype=self.out_dtype)
        elif out.shape != (B, K, nc) or out.dtype != self.out_dtype or not out.is_contiguous() \
                or out.device.type != "cpu":
            raise ValueError("out must be a contiguous CPU %s [B, K, nceptrums] tensor" % self.out_dtype)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().mfcc_b200_run_host(
                self._h.ptr, C.c_void_p(x.data_ptr()), N, N, B, K, C.c_void_p(out.data_ptr()),
                K * nc))
        return out[0] if single else out

    # ------------------------------------------------------------------ streaming
    def reset(self, n_streams: int = 0):
        """The ``reset`` signal (mfcc.py:28, :116): every stream back to power-on
        state.  With n_streams > 0 allocates the streaming state."""
        if n_streams:
            nbytes = C.c_int64()
            _lib.check(_lib.lib().mfcc_b200_state_bytes(self._h.ptr, n_streams, C.byref(nbytes)))
            self._state = torch.zeros((n_streams, nbytes.value // (2 * n_streams)),
                                      dtype=torch.int16, device=self.device)
        elif self._state is not None:
            self._state.zero_()
        return self

    def step(self, hop, out=None, first_stream=None):
        """One hop (nfft//3 samples) per stream in, one cepstrum column per stream
        out: ``[S, hop]`` int16 -> ``[S, nceptrums]`` int16.  The state after reset()
        is the core with nfft-hop zeros already pushed in.

        With ``first_stream`` the rows are the hops of streams ``first_stream ..
        first_stream + S - 1`` of a state allocated by ``reset(n_streams)``: groups of
        streams can then be stepped on different CUDA streams, so that the copy-in of
        one group overlaps the kernel of another.

        ``hop`` and ``out`` may also be PINNED host tensors: under unified addressing the
        kernel then reads the hops and writes the columns over PCIe itself (zero-copy),
        which removes the two copy operations from a latency-critical step."""
        if isinstance(hop, torch.Tensor) and hop.device.type == "cpu":
            if not hop.is_pinned() or hop.dtype != torch.int16:
                raise ValueError("a host hop tensor must be pinned int16")
            if out is not None and out.device.type == "cpu" and not out.is_pinned():
                raise ValueError("a host out tensor must be pinned")
        else:
            _check_samples(hop, self.device)
        if hop.dim() != 2 or hop.shape[1] != self.cfg.hop:
            raise ValueError("hop must be [S, %d]" % self.cfg.hop)
        S = hop.shape[0]
        if first_stream is None:
            if self._state is None or self._state.shape[0] != S:
                self.reset(S)
            state = self._state
        else:
            if self._state is None or first_stream < 0 or first_stream + S > self._state.shape[0]:
                raise ValueError("streams %d..%d are outside the state allocated by reset()"
                                 % (first_stream, first_stream + S - 1))
            state = self._state[first_stream:first_stream + S]
        if not hop.is_contiguous():
            hop = hop.contiguous()
        if out is None:
            out = torch.empty((S, self.nceptrums), dtype=self.out_dtype, device=self.device)
        elif out.shape != (S, self.nceptrums) or out.dtype != self.out_dtype or not out.is_contiguous():
            raise ValueError("out must be contiguous %s [%d, %d]" % (self.out_dtype, S, self.nceptrums))
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().mfcc_b200_step(
                self._h.ptr, C.c_void_p(state.data_ptr()), C.c_void_p(hop.data_ptr()), S,
                C.c_void_p(out.data_ptr()), _stream_ptr(self.device)))
        return out
