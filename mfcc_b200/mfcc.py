"""``MFCC``: the reference's top-level module as a callable GPU operator.

Mirrors ``MFCC(width=16, nfft=512, samplerate=16e3, nfilters=16, nceptrums=16)`` of
/root/reference/mfcc/core/mfcc.py:19-30: same keyword names and defaults, ``reset()``
for the ``reset`` signal, and the stage modules reachable as attributes
(``frame, window, fft_stream, powspec, filterbank, log2, dct_stream, discard``,
mfcc.py:106-114).  ``sink``/``source`` become the argument and the result of
``__call__``: int16 samples in, ``[K, nceptrums]`` int16 cepstra out, one row per
``last`` of the source stream.

The parameter sets of the reference's tops (nfft 512, 16 or 32 filters) run the fused
kernel; any other set the reference class accepts (power-of-two nfft 32..4096, 8..1024
filters, any sample rate) runs the same chain as one general stage kernel per module.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from .config import MFCCConfig, MODE_INT, MODE_FP32
from .handle import Handle


def _stream_ptr(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


class _PinnedBlock:
    """Owner of one mfcc_b200_host_alloc block (freed when the last tensor view dies)."""

    def __init__(self, nbytes, write_combined):
        p = C.c_void_p()
        _lib.check(_lib.lib().mfcc_b200_host_alloc(C.byref(p), nbytes, 1 if write_combined else 0))
        self.ptr, self.nbytes = p.value, nbytes

    def __del__(self):
        try:
            _lib.lib().mfcc_b200_host_free(C.c_void_p(self.ptr))
        except Exception:       # noqa: BLE001 - interpreter shutdown
            pass


def pinned_empty(shape, dtype=torch.int16, write_combined=False):
    """A page-locked host tensor from the library's own allocator (mfcc_b200_host_alloc),
    optionally write-combined (for sample buffers the CPU only fills).  Needs a CUDA device."""
    n = 1
    for d in shape:
        n *= int(d)
    nbytes = max(1, n) * torch.empty((), dtype=dtype).element_size()
    blk = _PinnedBlock(nbytes, write_combined)
    buf = (C.c_char * nbytes).from_address(blk.ptr)
    buf._owner = blk                                 # keeps the block alive as long as the tensor's buffer
    return torch.frombuffer(buf, dtype=dtype, count=n).reshape(tuple(shape))


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
        from . import stages as S   # stage operators share the integer-mode handle
        self._composed = False
        try:
            with torch.cuda.device(self.device):
                self._h = Handle(self.cfg)
        except (ValueError, _lib.MfccError) as e:
            # the fused kernels are built for the sets the reference's tops use (nfft 512,
            # 16 or 32 filters, mel segments of at most 21 / 42 bins); any other set the
            # reference class accepts runs as the chain of general stage kernels
            if isinstance(e, _lib.MfccError) and e.code != _lib.EUNSUPPORTED:
                raise
            if mode != "int":
                raise ValueError("fp32 mode covers nfft=512 with 16 or 32 filters (%s)" % e) from None
            self._h = None
            self._composed = True
        if self._composed:
            self.preemph = S.Preemph(width, device=self.device)
            self.frame = S.Frame(width, nfft, nfft // 3, nfft, device=self.device)
            self.window = S.WindowHamming(width, nfft, 8, device=self.device)
            self.fft_stream = S.FftStream(width, nfft, device=self.device)
            self.powspec = S.PowerSpectrum(width, 30, device=self.device)
            self.filterbank = S.FilterBank(30, 16, None, 18, samplerate, nfft, nfilters,
                                           device=self.device, _general=True)
            self.log2 = S.Log2Fix(16, 15, device=self.device, _general=True)
            self.dct_stream = S.DCTStream(width, nfilters, device=self.device)
            self.discard = S.Discard(width, 0, nceptrums)
            self._state = None
            return
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
        if self._composed:
            return self._call_composed(x if not single else x[None], K, out, single, taps)
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

    # ------------------------------------------------------------------ other parameter sets
    def _chain(self, frames):
        """frames int16 [..., nfft] -> every later stage, as int16 / int32 tensors."""
        t = {}
        t["window"] = self.window(frames)
        t["fft"] = self.fft_stream(t["window"])
        t["power"] = self.powspec(t["fft"])
        mel = self.filterbank(t["power"])                       # int32 holding 16-bit unsigned values
        lg = self.log2(mel.to(torch.int64))                     # int32, Q4.11
        t["filter"], t["log"] = mel.to(torch.int16), lg.to(torch.int16)      # same bits, the fused path's dtypes
        t["dct"] = self.dct_stream(t["log"])
        return self.discard(t["dct"]), t

    def _call_composed(self, x, K, out, single, taps):
        """MFCC with constructor arguments the fused kernel is not built for: the chain of
        mfcc.py:38-104 as one general stage kernel per module (stages.py).  Each row is
        zero-extended at the INPUT of the chain (software/main.c:135-144), given a leading
        zero (the reset value of the pre-emphasis register, preemph.py:18-24) and
        pre-emphasised in one launch over the flattened batch; Frame runs per row."""
        if taps and x.shape[0] != 1:
            raise ValueError("taps need a single clip")
        B, N = x.shape
        nfft, hop = self.nfft, self.cfg.hop
        if K == 0:
            return (out[0] if single else out, {}) if taps else (out[0] if single else out)
        need = max(N, hop * (K - 1) + nfft)
        rows = torch.zeros((B, need + 1), dtype=torch.int16, device=self.device)
        rows[:, 1:N + 1] = x
        y = self.preemph(rows.reshape(-1)).reshape(B, need + 1)[:, 1:]
        frames = torch.stack([self.frame(y[b].contiguous(), n_frames=K) for b in range(B)])
        res, t = self._chain(frames)
        out.copy_(res)
        if taps:
            t = {k: v[0] for k, v in t.items()}
            t["frame"] = frames[0]
            return out[0], t
        return out[0] if single else out

    def host_config(self, chunk_mb=None, slots=None):
        """Staging pipeline of run_host (mfcc_b200_host_config): chunk size in MiB of
        samples and the number of chunks in flight (2..8)."""
        if self._composed:
            return self
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().mfcc_b200_host_config(
                self._h.ptr, 0 if chunk_mb is None else int(chunk_mb * (1 << 20)), 0 if slots is None else int(slots)))
        return self

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
        if self._composed:
            res = self(x.to(self.device, non_blocking=True), n_frames=K)
            out.copy_(res.reshape(B, K, nc))
            return out[0] if single else out
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().mfcc_b200_run_host(
                self._h.ptr, C.c_void_p(x.data_ptr()), N, N, B, K, C.c_void_p(out.data_ptr()),
                K * nc))
        return out[0] if single else out

    # ------------------------------------------------------------------ streaming
    def reset(self, n_streams: int = 0):
        """The ``reset`` signal (mfcc.py:28, :116): every stream back to power-on
        state.  With n_streams > 0 allocates the streaming state."""
        if self._composed:
            if n_streams:
                self._state = torch.zeros((n_streams, self.nfft - self.cfg.hop + 1), dtype=torch.int16,
                                          device=self.device)
            elif self._state is not None:
                self._state.zero_()
            return self
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
        if self._composed:
            # state row = the pre-emphasis predecessor and the last nfft - hop samples
            raw = torch.cat([state, hop.to(self.device)], dim=1)             # [S, nfft + 1]
            frames = self.preemph(raw.reshape(-1)).reshape(S, self.nfft + 1)[:, 1:].contiguous()
            state.copy_(raw[:, self.cfg.hop:])
            res, _t = self._chain(frames)
            out.copy_(res)
            return out
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().mfcc_b200_step(
                self._h.ptr, C.c_void_p(state.data_ptr()), C.c_void_p(hop.data_ptr()), S,
                C.c_void_p(out.data_ptr()), _stream_ptr(self.device)))
        return out
