"""``LibrosaMFCC``: the librosa feature set of the reference's comparison script on the GPU.

/root/reference/software/genlibrosa.py:21-28 computes, per clip,

    spec = librosa.feature.mfcc(samples, sr=16000, hop_length=170, n_mfcc=32)
    spec = sklearn.preprocessing.scale(spec, axis=1).astype(np.int16)

to compare the FPGA's cepstra with.  ``LibrosaMFCC`` takes the same keyword names
(``sr, n_mfcc, hop_length, n_fft, n_mels``) and computes the first line for a batch of
int16 clips; ``mfcc_b200.post.scale`` is the second.  Results are float32 and TIME-MAJOR
``[K, n_mfcc]`` like every other operator of this package (librosa returns ``[n_mfcc, K]``).
There is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib, tables
from .mfcc import _check_samples, _stream_ptr


class LibrosaMFCC:
    def __init__(self, sr=16000, n_mfcc=32, hop_length=170, n_fft=2048, n_mels=128, center=True,
                 amin=1e-10, top_db=80.0, device=None):
        if n_fft != 2048:
            raise ValueError("CUDA kernels cover n_fft=2048 (librosa's default)")
        if not torch.cuda.is_available():
            raise RuntimeError("LibrosaMFCC needs a CUDA device: there is no CPU fallback")
        self.sr, self.n_mfcc, self.hop_length, self.n_fft, self.n_mels = sr, n_mfcc, hop_length, n_fft, n_mels
        self.center, self.amin, self.top_db = bool(center), float(amin), top_db
        self.device = torch.device(device if device is not None else
                                   "cuda:%d" % torch.cuda.current_device())
        self._keep = dict(
            # int16 -> [-1, 1) like librosa.load of a 16-bit file, folded into the window
            window=np.ascontiguousarray(tables.lr_window(n_fft) / 32768.0, dtype=np.float32),
            mel=np.ascontiguousarray(tables.lr_mel_filters(sr, n_fft, n_mels), dtype=np.float32),
            dct=np.ascontiguousarray(tables.lr_dct_matrix(n_mfcc, n_mels), dtype=np.float32))
        c = _lib.LibrosaCfg()
        c.struct_size = C.sizeof(_lib.LibrosaCfg)
        c.n_fft, c.hop, c.n_mels, c.n_mfcc, c.center = n_fft, hop_length, n_mels, n_mfcc, int(self.center)
        c.amin = self.amin
        c.top_db = -1.0 if top_db is None else float(top_db)
        c.window = self._keep["window"].ctypes.data
        c.mel = self._keep["mel"].ctypes.data
        c.dct = self._keep["dct"].ctypes.data
        h = C.c_void_p()
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().mfcc_b200_librosa_create(C.byref(c), C.byref(h)))
        self._h = h
        self._work = None

    def __del__(self):
        try:
            h = getattr(self, "_h", None)
            if h:
                _lib.lib().mfcc_b200_librosa_destroy(h)
                self._h = None
        except Exception:       # interpreter shutdown: module globals may be gone
            pass

    def config(self, chunk_mb):
        """Size of the log-mel rows kept between the two passes (mfcc_b200_librosa_config)."""
        _lib.check(_lib.lib().mfcc_b200_librosa_config(self._h, int(chunk_mb * (1 << 20))))
        return self

    def dct_path(self, tensor_cores=True):
        """Default shape only: DCT product on the tensor cores (default) or on the FMA pipes."""
        _lib.check(_lib.lib().mfcc_b200_librosa_dct_path(self._h, 1 if tensor_cores else 0))
        return self

    def num_frames(self, n_samples: int) -> int:
        """librosa's frame count: ``1 + n // hop_length`` for centred frames."""
        return int(_lib.lib().mfcc_b200_librosa_frames(self._h, n_samples))

    def __call__(self, samples, taps=False, out=None):
        """samples: int16 ``[N]`` or ``[B, N]`` on the GPU -> float32 ``[K, n_mfcc]`` /
        ``[B, K, n_mfcc]``.  taps=True (one clip) also returns dict(power, logmel)."""
        _check_samples(samples, self.device)
        single = samples.dim() == 1
        x = samples.reshape(1, -1) if single else samples
        if x.stride(1) != 1:
            x = x.contiguous()
        B, N = x.shape
        if N < 1:
            raise ValueError("a clip needs at least one sample")
        K = self.num_frames(N)
        L = _lib.lib()
        need = C.c_int64()
        _lib.check(L.mfcc_b200_librosa_workspace(self._h, B, K, C.byref(need)))
        if self._work is None or self._work.numel() < need.value:
            self._work = torch.empty(need.value, dtype=torch.uint8, device=self.device)
        if out is None:
            out = torch.empty((B, K, self.n_mfcc), dtype=torch.float32, device=self.device)
        elif out.shape != (B, K, self.n_mfcc) or out.dtype != torch.float32 or not out.is_contiguous():
            raise ValueError("out must be contiguous float32 [%d, %d, %d]" % (B, K, self.n_mfcc))
        t, tp = None, None
        if taps:
            if B != 1:
                raise ValueError("taps need a single clip")
            t = dict(power=torch.empty((K, self.n_fft // 2 + 1), dtype=torch.float32, device=self.device),
                     logmel=torch.empty((K, self.n_mels), dtype=torch.float32, device=self.device))
            tp = _lib.LibrosaTaps(t["power"].data_ptr(), t["logmel"].data_ptr())
        with torch.cuda.device(self.device):
            _lib.check(L.mfcc_b200_librosa_run(
                self._h, C.c_void_p(x.data_ptr()), x.stride(0), N, B, K,
                C.c_void_p(self._work.data_ptr()), C.c_void_p(out.data_ptr()), out.stride(0),
                C.byref(tp) if tp is not None else None, _stream_ptr(self.device)))
        res = out[0] if single else out
        return (res, t) if taps else res
