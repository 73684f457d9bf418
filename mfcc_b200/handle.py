"""Opaque library handle: configuration + ROM tables uploaded to the GPU once."""
from __future__ import annotations

import ctypes as C
import math

import numpy as np

from . import _lib, tables
from .config import MFCCConfig, MODE_FP32


class Handle:
    """Owns one ``mfcc_b200_create`` handle for the current CUDA device."""

    def __init__(self, cfg: MFCCConfig):
        self.cfg = cfg
        L = _lib.lib()
        if cfg.nfft != 512 or cfg.nfilters not in (16, 32):
            # the same parameter sets MFCC can generate in the reference tops
            raise ValueError("CUDA kernels cover nfft=512 and nfilters in {16, 32}")
        pts = tables.mel_points(cfg.nfilters, cfg.nfft, cfg.samplerate)
        w, _seg, close = tables.mel_weights(cfg.nfilters, cfg.nfft, cfg.samplerate,
                                            cfg.pow_width_out)
        # maxvalrange - (gain + width_output), filterbank.py:71, :122
        maxvalrange = int(math.log2(pts[-1] - pts[-3])) + 2 * cfg.pow_width_out
        self._keep = dict(
            curve=np.ascontiguousarray(tables.window_curve(cfg.nfft, cfg.window_precision),
                                       dtype=np.int32),
            tw_fft=np.ascontiguousarray(tables.twiddle_table(cfg.nfft), dtype=np.int32),
            tw_dct=np.ascontiguousarray(tables.twiddle_table(cfg.dct_size), dtype=np.int32),
            mel_w=np.ascontiguousarray(w, dtype=np.uint32),
            mel_close=np.ascontiguousarray(close, dtype=np.uint8),
        )
        k = self._keep
        c = _lib.Cfg()
        c.struct_size = C.sizeof(_lib.Cfg)
        c.mode = cfg.mode
        c.nfft, c.hop, c.nfilters, c.nceptrums = cfg.nfft, cfg.hop, cfg.nfilters, cfg.nceptrums
        c.window_shift = cfg.window_precision + 1
        c.pow_shift = 2 * cfg.width - cfg.pow_width_out
        c.mel_wbits = cfg.pow_width_out
        c.mel_lo_bit = maxvalrange - (cfg.mel_gain + cfg.mel_width_out)
        c.mel_out_bits = cfg.mel_width_out
        c.log_precision = cfg.log_precision
        c.log_out_bits = cfg.log_width_out
        c.window_curve = k["curve"].ctypes.data
        c.tw_fft = k["tw_fft"].ctypes.data
        c.tw_dct = k["tw_dct"].ctypes.data
        c.mel_weight = k["mel_w"].ctypes.data
        c.mel_close = k["mel_close"].ctypes.data
        c.f_preemph = 0.96875
        if cfg.mode == MODE_FP32:
            k["f_window"] = np.ascontiguousarray(tables.f_window(cfg.nfft), dtype=np.float32)
            k["f_mel"] = np.ascontiguousarray(
                tables.f_mel_filters(cfg.nfilters, cfg.nfft, cfg.samplerate), dtype=np.float32)
            k["f_dct"] = np.ascontiguousarray(tables.f_dct_matrix(cfg.nfilters), dtype=np.float32)
            c.f_window = k["f_window"].ctypes.data
            c.f_mel = k["f_mel"].ctypes.data
            c.f_dct = k["f_dct"].ctypes.data
            c.f_mel_floor = float(cfg.mel_floor)
        self._c = c
        h = C.c_void_p()
        _lib.check(L.mfcc_b200_create(C.byref(c), C.byref(h)))
        self.ptr = h

    def close(self):
        if getattr(self, "ptr", None):
            _lib.lib().mfcc_b200_destroy(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
