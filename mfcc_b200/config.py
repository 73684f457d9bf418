"""Configuration of the MFCC path.

Mirrors the constructor arguments of the reference top level
``MFCC(width, nfft, samplerate, nfilters, nceptrums)``
(/root/reference/mfcc/core/mfcc.py:19-26) and the fixed values it hands to its
stages (/root/reference/mfcc/core/mfcc.py:38-88).
"""
from __future__ import annotations

import math
from dataclasses import dataclass


MODE_INT = 0     # bit-exact emulation of the FPGA fixed-point pipeline
MODE_FP32 = 1    # float32 pipeline following notebook/MFCC-INT.ipynb


@dataclass(frozen=True)
class MFCCConfig:
    # names and defaults of MFCC.__init__, mfcc/core/mfcc.py:20-21
    width: int = 16
    nfft: int = 512
    samplerate: float = 16e3
    nfilters: int = 16
    nceptrums: int = 16
    # values MFCC.elaborate passes down, mfcc/core/mfcc.py:38-88
    window_precision: int = 8       # WindowHamming(precision=8)        :47-49
    pow_width_out: int = 30         # PowerSpectrum(width_output=30)    :60-62
    mel_width_out: int = 16         # FilterBank(width_output=16, ...)  :69-75
    mel_gain: int = 18              # FilterBank(gain=18)
    log_width_out: int = 15         # Log2Fix(filterbank.width_output, 15)  :82
    mode: int = MODE_INT
    mel_floor: float = 0.0          # fp32 mode: mel energies are clamped to this before log2 (0 = off)

    def __post_init__(self):
        if self.width != 16:
            raise ValueError("only width=16 is supported (the reference tops use 16)")
        if self.nfft < 8 or self.nfft & (self.nfft - 1):
            # window.py:13 asserts nfft >= 8; fft.py:351-353 wants a power of two
            raise ValueError("nfft must be a power of two >= 8, not %r" % (self.nfft,))
        if self.nfilters < 2 or self.nfilters & (self.nfilters - 1):
            # DCTStream builds FFT(size=4*nfilters): fft.py:351-353
            raise ValueError("nfilters must be a power of two >= 2, not %r" % (self.nfilters,))
        if not (1 <= self.nceptrums <= self.nfilters):
            raise ValueError("nceptrums must be in 1..nfilters")
        if self.mode not in (MODE_INT, MODE_FP32):
            raise ValueError("mode must be MODE_INT or MODE_FP32")
        if not self.mel_floor >= 0.0:
            raise ValueError("mel_floor must be >= 0")

    @property
    def hop(self) -> int:
        """Frame step, ``stepsize=self.nfft//3`` (mfcc/core/mfcc.py:43)."""
        return self.nfft // 3

    @property
    def nbins(self) -> int:
        """Bins emitted by FftStream: 0 .. nfft/2-1 (mfcc/core/fft_stream.py:24,28)."""
        return self.nfft // 2

    @property
    def log_precision(self) -> int:
        """Fraction bits of Log2Fix (mfcc/core/log.py:114)."""
        return self.log_width_out - math.ceil(math.log2(self.mel_width_out))

    @property
    def dct_size(self) -> int:
        """DCTStream runs FFT(size=nfft*4) with nfft=nfilters (dct_stream.py:18)."""
        return 4 * self.nfilters

    def num_frames(self, n_samples: int) -> int:
        """Full frames in a clip of n_samples: the core emits frame k once samples
        hop*k .. hop*k+nfft-1 have arrived (mfcc/core/frame.py:117-148)."""
        if n_samples < self.nfft:
            return 0
        return (n_samples - self.nfft) // self.hop + 1

    def num_frames_host(self, n_samples: int) -> int:
        """Frames the reference host tool writes per file: it sends nfft samples,
        then hop per round, zero-filled at EOF, one output row per round
        (software/main.c:95, :128-166)."""
        if n_samples < self.nfft:
            return 1
        return (n_samples - self.nfft) // self.hop + 2
