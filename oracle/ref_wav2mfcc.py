"""TEST INFRASTRUCTURE -- the reference's own wav2mfcc host tool, compiled and run.

oracle/_ref/libwav2mfcc_ref.so is /root/reference/software/main.c + its bundled libwav,
unmodified, with the USB layer replaced by oracle/ft601_bridge.c (`make -C oracle ref`).
`FpgaModel` answers the tool's ft601_write / ft601_read calls the way the FPGA top does
(mfcc/targets/wav2mfcc.py:27-44): a word with bit 31 set resets the core, any other word
pushes its low 16 bits into MFCC.sink, and every completed frame comes back as one 32-bit
word per cepstral coefficient.  The arithmetic behind the model is whatever `compute` is
(the C oracle in the CPU tests).  Only tests/ import this.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_ref", "libwav2mfcc_ref.so")
REF_SRC = "/root/reference/software/main.c"
_IO = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int)


def available() -> bool:
    return os.path.exists(_SO) or os.path.exists(REF_SRC)


def build() -> str:
    if os.path.exists(REF_SRC):
        subprocess.check_call(["make", "-C", _HERE, "-s", "ref"])
    if not os.path.exists(_SO):
        raise FileNotFoundError(_SO)
    return _SO


class FpgaModel:
    """compute(samples int16 [n], n_frames) -> int16 [n_frames, ncepstrums] for a stream
    that starts at reset."""

    def __init__(self, compute, ncepstrums=32, nfft=512, hop=170):
        self.compute, self.nc, self.nfft, self.hop = compute, ncepstrums, nfft, hop
        self.resets = 0
        self.words_in = 0
        self._reset()

    def _reset(self):
        self.samples = []
        self.sent = 0                       # frames already returned

    def write(self, data, size):
        words = np.frombuffer(C.string_at(data, size), dtype="<u4")
        self.words_in += len(words)
        for w in words:
            if w & 0x80000000:              # wav2mfcc.py:27: reset = valid & data[-1]
                self._reset()
                self.resets += 1
            else:
                self.samples.append(np.int16(np.uint16(w & 0xffff)))
        return 0

    def read(self, data, size):
        n = size // 4
        assert n == self.nc, "the tool reads one frame of cepstra per round"
        k = self.sent
        need = self.hop * k + self.nfft
        if len(self.samples) < need:
            return -200                     # the hardware would block: FT601_ERROR_OTHER
        x = np.asarray(self.samples[:need], dtype=np.int16)
        row = np.asarray(self.compute(x, k + 1))[k]
        out = (row.astype(np.int16).view(np.uint16).astype("<u4"))
        C.memmove(data, out.ctypes.data, 4 * n)
        self.sent += 1
        return 0


def run_tool(wavdir: str, model: FpgaModel) -> int:
    """`./mfcc <wavdir>` (software/main.c:249-273) against the model.  Returns main()'s value."""
    L = C.CDLL(build())
    w, r = _IO(model.write), _IO(model.read)
    L.ft601_bridge_set.argtypes = [_IO, _IO]
    L.ft601_bridge_set(w, r)
    L.wav2mfcc_main.restype = C.c_int
    L.wav2mfcc_main.argtypes = [C.c_int, C.POINTER(C.c_char_p)]
    argv = (C.c_char_p * 3)(b"mfcc", os.fsencode(wavdir), None)
    rc = L.wav2mfcc_main(2, argv)
    L.ft601_bridge_set(_IO(0), _IO(0))
    return int(rc)
