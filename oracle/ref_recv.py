"""TEST INFRASTRUCTURE -- the reference's own host receiver, compiled and called.

/root/reference/software/cepstrum.c and serial.c (the C side of the microphone build:
column receiver with 0xa55a re-alignment, 93-column sliding window, voice-activity gate)
are compiled unmodified into oracle/_ref/libcepstrum_ref.so by `make -C oracle ref`
(only where /root/reference exists; the built file travels to the GPU box).  Bytes are
fed to it through a pipe in place of the serial port.  Only tests/ import this.
"""
from __future__ import annotations

import ctypes as C
import fcntl
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_ref", "libcepstrum_ref.so")
REF_SRC = "/root/reference/software/cepstrum.c"


class Circular(C.Structure):
    """struct circular_s, software/cepstrum.h:12-18."""
    _fields_ = [("array", C.POINTER(C.c_int16)), ("head", C.POINTER(C.c_int16)),
                ("size", C.c_size_t), ("count", C.c_size_t)]


def available() -> bool:
    return os.path.exists(_SO) or os.path.exists(REF_SRC)


def build() -> str:
    if os.path.exists(REF_SRC):
        subprocess.check_call(["make", "-C", _HERE, "-s", "ref"])
    if not os.path.exists(_SO):
        raise FileNotFoundError(_SO)
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(build())
        L.cepstrum_get_window.restype = C.c_int
        L.cepstrum_get_window.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int]
        L.cepstrum_refill_window.restype = C.c_int
        L.cepstrum_refill_window.argtypes = [C.c_int, C.POINTER(Circular), C.c_int, C.c_int, C.c_void_p]
        L.cepstrum_eval_power.restype = C.c_bool
        L.cepstrum_eval_power.argtypes = [C.POINTER(Circular), C.c_int]
        _lib = L
    return _lib


class Port:
    """A pipe standing in for /dev/ttyUSBx: write() what the board would send."""

    def __init__(self):
        self.r, self.w = os.pipe()
        fcntl.fcntl(self.w, 1031, 1 << 20)                  # F_SETPIPE_SZ

    def write(self, data: bytes):
        os.write(self.w, bytes(data))

    def close(self):
        for fd in (self.r, self.w):
            try:
                os.close(fd)
            except OSError:
                pass


def get_window(data: bytes, ncepstrums: int, nframes: int) -> np.ndarray:
    """cepstrum_get_window (software/cepstrum.c:68-90) on a byte string."""
    port = Port()
    try:
        port.write(data)
        buf = np.zeros((nframes, ncepstrums), dtype=np.int16)
        rc = lib().cepstrum_get_window(port.r, buf.ctypes.data, ncepstrums, nframes)
        if rc != 0:
            raise RuntimeError("cepstrum_get_window: %d" % rc)
        return buf
    finally:
        port.close()


class Receiver:
    """The main loop of the microphone host (software/cepstrum.c:92-183 as driven by the
    viewer): a circular window of `nframes` columns, refilled from the port, then
    cepstrum_eval_power."""

    def __init__(self, ncepstrums=16, nframes=93):
        self.nc, self.nf = ncepstrums, nframes
        self.store = np.zeros(nframes * ncepstrums, dtype=np.int16)
        self.circ = Circular()
        self.circ.array = self.store.ctypes.data_as(C.POINTER(C.c_int16))
        self.circ.head = self.circ.array
        self.circ.size = nframes * ncepstrums
        self.circ.count = 0
        self.port = Port()

    def push(self, frame_bytes: bytes) -> int:
        """Send bytes, let the receiver take what is available; returns new columns."""
        self.port.write(frame_bytes)
        return int(lib().cepstrum_refill_window(self.port.r, C.byref(self.circ), self.nc, self.nf, None))

    def active(self) -> bool:
        return bool(lib().cepstrum_eval_power(C.byref(self.circ), self.nc))

    def close(self):
        self.port.close()
