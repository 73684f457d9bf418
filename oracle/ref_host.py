"""TEST INFRASTRUCTURE -- CPU restatement of the reference HOST tools around the path:
the wav2mfcc round protocol (software/main.c:100-177), the lifter (software/lift.py:12-26),
the standardisation of software/genlibrosa.py:25-28, the UART column framing
(mfcc/misc/magic.py:27-39, mfcc/targets/mic2mfcc.py:53-74) and the voice-activity gate
(software/cepstrum.c:13, :161-183).  Only tests/ may import it."""
from __future__ import annotations

import numpy as np

NFFT, STEPSIZE = 512, 170


def host_rounds(samples):
    """The sample stream ``mfcc_convert`` sends and the number of rounds (= rows written):
    NFFT samples in round 0, STEPSIZE in every later one, the buffer zeroed first
    (main.c:128-133), a short read sets eof but the round is still sent (:135-144) and
    answered (:156-165); the loop ends after the round in which eof was set (:123)."""
    samples = np.asarray(samples, dtype=np.int16)
    pos, sent, rounds, eof, index = 0, [], 0, False, 0
    while not eof:
        amount = NFFT if index == 0 else STEPSIZE
        index += 1
        buf = np.zeros(amount, dtype=np.int16)
        for i in range(amount):
            if pos < len(samples):
                buf[i] = samples[pos]
                pos += 1
            else:
                eof = True
        sent.append(buf)
        rounds += 1
    return np.concatenate(sent), rounds


def wav2mfcc(samples, run):
    """Rows the tool writes for one file.  `run(stream, n_frames)` is the integer chain
    (an oracle): the FPGA answers round r with frame r of the stream sent so far."""
    stream, rounds = host_rounds(samples)
    return run(stream, rounds)


def lifter(cepstra, L=22):
    """software/lift.py:12-26: ``(1 + (L/2) sin(pi n / L)) * cepstra`` in float64; the
    script then casts with ``astype(np.int16)`` (:39)."""
    cepstra = np.asarray(cepstra)
    if L > 0:
        n = np.arange(cepstra.shape[1])
        lift = 1 + (L / 2.) * np.sin(np.pi * n / L)
        return lift * cepstra
    return cepstra


def lifter_coeffs(ncoeff, L=22):
    n = np.arange(ncoeff)
    return 1 + (L / 2.) * np.sin(np.pi * n / L)


def to_int16(a):
    """What ``ndarray.astype(np.int16)`` does to float64 on this platform: truncate toward
    zero, keep the low 16 bits."""
    t = np.trunc(np.asarray(a, dtype=np.float64)).astype(np.int64)
    return ((t + 32768) % 65536 - 32768).astype(np.int16)


def scale_time(x):
    """``sklearn.preprocessing.scale(spec, axis=1)`` on ``[ncoeff][frames]``
    (genlibrosa.py:25) restated for the ``[frames][ncoeff]`` layout of this repo: per
    coefficient, subtract the mean over time and divide by the population standard
    deviation (a zero deviation divides by 1)."""
    x = np.asarray(x, dtype=np.float64)
    mean = x.mean(axis=0)
    std = x.std(axis=0)
    std = np.where(std < 10 * np.finfo(np.float64).eps, 1.0, std)
    return (x - mean) / std


def uart_frame(cols, magic=0xa55a):
    """Bytes on the serial line for columns ``[n][ncoeff]`` int16: per column the magic
    word then the coefficients, every 16-bit word high byte first (magic.py:27-39 puts
    the magic in front of each `last`-delimited packet, mic2mfcc.py:58-62 sends
    ``tail.source.data`` high byte then low byte)."""
    cols = np.asarray(cols, dtype=np.int16)
    n, nc = cols.shape
    words = np.empty((n, nc + 1), dtype=">u2")
    words[:, 0] = magic
    words[:, 1:] = cols.view(np.uint16)
    return words.tobytes()


def vad_power(window_cols, ncepstrums=16):
    """``cepstrum_eval_power`` (cepstrum.c:161-183) on a full ring given oldest column
    first: sum of c0^2 over elements size/3 <= i < 2*size/3 in steps of ncepstrums, in a C
    ``int`` (wraps modulo 2^32); active when >= POWER_THRESHOLD = 1e8 (:13)."""
    w = np.asarray(window_cols, dtype=np.int64)
    size = w.size
    flat = w.reshape(-1)
    power = 0
    for i in range(size // 3, 2 * size // 3, ncepstrums):
        power += int(flat[i]) * int(flat[i])
    power = (power + 2 ** 31) % 2 ** 32 - 2 ** 31
    return power, power >= 100000000
