"""Constant tables of the MFCC path, built the way the reference builds them when
its hardware description is elaborated.

Nothing here runs per sample; the arrays are uploaded to the GPU once when a
handle is created (see include/mfcc_b200.h, ``mfcc_b200_create``).

Reference:
  window    /root/reference/mfcc/core/window.py:22-43 (table), :59-115 (rebuild)
  twiddles  /root/reference/mfcc/misc/fft.py:28-36 (ROM), :43-59 (decode)
  mel       /root/reference/mfcc/core/filterbank.py:9-34 (points, steps),
            :90-115 (weight generator)
"""
from __future__ import annotations

import numpy as np
from scipy.signal import get_window


# --------------------------------------------------------------------------- window

def window_rom(nfft: int = 512, precision: int = 8):
    """Quarter table of the Hamming window, every second point, first value removed.

    window.py:22-43: ``winfull = trunc(hamm(nfft, periodic) * (2**(precision+1)-1))``;
    ``mem = winfull[:nfft//4][1::2] - mem[0]``; ``off_lst = 2*(winfull[nfft//4]-off_fst)``.
    Returns (mem[nfft//8] int64, off_fst, off_lst).
    """
    maxheight = 2 ** (precision + 1) - 1
    win = get_window("hamm", nfft, fftbins=True)
    winfull = (win * maxheight).astype(int)
    mem = np.array(winfull[: nfft // 4][1::2], dtype=np.int64)
    off_fst = int(mem[0])
    mem -= off_fst
    if mem.max() >= 2 ** precision:
        raise ValueError("window table does not fit in 'precision' bits")  # window.py:37
    off_lst = int(2 * (winfull[nfft // 4] - off_fst))
    return mem, off_fst, off_lst


def window_curve(nfft: int = 512, precision: int = 8) -> np.ndarray:
    """The nfft multipliers the hardware applies, one per sample position.

    Rebuilt from the quarter table with the two symmetries and the even-point
    interpolation of window.py:59-115. The curve is the same for every frame:
    ``point_r`` holds P(nfft-1) = mem[0] = 0 when a frame starts, which is also its
    reset value.
    """
    mem, off_fst, off_lst = window_rom(nfft, precision)
    nbits = nfft.bit_length() - 1
    abits = nbits - 3
    amask = (1 << abits) - 1
    cmask = (1 << (precision + 1)) - 1
    curve = np.zeros(nfft, dtype=np.int64)
    point_r = 0
    for n in range(nfft):
        msb = (n >> (nbits - 1)) & 1
        dirb = (n >> (nbits - 2)) & 1
        addr = (n >> 1) & amask
        if dirb:
            addr = (~addr) & amask
        data = int(mem[addr])
        point = ((off_lst - data) if (msb ^ dirb) else data) & cmask
        if n & 1:
            curve[n] = (off_fst + point) & cmask
            point_r = point
        else:
            curve[n] = (off_fst + ((point + point_r) >> 1)) & cmask
    return curve


# --------------------------------------------------------------------------- twiddles

def twiddle_rom(size: int, width: int = 16):
    """Quarter-circle ROM of TwiddleROM(size, width): size/4 entries of
    ``round(2**(width-2) * exp(-1j*p))``, p = linspace(0, pi/2, size/4, endpoint=False)
    (fft.py:28-36). np.round is round-half-even on each component."""
    p = np.linspace(start=0, stop=np.pi / 2, num=size // 4, endpoint=False)
    z = np.round((1 << (width - 2)) * np.exp(-1j * p))
    return z.real.astype(np.int64), z.imag.astype(np.int64)


def twiddle_table(size: int, width: int = 16) -> np.ndarray:
    """Decoded twiddles W(k), k in [0, size/2), as int32 [size/2, 2] = (re, im).

    fft.py:43-59 (non-inverted branch): k < size/4 reads (re[k], im[k]); otherwise,
    with i = k - size/4, it reads (im[i], -re[i])."""
    re, im = twiddle_rom(size, width)
    q = size // 4
    out = np.zeros((size // 2, 2), dtype=np.int32)
    out[:q, 0] = re
    out[:q, 1] = im
    out[q:, 0] = im
    out[q:, 1] = -re
    return out


# --------------------------------------------------------------------------- mel

def freq_to_mel(freq):
    return 2595.0 * np.log10(1.0 + freq / 700.0)          # filterbank.py:9-10


def mel_to_freq(mels):
    return 700.0 * (10.0 ** (mels / 2595.0) - 1.0)         # filterbank.py:12-13


def mel_points(nfilters: int = 32, nfft: int = 512, samplerate: float = 16e3) -> np.ndarray:
    """nfilters+2 bin indices, filterbank.py:15-20 called as at :62
    (fmin=0, fmax=samplerate/2)."""
    mels = np.linspace(freq_to_mel(0.0), freq_to_mel(samplerate / 2), num=nfilters + 2)
    freqs = mel_to_freq(mels)
    return np.floor((nfft + 1) / samplerate * freqs).astype(int)


def mel_steps(points, wsize: int = 30):
    """Per-segment increments of the weight accumulator, filterbank.py:22-34.
    Python ints (60 bits)."""
    max_acc = 1 << (2 * wsize)
    steps = []
    for i in range(len(points) - 1):
        diff = int(points[i + 1]) - int(points[i]) - 1
        steps.append((max_acc // diff) - 1 if diff else max_acc - 1)
    return steps


def mel_weights(nfilters: int = 32, nfft: int = 512, samplerate: float = 16e3,
                wsize: int = 30):
    """Run the weight generator of FilterBank over one frame (filterbank.py:90-115).

    Returns (w, seg, close):
      w[b]     uint32  multiplier applied to bin b (top wsize bits of the accumulator)
      seg[b]   int32   segment (ROM address) bin b belongs to
      close[b] uint8   1 when bin b ends its segment: the weight reached all-ones
                       ("highest") or b is the last bin of the frame
    The generator is input-independent, so this is a constant of the configuration.
    """
    nbins = nfft // 2
    pts = mel_points(nfilters, nfft, samplerate)
    steps = mel_steps(pts, wsize)
    top = (1 << wsize) - 1
    amask = (1 << (2 * wsize)) - 1
    w = np.zeros(nbins, dtype=np.uint32)
    seg = np.zeros(nbins, dtype=np.int32)
    close = np.zeros(nbins, dtype=np.uint8)
    acc = 0
    adr = 0
    for b in range(nbins):
        wb = acc >> wsize
        highest = wb == top
        last = b == nbins - 1
        w[b] = wb
        seg[b] = adr
        close[b] = 1 if (highest or last) else 0
        if highest or last:
            adr = 0 if last else adr + 1
            acc = 0
        else:
            if adr >= len(steps):
                raise ValueError("filter ROM overrun: configuration not realisable")
            acc = (acc + steps[adr]) & amask
    return w, seg, close


def mel_segment_bounds(nfilters: int = 32, nfft: int = 512, samplerate: float = 16e3,
                       wsize: int = 30) -> np.ndarray:
    """int32 [nseg, 2] = (first bin, closing bin) of every segment of one frame,
    derived from mel_weights. Segment g feeds the rising half of filter g and the
    falling half of filter g-1 (filterbank.py:119-142)."""
    _, seg, close = mel_weights(nfilters, nfft, samplerate, wsize)
    nseg = int(seg.max()) + 1
    out = np.zeros((nseg, 2), dtype=np.int32)
    for g in range(nseg):
        idx = np.nonzero(seg == g)[0]
        out[g, 0] = idx[0]
        out[g, 1] = idx[-1]
        assert close[idx[-1]] == 1 and not close[idx[:-1]].any()
    return out


# ----------------------------------------------------------------------------- fp32 mode
# Tables of the float pipeline of notebook/MFCC-INT.ipynb (cells 4, 7, 10), built in
# float64 and handed to the library as float32.

def f_window(nfft: int = 512) -> np.ndarray:
    """Periodic Hamming, ``get_window("hamm", nfft, fftbins=True)`` (cell 4):
    0.54 - 0.46 cos(2 pi n / nfft)."""
    n = np.arange(nfft)
    return 0.54 - 0.46 * np.cos(2.0 * np.pi * n / nfft)


def f_mel_filters(nfilters: int = 32, nfft: int = 512, samplerate: float = 16e3) -> np.ndarray:
    """Dense [nfilters][nfft/2+1] triangles of ``get_filters`` (cell 7): linspace(0, 1)
    up to the centre point, linspace(1, 0) down, on the same points as the RTL."""
    p = mel_points(nfilters, nfft, samplerate)
    f = np.zeros((nfilters, nfft // 2 + 1))
    for n in range(nfilters):
        f[n, p[n]:p[n + 1]] = np.linspace(0, 1, p[n + 1] - p[n])
        f[n, p[n + 1]:p[n + 2]] = np.linspace(1, 0, p[n + 2] - p[n + 1])
    return f


def f_dct_matrix(n: int = 32) -> np.ndarray:
    """Orthonormal DCT-II matrix D[k][m] (``scipy.fft.dct(type=2, norm="ortho")``, cell 10)."""
    k = np.arange(n)[:, None]
    m = np.arange(n)[None, :]
    d = np.cos(np.pi * (m + 0.5) * k / n) * np.sqrt(2.0 / n)
    d[0] *= np.sqrt(0.5)
    return d


# ----------------------------------------------------------------------------- librosa set
# Tables of librosa.feature.mfcc with the arguments of software/genlibrosa.py:21-24 and
# librosa's defaults for the rest (librosa is not part of /root/reference; these restate
# its published definitions).  Built in float64, handed to the library as float32.

def lr_hz_to_mel(f):
    """Slaney mel scale: linear (200/3 Hz per mel) below 1 kHz, logarithmic above with
    27 mels per factor 6.4."""
    f = np.asarray(f, dtype=np.float64)
    lin = f * (3.0 / 200.0)
    with np.errstate(divide="ignore"):
        log = 15.0 + np.log(f / 1000.0) * (27.0 / np.log(6.4))
    return np.where(f >= 1000.0, log, lin)


def lr_mel_to_hz(m):
    m = np.asarray(m, dtype=np.float64)
    return np.where(m >= 15.0, 1000.0 * np.exp((m - 15.0) * (np.log(6.4) / 27.0)), m * (200.0 / 3.0))


def lr_mel_filters(samplerate: float = 16000, n_fft: int = 2048, n_mels: int = 128,
                   fmin: float = 0.0, fmax=None) -> np.ndarray:
    """Dense [n_mels][n_fft/2+1] triangular filters on Slaney-spaced centres, each scaled
    to constant area (2 / bandwidth)."""
    fmax = samplerate / 2.0 if fmax is None else float(fmax)
    bins = np.arange(n_fft // 2 + 1) * (samplerate / n_fft)
    edges = lr_mel_to_hz(np.linspace(lr_hz_to_mel(fmin), lr_hz_to_mel(fmax), n_mels + 2))
    out = np.zeros((n_mels, bins.size))
    for i in range(n_mels):
        lo, mid, hi = edges[i], edges[i + 1], edges[i + 2]
        rise = (bins - lo) / (mid - lo)
        fall = (hi - bins) / (hi - mid)
        out[i] = np.clip(np.minimum(rise, fall), 0.0, None) * (2.0 / (hi - lo))
    return out


def lr_window(n_fft: int = 2048) -> np.ndarray:
    """Periodic Hann, 0.5 - 0.5 cos(2 pi n / n_fft)."""
    return 0.5 - 0.5 * np.cos(2.0 * np.pi * np.arange(n_fft) / n_fft)


def lr_dct_matrix(n_mfcc: int = 32, n_mels: int = 128) -> np.ndarray:
    """First n_mfcc rows of the orthonormal DCT-II of length n_mels."""
    k = np.arange(n_mfcc)[:, None]
    m = np.arange(n_mels)[None, :]
    d = np.cos(np.pi * (m + 0.5) * k / n_mels) * np.sqrt(2.0 / n_mels)
    d[0] *= np.sqrt(0.5)
    return d
