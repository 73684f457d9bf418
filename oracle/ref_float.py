"""TEST INFRASTRUCTURE -- float64 CPU oracle for the fp32 mode of the MFCC path.

Restates the float prototype of the reference, notebook/MFCC-INT.ipynb cells 2-10
(citations give the cell and the line inside the .ipynb JSON file), which runs
numpy/scipy float math on un-normalised int16 audio.  Only tests/, smoke() and
bench.py's CPU leg may import it; the product never does.

Pinning: tests/golden/make_float_golden.py EXECUTES the notebook's own code cells on
the bundled clip and stores their outputs (tests/golden/notebook_float.npz);
tests/test_float_oracle.py requires this restatement to reproduce them to 1e-9.
"""
from __future__ import annotations

import numpy as np
from scipy.signal import get_window


def preemph(x, coeff=0.96875):
    """cell 2 (:85-86): y[0] = x[0], y[n] = x[n] - 0.96875 x[n-1]."""
    x = np.asarray(x, dtype=np.float64)
    if x.size == 0:
        return x
    return np.append(x[0], x[1:] - coeff * x[:-1])


def num_frames(n, nfft=512, hop=170):
    """cell 3 (:138): int((len - FFT_size) / hop_len) + 1."""
    return 0 if n < nfft else int((n - nfft) / hop) + 1


def frame(y, nfft=512, hop=170, n_frames=None):
    """cell 3 (:137-142).  n_frames beyond the full frames zero-extends the stream
    (what the reference hosts feed, software/main.c:135-144)."""
    y = np.asarray(y, dtype=np.float64)
    K = num_frames(len(y), nfft, hop) if n_frames is None else n_frames
    need = hop * (K - 1) + nfft if K > 0 else 0
    if need > len(y):
        y = np.concatenate([y, np.zeros(need - len(y))])
    out = np.zeros((K, nfft))
    for n in range(K):
        out[n] = y[n * hop:n * hop + nfft]
    return out


def window(nfft=512):
    """cell 4 (:227): periodic Hamming."""
    return get_window("hamm", nfft, fftbins=True)


def spectrum(frames, nfft=512):
    """cell 5 (:262-268): fft / 512, bins 0..nfft/2 (257 of them)."""
    return np.fft.fft(frames, axis=1)[:, :nfft // 2 + 1] / 512


def freq_to_mel(freq):
    return 2595.0 * np.log10(1.0 + freq / 700.0)             # cell 7 (:291)


def mel_to_freq(mels):
    return 700.0 * (10.0 ** (mels / 2595.0) - 1.0)            # cell 7 (:294)


def filter_points(nfilters=32, nfft=512, samplerate=16000):
    """cell 7 (:296-303), called with fmin=0, fmax=samplerate/2 (cell 8)."""
    mels = np.linspace(freq_to_mel(0), freq_to_mel(samplerate / 2), num=nfilters + 2)
    return np.floor((nfft + 1) / samplerate * mel_to_freq(mels)).astype(int)


def filters(nfilters=32, nfft=512, samplerate=16000):
    """cell 7 (:305-312): dense [nfilters][nfft/2+1] linspace triangles."""
    p = filter_points(nfilters, nfft, samplerate)
    f = np.zeros((len(p) - 2, nfft // 2 + 1))
    for n in range(len(p) - 2):
        f[n, p[n]:p[n + 1]] = np.linspace(0, 1, p[n + 1] - p[n])
        f[n, p[n + 1]:p[n + 2]] = np.linspace(1, 0, p[n + 2] - p[n + 1])
    return f


def dct_matrix(n):
    """Orthonormal DCT-II, what scipy.fft.dct(type=2, norm="ortho") applies (cell 10, :464)."""
    k = np.arange(n)[:, None]
    m = np.arange(n)[None, :]
    d = np.cos(np.pi * (m + 0.5) * k / n) * np.sqrt(2.0 / n)
    d[0] *= np.sqrt(0.5)
    return d


def mfcc(x, nfft=512, samplerate=16000, nfilters=32, nceptrums=32, n_frames=None, mel_floor=0.0,
         taps=False):
    """Cells 2-10 in order.  x: int16 samples (not normalised).  Returns float64
    [K, nceptrums].  mel_floor > 0 clamps the mel energies before log2 (the notebook has
    no floor: silence gives -inf)."""
    y = preemph(x)
    if n_frames is not None:
        need = (nfft // 3) * (n_frames - 1) + nfft if n_frames > 0 else 0
        if need > len(y):
            # zero samples appended to the INPUT stream: the first one still carries
            # the pre-emphasis tail of the last real sample
            xe = np.concatenate([np.asarray(x, dtype=np.float64), np.zeros(need - len(y))])
            y = preemph(xe)
    fr = frame(y, nfft, nfft // 3, n_frames)
    win = fr * window(nfft)
    spec = spectrum(win, nfft)
    power = np.square(np.abs(spec))                            # cell 6 (:268)
    mel = power @ filters(nfilters, nfft, samplerate).T        # cell 8 (:376), transposed
    if mel_floor > 0:
        mel = np.maximum(mel, mel_floor)
    with np.errstate(divide="ignore"):
        lg = np.log2(mel)                                      # cell 9 (:415)
    dc = lg @ dct_matrix(nfilters).T                           # cell 10 (:464)
    out = dc[:, :nceptrums]
    if taps:
        return out, dict(preemph=y, frame=fr, window=win, spec=spec, power=power, mel=mel, log=lg,
                         dct=dc)
    return out
