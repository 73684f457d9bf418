"""TEST INFRASTRUCTURE -- float64 restatement of the librosa call the reference makes:

    librosa.feature.mfcc(samples, sr=16000, hop_length=170, n_mfcc=32)
                                                    (/root/reference/software/genlibrosa.py:21-24)

librosa itself is a third-party dependency that is NOT in /root/reference and not
installed here (the script passes `y` positionally, which librosa accepts up to 0.9.x, so
0.9 defaults are restated: n_fft 2048, hann window, center=True with reflect padding,
power 2, 128 Slaney mel filters 0..sr/2 with Slaney area normalisation, power_to_db with
ref 1, amin 1e-10, top_db 80, orthonormal DCT-II).  Parity against librosa is therefore
UNPINNED on this box; tests/test_librosa_mode.py cross-checks this restatement against
torchaudio's independent implementation of the same definition.  Only tests/ import this.
"""
from __future__ import annotations

import numpy as np


def hz_to_mel(f):
    """Slaney scale (librosa.hz_to_mel, htk=False): linear below 1 kHz, log above."""
    f = np.asanyarray(f, dtype=np.float64)
    f_sp = 200.0 / 3
    mels = f / f_sp
    min_log_hz, min_log_mel, logstep = 1000.0, 1000.0 / f_sp, np.log(6.4) / 27.0
    return np.where(f >= min_log_hz, min_log_mel + np.log(np.maximum(f, 1e-300) / min_log_hz) / logstep, mels)


def mel_to_hz(m):
    m = np.asanyarray(m, dtype=np.float64)
    f_sp = 200.0 / 3
    min_log_hz, min_log_mel, logstep = 1000.0, 1000.0 / f_sp, np.log(6.4) / 27.0
    return np.where(m >= min_log_mel, min_log_hz * np.exp(logstep * (m - min_log_mel)), f_sp * m)


def mel_filters(sr=16000, n_fft=2048, n_mels=128, fmin=0.0, fmax=None):
    """librosa.filters.mel(..., htk=False, norm='slaney'): [n_mels][1 + n_fft/2]."""
    fmax = sr / 2.0 if fmax is None else fmax
    fftfreqs = np.linspace(0, sr / 2.0, 1 + n_fft // 2)
    mel_f = mel_to_hz(np.linspace(hz_to_mel(fmin), hz_to_mel(fmax), n_mels + 2))
    fdiff = np.diff(mel_f)
    ramps = mel_f[:, None] - fftfreqs[None, :]
    w = np.zeros((n_mels, 1 + n_fft // 2))
    for i in range(n_mels):
        lower = -ramps[i] / fdiff[i]
        upper = ramps[i + 2] / fdiff[i + 1]
        w[i] = np.maximum(0, np.minimum(lower, upper))
    enorm = 2.0 / (mel_f[2:n_mels + 2] - mel_f[:n_mels])
    return w * enorm[:, None]


def hann(n):
    """scipy.signal.get_window('hann', n, fftbins=True)."""
    return 0.5 - 0.5 * np.cos(2 * np.pi * np.arange(n) / n)


def dct_matrix(n_out, n_in):
    """Rows 0..n_out-1 of the orthonormal DCT-II of length n_in."""
    k = np.arange(n_out)[:, None]
    m = np.arange(n_in)[None, :]
    d = np.cos(np.pi * (m + 0.5) * k / n_in) * np.sqrt(2.0 / n_in)
    d[0] *= np.sqrt(0.5)
    return d


def num_frames(n, hop=170):
    return 1 + n // hop


def mfcc(x, sr=16000, hop=170, n_mfcc=32, n_fft=2048, n_mels=128, amin=1e-10, top_db=80.0, taps=False,
         center=True):
    """x: int16 samples.  Returns float64 [frames][n_mfcc] (librosa returns the transpose).
    center=False (librosa's stft option): frames start at t*hop, no padding."""
    y = np.asarray(x, dtype=np.float64) / 32768.0               # librosa.load of a 16-bit file
    if center:
        yp = np.pad(y, n_fft // 2, mode="reflect")
        K = num_frames(len(y), hop)
    else:
        yp = y
        K = 1 + (len(y) - n_fft) // hop if len(y) >= n_fft else 0
    idx = hop * np.arange(K)[:, None] + np.arange(n_fft)[None, :]
    frames = yp[idx] * hann(n_fft)
    S = np.abs(np.fft.rfft(frames, axis=1)) ** 2
    mel = S @ mel_filters(sr, n_fft, n_mels).T
    db = 10.0 * np.log10(np.maximum(amin, mel))
    dbc = np.maximum(db, db.max() - top_db)
    out = dbc @ dct_matrix(n_mfcc, n_mels).T
    if taps:
        return out, dict(power=S, mel=mel, db=db, dbc=dbc)
    return out
