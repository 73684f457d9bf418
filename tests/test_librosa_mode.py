"""The librosa parameter set of software/genlibrosa.py:21-24.

librosa is not in /root/reference and not installed here, so oracle/ref_librosa.py is a
restatement of its published definitions and parity with librosa itself is UNPINNED.
What these tests do hold it to:
  * an independent implementation of the same definition (torchaudio's MFCC transform
    configured for Slaney filters, reflect padding, power_to_db/top_db, ortho DCT),
  * closed-form anchors of the mel scale, the filters and the DCT,
  * the CUDA float32 path against the float64 restatement, tolerance below.

Stated tolerance of the CUDA path: per coefficient |got - ref| <= 5e-3 + 1e-4 |ref|
(values are decibel sums of magnitude up to ~1e3; float32 spectrum, float32 log10)."""
import os

import numpy as np
import pytest

from oracle import ref_librosa as L
from oracle import ref_host as H

ATOL, RTOL = 5e-3, 1e-4


def test_mel_scale_and_filter_anchors():
    assert L.hz_to_mel(1000.0) == pytest.approx(15.0)
    assert L.hz_to_mel(6400.0) == pytest.approx(42.0)          # 27 mels per factor 6.4
    assert L.mel_to_hz(L.hz_to_mel([0, 440, 999.9, 1000, 4000, 8000])) == pytest.approx([0, 440, 999.9, 1000, 4000, 8000])
    m = L.mel_filters()
    assert m.shape == (128, 1025) and (m >= 0).all()
    # constant-area triangles: integral over frequency = 1 (bin spacing 7.8125 Hz), to sampling error
    area = m.sum(axis=1) * (16000 / 2048)
    assert np.abs(area[20:] - 1).max() < 0.02
    peaks = m.argmax(axis=1)
    assert (np.diff(peaks) > 0).all()
    d = L.dct_matrix(128, 128)
    assert np.allclose(d @ d.T, np.eye(128), atol=1e-12)
    assert L.num_frames(178240) == 1049 and L.num_frames(169) == 1 and L.num_frames(170) == 2


def test_restatement_matches_torchaudio(wav):
    torch = pytest.importorskip("torch")
    ta = pytest.importorskip("torchaudio")
    t = ta.transforms.MFCC(sample_rate=16000, n_mfcc=32, dct_type=2, norm="ortho", log_mels=False,
                           melkwargs=dict(n_fft=2048, hop_length=170, n_mels=128, center=True,
                                          pad_mode="reflect", power=2.0, norm="slaney", mel_scale="slaney"))
    rng = np.random.default_rng(3)
    for x in (wav[:30000], np.clip(rng.normal(0, 4000, 9000), -32768, 32767).astype(np.int16)):
        got = t(torch.from_numpy(x.astype(np.float32) / 32768.0)).numpy().T
        ref = L.mfcc(x)
        assert got.shape == ref.shape
        assert np.abs(got - ref).max() < 2e-3                  # torchaudio works in float32


def test_restatement_matches_the_transformers_port_of_librosa(wav):
    """A second independent implementation of the same definition: transformers.audio_utils is the
    numpy port of librosa's mel_filter_bank / stft / power_to_db that the Whisper feature extractor is
    built on.  Filters to rounding error, mel energies and the clamped dB rows (top_db = 80 under the
    clip maximum) at its float32 resolution; the ortho DCT-II of those rows is scipy's, which is what
    librosa.feature.mfcc itself calls."""
    au = pytest.importorskip("transformers.audio_utils")
    from scipy.fft import dct
    mf = au.mel_filter_bank(1025, 128, 0.0, 8000.0, 16000, norm="slaney", mel_scale="slaney")
    assert np.abs(mf.T - L.mel_filters()).max() < 1e-12
    rng = np.random.default_rng(5)
    for x in (wav[:40000], np.clip(rng.normal(0, 3000, 12000), -32768, 32767).astype(np.int16)):
        y = x.astype(np.float64) / 32768.0
        kw = dict(frame_length=2048, hop_length=170, fft_length=2048, power=2.0, center=True, pad_mode="reflect",
                  mel_filters=mf)
        ref, rt = L.mfcc(x, taps=True)
        mel = au.spectrogram(y, au.window_function(2048, "hann"), log_mel=None, **kw).T
        assert mel.shape == rt["mel"].shape and np.abs(mel - rt["mel"]).max() <= 1e-6 * rt["mel"].max()
        db = au.spectrogram(y, au.window_function(2048, "hann"), log_mel="dB", reference=1.0, min_value=1e-10,
                            db_range=80.0, **kw).T
        assert np.abs(db - rt["dbc"]).max() < 1e-4
        assert np.abs(dct(db.astype(np.float64), type=2, norm="ortho", axis=1)[:, :32] - ref).max() < 1e-3


def test_product_tables_match_restatement():
    from mfcc_b200 import tables as T
    assert np.allclose(T.lr_mel_filters(), L.mel_filters(), rtol=0, atol=1e-12)
    assert np.allclose(T.lr_window(), L.hann(2048), rtol=0, atol=1e-15)
    assert np.allclose(T.lr_dct_matrix(32, 128), L.dct_matrix(32, 128), rtol=0, atol=1e-15)
    assert np.allclose(T.lr_mel_filters(16000, 2048, 40), L.mel_filters(16000, 2048, 40), rtol=0, atol=1e-12)


# ------------------------------------------------------------------ CUDA
def _close(got, ref):
    err = np.abs(got.astype(np.float64) - ref)
    bound = ATOL + RTOL * np.abs(ref)
    assert (err <= bound).all(), "max err %g (bound %g)" % (err.max(), bound.flat[err.argmax()])


@pytest.mark.gpu
def test_cuda_librosa_set_against_restatement(wav):
    import torch
    import mfcc_b200
    op = mfcc_b200.LibrosaMFCC(sr=16000, hop_length=170, n_mfcc=32, device="cuda:0")
    x = wav
    got, t = op(torch.from_numpy(np.ascontiguousarray(x)).to("cuda:0"), taps=True)
    ref, rt = L.mfcc(x, taps=True)
    assert got.shape == ref.shape == (1049, 32) and got.dtype == torch.float32
    pw = t["power"].cpu().numpy().astype(np.float64)
    assert np.abs(pw - rt["power"]).max() <= 1e-5 * rt["power"].max()
    db = t["logmel"].cpu().numpy().astype(np.float64)
    audible = rt["db"] > rt["db"].max() - 80
    assert np.abs(db - rt["db"])[audible].max() < 2e-3
    _close(got.cpu().numpy(), ref)


@pytest.mark.gpu
def test_cuda_librosa_batch_short_clips_and_silence(wav):
    import torch
    import mfcc_b200
    op = mfcc_b200.LibrosaMFCC(device="cuda:0")
    rng = np.random.default_rng(11)
    for n in (1, 2, 169, 300, 1024, 1025, 5000):               # reflect padding folds many times below 1025
        x = np.stack([wav[20000:20000 + n], np.clip(rng.normal(0, 5000, n), -32768, 32767).astype(np.int16),
                      np.zeros(n, dtype=np.int16)])
        got = op(torch.from_numpy(x).to("cuda:0")).cpu().numpy()
        assert got.shape == (3, 1 + n // 170, 32)
        for i in range(2):
            if n >= 169 or i == 1:
                _close(got[i], L.mfcc(x[i]))
        # digital silence: every mel energy sits at amin = -100 dB -> c0 = -100 sqrt(128)
        assert np.allclose(got[2][:, 0], -100.0 * np.sqrt(128.0), atol=1e-2)
        assert np.abs(got[2][:, 1:]).max() < 1e-2


@pytest.mark.gpu
def test_cuda_librosa_dct_on_tensor_cores_and_on_fma_pipes(wav):
    """The default shape forms the DCT product with tensor-core MMAs (TF32 with head / tail splits,
    lr_dct_mma); the FMA-pipe kernel lr_dct_c stays selectable.  Both against the float64
    restatement inside the stated tolerance, on a batch whose frame count is not a multiple of the
    32-frame group (ragged last group, rows of two clips in one group, a silent clip whose rows sit
    at the dB floor), and close to each other at float32 resolution."""
    import torch
    import mfcc_b200
    op = mfcc_b200.LibrosaMFCC(device="cuda:0")
    rng = np.random.default_rng(3)
    n = 170 * 45 + 13                                            # 46 frames per clip
    x = np.stack([wav[30000:30000 + n], np.clip(rng.normal(0, 3000, n), -32768, 32767).astype(np.int16),
                  np.zeros(n, dtype=np.int16), wav[90000:90000 + n],
                  (8000 * np.sin(np.arange(n) * 0.3)).astype(np.int16)])
    xd = torch.from_numpy(x).to("cuda:0")
    got = {}
    for tc in (True, False):
        op.dct_path(tc)
        got[tc] = op(xd).cpu().numpy()
        assert got[tc].shape == (5, 46, 32)
        for i in (0, 1, 3, 4):
            _close(got[tc][i], L.mfcc(x[i]))
        assert np.allclose(got[tc][2][:, 0], -100.0 * np.sqrt(128.0), atol=1e-2)
    op.dct_path(True)
    scale = np.abs(got[False]).max()
    assert np.abs(got[True] - got[False]).max() <= 2e-5 * scale
    # pitched output rows and a single clip through the same kernel
    one = op(xd[0]).cpu().numpy()
    assert np.array_equal(one, got[True][0])


@pytest.mark.gpu
def test_cuda_librosa_other_sizes_and_scale_chain(wav):
    import torch
    import mfcc_b200
    from mfcc_b200 import post
    x = wav[60000:100000]
    xd = torch.from_numpy(np.ascontiguousarray(x)).to("cuda:0")
    for kw in (dict(n_mfcc=20, n_mels=128, hop_length=512), dict(n_mfcc=13, n_mels=40, hop_length=160),
               dict(n_mfcc=64, n_mels=64, hop_length=170, top_db=None)):
        op = mfcc_b200.LibrosaMFCC(device="cuda:0", **kw)
        ref = L.mfcc(x, hop=kw["hop_length"], n_mfcc=kw["n_mfcc"], n_mels=kw["n_mels"],
                     top_db=kw.get("top_db", 80.0) if kw.get("top_db", 80.0) is not None else 1e9)
        _close(op(xd).cpu().numpy(), ref)
    # uncentred frames (stft center=False), another rate (different filter bank -> run-time step counts)
    op = mfcc_b200.LibrosaMFCC(device="cuda:0", center=False, hop_length=256)
    got = op(xd).cpu().numpy()
    assert got.shape == (1 + (len(x) - 2048) // 256, 32)
    _close(got, L.mfcc(x, hop=256, center=False))
    op = mfcc_b200.LibrosaMFCC(sr=22050, device="cuda:0")
    _close(op(xd).cpu().numpy(), L.mfcc(x, sr=22050))
    # genlibrosa.py:25-28: scale over time, then int16
    op = mfcc_b200.LibrosaMFCC(device="cuda:0")
    got = post.scale(op(xd)).cpu().numpy()
    want = H.scale_time(L.mfcc(x))
    assert got.dtype == np.int16
    tr = np.trunc(want)
    near = np.abs(want - np.round(want)) < 1e-3                # float32 noise may cross a truncation boundary
    assert np.array_equal(got[~near], tr[~near].astype(np.int16))
    assert np.abs(got - tr).max() <= 1


@pytest.mark.gpu
def test_cuda_librosa_argument_errors():
    import torch
    import mfcc_b200
    with pytest.raises(ValueError):
        mfcc_b200.LibrosaMFCC(n_fft=1024, device="cuda:0")
    with pytest.raises(mfcc_b200._lib.MfccError):
        mfcc_b200.LibrosaMFCC(n_mfcc=65, device="cuda:0")
    op = mfcc_b200.LibrosaMFCC(device="cuda:0")
    with pytest.raises(TypeError):
        op(torch.zeros(100, device="cuda:0"))
    with pytest.raises(ValueError):
        op(torch.zeros(0, dtype=torch.int16, device="cuda:0"))
