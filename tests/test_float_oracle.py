"""fp32 mode: the float64 oracle (oracle/ref_float.py) against outputs of the
reference's float prototype executed by tests/golden/make_float_golden.py
(notebook/MFCC-INT.ipynb cells 2-10), and the CUDA fp32 path against both.

Stated tolerance of the fp32 CUDA mode: per coefficient
    |got - ref| <= 2e-3 + 1e-4 * |ref|
on every frame whose mel energies are all >= 2**-10 (log2 of a smaller energy
amplifies float32 rounding of the spectrum without bound)."""
import os

import numpy as np
import pytest

from oracle import ref_float as F

HERE = os.path.dirname(os.path.abspath(__file__))
ATOL, RTOL, MEL_MIN = 2e-3, 1e-4, 2.0 ** -10


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(HERE, "golden", "notebook_float.npz"))


def clip(gold, name, wav):
    return wav if name == "wav" else gold["synth_samples"]


@pytest.mark.parametrize("name", ["wav", "synth"])
def test_oracle_reproduces_the_notebook(gold, wav, name):
    x = clip(gold, name, wav)
    out, t = F.mfcc(x, taps=True)
    a, b = gold[name + "_slice"]
    n = len(gold[name + "_emphasis"])
    assert np.array_equal(t["preemph"][:n], gold[name + "_emphasis"])
    assert np.allclose(t["window"][a:b], gold[name + "_window"], rtol=0, atol=1e-9)
    assert np.allclose(t["spec"][a:b], gold[name + "_fft"], rtol=1e-12, atol=1e-9)
    assert np.allclose(t["power"][a:b], gold[name + "_power"], rtol=1e-11, atol=1e-9)
    assert np.allclose(t["mel"][a:b], gold[name + "_mel"], rtol=1e-11, atol=1e-9)
    assert np.allclose(t["log"][a:b], gold[name + "_log"], rtol=0, atol=1e-10)
    assert out.shape == gold[name + "_dct"].shape
    assert np.abs(out - gold[name + "_dct"]).max() < 1e-9
    assert np.array_equal(F.filter_points(), gold[name + "_filter_points"])
    assert np.array_equal(F.filters(), gold[name + "_filters"])


def test_oracle_frame_count_and_padding(wav):
    assert F.num_frames(178240) == 1046 and F.num_frames(511) == 0 and F.num_frames(512) == 1
    x = wav[30000:30700]
    full = F.mfcc(np.concatenate([x, np.zeros(512 + 170 * 3 - 700, dtype=np.int16)]), mel_floor=1e-3)
    assert np.allclose(F.mfcc(x, n_frames=4, mel_floor=1e-3), full)


def test_product_tables_match_oracle():
    from mfcc_b200 import tables as T
    assert np.allclose(T.f_window(512), F.window(512), rtol=0, atol=1e-7)
    assert np.allclose(T.f_mel_filters(32, 512, 16e3), F.filters(32, 512, 16000), rtol=0, atol=1e-7)
    assert np.allclose(T.f_dct_matrix(32), F.dct_matrix(32), rtol=0, atol=1e-7)
    assert np.allclose(T.f_dct_matrix(16), F.dct_matrix(16), rtol=0, atol=1e-7)


# ------------------------------------------------------------------ CUDA fp32 mode
def _check(got, ref, mel):
    ok = (mel >= MEL_MIN).all(axis=1)
    assert ok.sum() > 0.5 * len(ok)
    err = np.abs(got[ok] - ref[ok])
    bound = ATOL + RTOL * np.abs(ref[ok])
    assert (err <= bound).all(), "max err %g at bound %g" % (err.max(), bound.flat[err.argmax()])
    return err.max()


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["wav", "synth"])
def test_cuda_fp32_against_notebook_and_oracle(gold, wav, name):
    import torch
    import mfcc_b200
    x = clip(gold, name, wav)
    op = mfcc_b200.MFCC(nfilters=32, nceptrums=32, mode="fp32", device="cuda:0")
    got = op(torch.from_numpy(np.ascontiguousarray(x)).to("cuda:0")).cpu().numpy()
    assert got.dtype == np.float32 and got.shape == gold[name + "_dct"].shape
    ref, t = F.mfcc(x, taps=True)
    _check(got.astype(np.float64), gold[name + "_dct"], t["mel"])     # the notebook's own output
    _check(got.astype(np.float64), ref, t["mel"])


@pytest.mark.gpu
@pytest.mark.parametrize("nf,nc", [(32, 16), (16, 16), (16, 7)])
def test_cuda_fp32_other_filter_sets(wav, nf, nc):
    import torch
    import mfcc_b200
    x = wav[40000:80000]
    op = mfcc_b200.MFCC(nfilters=nf, nceptrums=nc, mode="fp32", device="cuda:0")
    got = op(torch.from_numpy(np.ascontiguousarray(x)).to("cuda:0")).cpu().numpy()
    ref, t = F.mfcc(x, nfilters=nf, nceptrums=nc, taps=True)
    assert got.shape == ref.shape
    _check(got.astype(np.float64), ref, t["mel"])


@pytest.mark.gpu
@pytest.mark.parametrize("K", [1, 7, 8, 9, 15, 16, 17, 31, 32, 33, 47, 48, 49, 65])
def test_cuda_fp32_frame_counts_around_the_tile_and_the_frame_pairs(wav, K):
    """The fp32 kernel cuts a clip into 32-frame tiles and a warp carries frames f and f + 8
    together; every frame count around those sizes, two clips per launch, each frame equal to the
    same frame of a one-clip launch (the tiling is invisible) and inside the tolerance."""
    import torch
    import mfcc_b200
    n = 512 + 170 * (K - 1) + 3
    x = np.stack([wav[40000:40000 + n], wav[90000:90000 + n]])
    op = mfcc_b200.MFCC(nfilters=32, nceptrums=32, mode="fp32", mel_floor=1e-3, device="cuda:0")
    got = op(torch.from_numpy(np.ascontiguousarray(x)).to("cuda:0")).cpu().numpy()
    assert got.shape == (2, K, 32)
    for i in range(2):
        one = op(torch.from_numpy(np.ascontiguousarray(x[i])).to("cuda:0")).cpu().numpy()
        assert np.array_equal(one, got[i])
        ref, t = F.mfcc(x[i], mel_floor=1e-3, taps=True)
        _check(got[i].astype(np.float64), ref, t["mel"])


@pytest.mark.gpu
def test_cuda_fp32_batch_ragged_floor_and_host(wav):
    import torch
    import mfcc_b200
    rng = np.random.default_rng(5)
    n = 512 + 170 * 40 + 57                                   # ragged tail, rows not 16-byte multiples
    x = np.stack([wav[1000:1000 + n], wav[90000:90000 + n],
                  np.clip(rng.normal(0, 3000, n), -32768, 32767).astype(np.int16),
                  np.zeros(n, dtype=np.int16)])
    op = mfcc_b200.MFCC(nfilters=32, nceptrums=32, mode="fp32", mel_floor=1e-3, device="cuda:0")
    K = op.num_frames(n) + 2                                  # two zero-extended frames
    got = op(torch.from_numpy(x).to("cuda:0"), n_frames=K).cpu().numpy()
    assert got.shape == (4, K, 32) and np.isfinite(got).all()
    for i in range(4):
        ref, t = F.mfcc(x[i], n_frames=K, mel_floor=1e-3, taps=True)
        if i < 3:
            _check(got[i].astype(np.float64), ref, t["mel"])
    # silence: every mel energy is clamped to the floor -> c0 = sqrt(32) * log2(floor), rest 0
    assert np.allclose(got[3][:, 0], np.sqrt(32) * np.log2(1e-3), atol=1e-3)
    assert np.abs(got[3][:, 1:]).max() < 1e-3
    host = op.run_host(torch.from_numpy(x), n_frames=K).numpy()
    assert np.array_equal(host, got)


@pytest.mark.gpu
def test_cuda_fp32_streaming_matches_batch(wav):
    import torch
    import mfcc_b200
    op = mfcc_b200.MFCC(nfilters=32, nceptrums=32, mode="fp32", mel_floor=1e-3, device="cuda:0")
    S, steps = 5, 12
    x = np.stack([wav[5000 * (i + 1):5000 * (i + 1) + 170 * steps] for i in range(S)])
    op.reset(S)
    cols = [op.step(torch.from_numpy(np.ascontiguousarray(x[:, 170 * k:170 * (k + 1)])).to("cuda:0")).cpu().numpy()
            for k in range(steps)]
    # the state after reset holds 342 zeros: prepend them to get the batch equivalent
    xb = np.concatenate([np.zeros((S, 342), dtype=np.int16), x], axis=1)
    want = op(torch.from_numpy(xb).to("cuda:0")).cpu().numpy()
    assert want.shape[1] == steps
    assert np.array_equal(np.stack(cols, axis=1), want)
