"""CPU tests: the oracle against the reference's own printed values, the two oracle
implementations against each other, and the product's tables against both."""
import hashlib
import json
import os

import numpy as np
import pytest

from mfcc_b200 import MFCCConfig, tables as T
from oracle import ref_c as RC
from oracle import ref_np as R

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def pins():
    return json.load(open(os.path.join(GOLD, "notebook_pins.json")))


@pytest.fixture(scope="module")
def stim():
    return json.load(open(os.path.join(GOLD, "stimuli.json")))


# ---- pins printed by the reference's notebooks -------------------------------------
def test_window_curve_matches_notebook(pins):
    want = np.array(pins["window_curve"])
    assert np.array_equal(R.window_curve(512, 8, 1)[0], want)          # oracle
    assert np.array_equal(T.window_curve(512, 8), want)                # product tables
    mem, off_fst, off_lst = R.window_coeffs(512, 8)
    assert mem == pins["window_small"]
    assert off_lst == pins["window_offsetlast"] == 470
    assert off_fst == 40 and pins["window_after"] == 275 == want[128]


def test_window_curve_is_frame_invariant():
    c = R.window_curve(512, 8, 3)
    assert np.array_equal(c[0], c[1]) and np.array_equal(c[1], c[2])


def test_mel_points_match_notebook(pins):
    want = np.array(pins["mel_points_32"])
    assert np.array_equal(R.get_filter_points(0, 8000.0, 32, 512, sample_rate=16e3), want)
    assert np.array_equal(T.mel_points(32, 512, 16e3), want)


def test_frame_count_matches_notebook(pins):
    assert R.num_frames(16000) == pins["frames_in_16000"] == MFCCConfig().num_frames(16000)


def test_host_frame_count():
    c = MFCCConfig()
    # software/main.c:95: ((nsamples - NFFT) / STEPSIZE) + 1 + 1
    for n in (512, 513, 681, 682, 683, 178240):
        assert c.num_frames_host(n) == (n - 512) // 170 + 2


# ---- tables: product == oracle -----------------------------------------------------
@pytest.mark.parametrize("size", [512, 128, 64])
def test_twiddles(size):
    t = T.twiddle_table(size)
    for k in range(size // 2):
        assert tuple(t[k]) == R.twiddle(size, k)
    if size == 512:
        re, im = R.twiddle_rom(512)
        assert re[:5] == [16384, 16383, 16379, 16373, 16364] and im[:5] == [0, -201, -402, -603, -804]


@pytest.mark.parametrize("nf", [16, 32])
def test_mel_tables_reproduce_filterbank(nf):
    """weights/segments from mfcc_b200.tables drive the same sums as the register-level
    FilterBank emulation."""
    from lane_model import mel_lanes
    import math
    rng = np.random.default_rng(nf)
    pw = rng.integers(0, 1 << 30, (6, 256))
    pw[0] = 0
    pw[1] = (1 << 30) - 1
    pts = R.get_filter_points(0, 8000.0, nf, 512, sample_rate=16e3)
    lo = int(math.log2(pts[-1] - pts[-3])) + 60 - 34
    want = R.filterbank(pw, ntap=nf)
    for k in range(len(pw)):
        assert np.array_equal(mel_lanes(pw[k], nf, lo_bit=lo), want[k])


# ---- FFT: loop == bank-level emulation of the Scheduler ------------------------------
@pytest.mark.parametrize("size", [64, 128, 512])
def test_fft_loop_equals_bank_emulation(size, stim):
    rng = np.random.default_rng(size)
    cases = [(rng.integers(-32768, 32768, size), rng.integers(-32768, 32768, size)),
             (np.full(size, -32768), np.full(size, 32767))]
    if size == 512:
        x = np.array(stim["fft512_hex"], dtype=np.uint16).astype(np.int16).astype(np.int64)
        cases.append((x, np.zeros(size, dtype=np.int64)))
    for xr, xi in cases:
        a = R.fft_int(xr, xi, size)
        b = R.fft_rtl_banks(xr, xi, size)
        assert np.array_equal(a[0][: size // 2], b[0]) and np.array_equal(a[1][: size // 2], b[1])
        c = RC.fft(xr[None, :], xi[None, :], size)
        assert np.array_equal(c[0][0], a[0]) and np.array_equal(c[1][0], a[1])


def test_fft_close_to_float(stim):
    x = np.array(stim["fft512_hex"], dtype=np.uint16).astype(np.int16).astype(np.int64)
    vr, vi = R.fft_stream(x[None, :])
    f = np.fft.fft(x) / 512
    assert np.abs(vr[0] - f.real[:256]).max() <= 4 and np.abs(vi[0] - f.imag[:256]).max() <= 4


def test_dct_close_to_float(stim):
    from scipy.fft import dct
    x = np.array(stim["dct16_hex"], dtype=np.uint16).astype(np.int16).astype(np.int64)
    got = R.dct_stream(x[None, :], 16)[0]
    want = dct(x.astype(float), type=2) / 64        # the bench compares with dct(tst) // 64
    assert np.abs(got - want).max() <= 3


def test_log2_known_value(stim):
    # log.py:143-162: Log2Fix(37, 20) on 2207315; precision = 20 - ceil(log2 37) = 14
    got = int(R.log2fix(np.array([stim["log2_value"]]), 37, 20)[0])
    assert abs(got / (1 << 14) - np.log2(stim["log2_value"])) < 2e-3
    m = np.arange(0, 65536, 13)
    a = R.log2fix(m, 16, 15)
    assert np.array_equal(RC.log2(m), a)
    ref = 2048 * np.log2(np.maximum(m, 1))
    assert np.abs(a - ref).max() < 4.0 and a[0] == 0


def test_filterbank_constant_input(stim):
    """filterbank.py:147-176 feeds a constant; the float notebook prints the matching
    row sums (MFCC.ipynb:920-951).  With unit weight sums s_m the output is
    (1234 * s_m) >> 1 for our widths; check the first rows of the survey table."""
    pw = np.full((1, 256), stim["filterbank_const"])
    out = R.filterbank(pw, ntap=32)[0]
    sums = [2, 2, 2.5, 2.5, 2.5, 2.5, 2.5, 3.5, 3.5, 3.5, 4, 4.5, 4.5, 5, 5.5, 5.5]
    for m, s in enumerate(sums):
        assert abs(out[m] - 1234 * s / 2) <= 1, (m, out[m])


# ---- whole chain: C oracle == numpy oracle ------------------------------------------
def test_chain_c_equals_numpy_on_wav(wav):
    cfg = RC.RefConfig(nfilters=32, nceptrums=32)
    x = wav[:40000]
    a, ta = RC.run(cfg, x, taps=True)
    b, tb = R.mfcc(x, nfilters=32, nceptrums=32, taps=True)
    assert a.shape == (233, 32) and np.array_equal(a, b)
    assert np.array_equal(ta["window"], tb["window"])
    assert np.array_equal(ta["fft"][..., 0], tb["fft_r"]) and np.array_equal(ta["fft"][..., 1], tb["fft_i"])
    assert np.array_equal(ta["power"], tb["power"]) and np.array_equal(ta["mel"], tb["filter"])
    assert np.array_equal(ta["log2"], tb["log"]) and np.array_equal(ta["dct"], tb["dct"])


def test_wav_regression_marker(wav):
    """First frame and SHA-1 of the whole clip: derived twice independently (the survey's
    scratch model and this oracle); a regression marker, not a reference vector."""
    out = RC.run(RC.RefConfig(), wav)
    assert out.shape == (1046, 32)
    assert list(out[0][:8]) == [4059, 1164, -93, -353, -597, -289, -221, -201]
    assert hashlib.sha1(out.astype("<i2").tobytes()).hexdigest() == \
        "6811dedb5eff92d39704422b2b2ae2734bf6b8ca"


@pytest.mark.parametrize("nf,nc", [(32, 32), (32, 16), (16, 16), (16, 5)])
def test_chain_edge_inputs(nf, nc):
    cfg = RC.RefConfig(nfilters=nf, nceptrums=nc)
    rng = np.random.default_rng(7)
    cases = {
        "noise": rng.integers(-32768, 32768, 3000),
        "alt": np.tile([-32768, 32767], 1200),
        "max": np.full(1500, 32767), "min": np.full(1500, -32768),
        "zero": np.zeros(1200), "short": rng.integers(-100, 100, 300),
        "imp": (np.arange(2000) == 700) * 32767,
    }
    for name, x in cases.items():
        x = x.astype(np.int16)
        for nfr in (None, 12):
            a = RC.run(cfg, x, n_frames=nfr)
            b = R.mfcc(x, nfilters=nf, nceptrums=nc, n_frames=nfr)
            assert a.shape == b.shape and np.array_equal(a, b), (name, nfr)


@pytest.mark.parametrize("nf,nc", [(32, 32), (32, 13), (16, 16)])
def test_timed_baseline_equals_checker(nf, nc, wav):
    """oracle/mfcc_ref_fast.c (gcc -O3 -march=native -fopenmp, the build bench.py times as
    cpu_baseline / --impl reference) must be the checker's arithmetic bit for bit."""
    cfg = RC.RefConfig(nfilters=nf, nceptrums=nc)
    rng = np.random.default_rng(11)
    n = 20000
    x = np.stack([
        rng.integers(-32768, 32768, n), (rng.standard_normal(n) * 2000).clip(-32768, 32767),
        np.tile([-32768, 32767], n // 2), np.tile([32767, 32767, -32768, -32768], n // 4),
        np.full(n, 32767), np.full(n, -32768), np.zeros(n), (np.arange(n) == 700) * 32767,
        wav[:n], np.ones(n)]).astype(np.int16)
    assert np.array_equal(RC.run(cfg, x, fast=True, threads=2), RC.run(cfg, x))
    # ragged: frames past the end of the clip read zeros; a single row; more threads than clips
    assert np.array_equal(RC.run(cfg, x[:, :700], n_frames=9, fast=True), RC.run(cfg, x[:, :700], n_frames=9))
    assert np.array_equal(RC.run(cfg, x[0], fast=True, threads=4), RC.run(cfg, x[0]))
    assert "-march=native" in RC.FAST_FLAGS and "-O3" in RC.FAST_FLAGS


def test_float_sanity(wav):
    """The integer chain tracks a float pipeline of the same structure (loose bound:
    guards against gross restatement errors, e.g. a wrong shift)."""
    from scipy.fft import dct
    x = wav[:20000].astype(np.float64)
    out, tp = R.mfcc(wav[:20000], nfilters=32, nceptrums=32, taps=True)
    y = np.concatenate([[x[0]], x[1:] - 0.96875 * x[:-1]])
    assert np.abs(tp["preemph"] - y).max() <= 1.0
    fr = R.frame(np.round(y), 512)
    f = np.fft.fft(fr * T.window_curve() / 512.0, axis=1)[:, :256] / 512
    assert np.abs(tp["fft_r"] - f.real).max() <= 12
    lg = tp["log"] / 2048.0
    want = dct(lg, type=2, axis=1) / 128 * 2048
    assert np.abs(tp["dct"] - want).max() <= 3
