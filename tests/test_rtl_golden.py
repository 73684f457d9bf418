"""Parity against vectors produced by EXECUTING the reference's own RTL source
(tests/golden/make_rtl_golden.py: /root/reference/mfcc/core/*.py run under
oracle/hdl_shim).  The CPU tests pin both oracles (numpy and C) stage by stage; the GPU
tests pin the CUDA path the same way, through the C-ABI."""
import glob
import os

import numpy as np
import pytest

from oracle import ref_np as R

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")
CASES = sorted(os.path.basename(p)[4:-4] for p in glob.glob(os.path.join(GOLD, "rtl_*.npz"))
               if not p.endswith(("rtl_stage_vectors.npz", "rtl_frame_vectors.npz", "rtl_size_vectors.npz", "rtl_param_vectors.npz")))


def load(case):
    return np.load(os.path.join(GOLD, "rtl_%s.npz" % case))


def s16(v):
    return ((np.asarray(v, dtype=np.int64) + 32768) % 65536) - 32768


def test_fixture_set_is_complete():
    assert len(CASES) >= 14, CASES
    assert os.path.exists(os.path.join(GOLD, "rtl_stage_vectors.npz"))


@pytest.mark.parametrize("case", CASES)
def test_numpy_oracle_matches_rtl_every_stage(case):
    g = load(case)
    nf, nc = int(g["nfilters"]), int(g["nceptrums"])
    K = g["out"].shape[0]
    out, t = R.mfcc(g["samples"], nfilters=nf, nceptrums=nc, n_frames=K, taps=True)
    assert np.array_equal(t["frame"], s16(g["frame"]))
    assert np.array_equal(t["window"], s16(g["window"]))
    assert np.array_equal(t["fft_r"], s16(g["fft_r"]))
    assert np.array_equal(t["fft_i"], s16(g["fft_i"]))
    assert np.array_equal(t["power"], g["power"].astype(np.int64) & 0x3fffffff)
    assert np.array_equal(t["filter"], g["filter"].astype(np.int64) & 0xffff)
    assert np.array_equal(t["log"], g["log"])
    assert np.array_equal(t["dct"], s16(g["dct"]))
    assert np.array_equal(out, s16(g["out"]))


@pytest.mark.parametrize("case", CASES)
def test_c_oracle_matches_rtl_every_stage(case):
    from oracle import ref_c as RC
    RC.build()
    g = load(case)
    nf, nc = int(g["nfilters"]), int(g["nceptrums"])
    K = g["out"].shape[0]
    out, t = RC.run(RC.RefConfig(nfilters=nf, nceptrums=nc), g["samples"], n_frames=K, taps=True)
    assert np.array_equal(t["window"], s16(g["window"]))
    assert np.array_equal(t["fft"][..., 0], s16(g["fft_r"]))
    assert np.array_equal(t["fft"][..., 1], s16(g["fft_i"]))
    assert np.array_equal(t["power"], g["power"].astype(np.uint32))
    assert np.array_equal(t["mel"], g["filter"].astype(np.uint16))
    assert np.array_equal(t["log2"], g["log"].astype(np.uint16))
    assert np.array_equal(t["dct"], s16(g["dct"]))
    assert np.array_equal(out, s16(g["out"]))


def test_stalled_run_equals_free_run():
    """Back-pressure in the RTL bench must not change any value: the fixture made with
    random stalls on both ends equals the oracle just like the free-running ones, and
    its inputs overlap wav_mid_16c's, whose common frames must agree."""
    a, b = load("wav_stalled"), load("wav_mid_16c")
    n = min(a["dct"].shape[0], b["dct"].shape[0])
    assert np.array_equal(a["dct"][:n], b["dct"][:n])
    assert np.array_equal(a["filter"][:n], b["filter"][:n])


def test_golden_covers_the_hard_cases():
    """The fixtures must exercise what makes the path intricate: the 16-bit mel wrap
    (SURVEY.md F7), log2(0), FFT 16-bit wraps and the zero-padded tail."""
    g = load("noise_full")
    # mel values whose unwrapped register exceeds 16 bits exist for full-scale noise
    pw = g["power"].astype(np.int64)
    wide = R.filterbank(pw, 30, 24, 10, 16e3, 512, 32)        # same first bit, 8 more kept
    assert (wide >= 65536).any() and np.array_equal(wide & 0xffff, g["filter"])
    assert (load("zeros")["filter"] == 0).all() and (load("zeros")["log"] == 0).all()
    assert load("short_padded")["samples"].size < 512 + 170 * 3
    alt = load("alt_fullscale")
    assert np.abs(alt["fft_r"]).max() > 0


# ------------------------------------------------------------------ single stages
@pytest.fixture(scope="module")
def sv():
    return np.load(os.path.join(GOLD, "rtl_stage_vectors.npz"))


def test_stage_fft_core_and_stream(sv):
    x = sv["fft512_in"]
    re, im = R.fft_int(x, np.zeros_like(x), 512)
    assert np.array_equal(re[:256], sv["fft512_out"][:, 0])
    assert np.array_equal(im[:256], sv["fft512_out"][:, 1])
    re, im = R.fft_stream(sv["fftstream_in"], 512)
    assert np.array_equal(re, sv["fftstream_out"][:, :, 0])
    assert np.array_equal(im, sv["fftstream_out"][:, :, 1])


def test_stage_dct(sv):
    assert np.array_equal(R.dct_stream(sv["dct16_in"], 16), s16(sv["dct16_out"]))
    assert np.array_equal(R.dct_stream(sv["dct32_in"], 32), s16(sv["dct32_out"]))
    assert np.array_equal(R.dct_stream(sv["dct32s_in"], 32), s16(sv["dct32s_out"]))


def test_stage_log2(sv):
    assert np.array_equal(R.log2fix(sv["log2_in"], 16, 15), sv["log2_out"])


def test_stage_power(sv):
    p = sv["power_in"]
    assert np.array_equal(R.power_spectrum(p[:, 0], p[:, 1], 16, 30), sv["power_out"])


@pytest.mark.parametrize("ntap", [16, 32])
def test_stage_filterbank(sv, ntap):
    got = R.filterbank(sv["mel%d_in" % ntap], 30, 16, 18, 16e3, 512, ntap)
    assert np.array_equal(got, sv["mel%d_out" % ntap])
    # the all-ones row overflows the 16 kept bits on the wide filters: the wrap is real
    assert (sv["mel%d_out" % ntap] < 65536).all()


def test_stage_window(sv):
    assert np.array_equal(R.window(sv["window_in"], 512, 8), sv["window_out"])


# ------------------------------------------------------------------ CUDA path
@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES)
def test_cuda_matches_rtl_every_stage(case):
    import torch
    import mfcc_b200
    g = load(case)
    nf, nc = int(g["nfilters"]), int(g["nceptrums"])
    K = g["out"].shape[0]
    op = mfcc_b200.MFCC(nfilters=nf, nceptrums=nc, device="cuda:0")
    x = torch.from_numpy(np.ascontiguousarray(g["samples"])).to("cuda:0")
    out, t = op(x, n_frames=K, taps=True)
    torch.cuda.synchronize()
    u16 = lambda a: a.cpu().numpy().view(np.uint16).astype(np.int64)
    assert np.array_equal(t["window"].cpu().numpy(), s16(g["window"]))
    assert np.array_equal(t["fft"].cpu().numpy()[..., 0], s16(g["fft_r"]))
    assert np.array_equal(t["fft"].cpu().numpy()[..., 1], s16(g["fft_i"]))
    assert np.array_equal(t["power"].cpu().numpy().view(np.uint32), g["power"].astype(np.uint32))
    assert np.array_equal(u16(t["filter"]), g["filter"])
    assert np.array_equal(u16(t["log"]), g["log"])
    assert np.array_equal(t["dct"].cpu().numpy(), s16(g["dct"]))
    assert np.array_equal(out.cpu().numpy(), s16(g["out"]))
    # the production instantiation (no taps, batched with a copy of itself)
    out2 = op(torch.stack([x, x]), n_frames=K)
    assert np.array_equal(out2[1].cpu().numpy(), s16(g["out"]))


@pytest.mark.gpu
def test_cuda_stage_operators_match_rtl_stage_vectors(sv):
    import torch
    import mfcc_b200 as M
    dev = "cuda:0"
    T = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a).astype(dt)).to(dev)
    got = M.FftStream(16, 512, device=dev)(T(sv["fftstream_in"], np.int16)).cpu().numpy()
    assert np.array_equal(got, sv["fftstream_out"])
    got = M.WindowHamming(16, 512, 8, device=dev)(T(sv["window_in"], np.int16)).cpu().numpy()
    assert np.array_equal(got, sv["window_out"])
    got = M.Log2Fix(16, 15, device=dev)(T(sv["log2_in"].astype(np.uint16).view(np.int16), np.int16))
    assert np.array_equal(got.cpu().numpy().view(np.uint16), sv["log2_out"])
    for ntap in (16, 32):
        fb = M.FilterBank(30, 16, None, 18, 16e3, 512, ntap, device=dev)
        got = fb(T(sv["mel%d_in" % ntap], np.int32)).cpu().numpy().view(np.uint16)
        assert np.array_equal(got, sv["mel%d_out" % ntap])
    d = M.DCTStream(16, 32, device=dev)
    got = d(T(sv["dct32_in"].astype(np.uint16).view(np.int16), np.int16)).cpu().numpy()
    assert np.array_equal(got, s16(sv["dct32_out"]))
    got = d(T(sv["dct32s_in"], np.int16)).cpu().numpy()                  # full signed range
    assert np.array_equal(got, s16(sv["dct32s_out"]))
    got = M.DCTStream(16, 16, device=dev)(T(sv["dct16_in"].astype(np.uint16).view(np.int16), np.int16))
    assert np.array_equal(got.cpu().numpy(), s16(sv["dct16_out"]))
    p = np.zeros((1, 256, 2), dtype=np.int64)                   # the operator takes whole frames
    p[0, :len(sv["power_in"])] = sv["power_in"]
    got = M.PowerSpectrum(16, 30, device=dev)(T(p, np.int16)).cpu().numpy().view(np.uint32)
    assert np.array_equal(got[0, :len(sv["power_out"])], sv["power_out"])


# ---------------------------------------------------------------- Frame / Preemph on their own
FRAME_SETS = {"400_160_512": (400, 160, 512), "512_170_512": (512, 170, 512),
              "96_96_128": (96, 96, 128), "100_30_128": (100, 30, 128), "25_8_32": (25, 8, 32)}


@pytest.fixture(scope="module")
def fvec():
    return np.load(os.path.join(GOLD, "rtl_frame_vectors.npz"))


@pytest.mark.parametrize("name", sorted(FRAME_SETS))
def test_oracle_frame_matches_rtl(fvec, name):
    """Frame(16, windowlen, stepsize, nfft) of the reference, run under random back-pressure
    (tests/golden/make_rtl_golden.py frame_vectors): the constructor defaults / bench case
    400-160-512 (frame.py:50, :233-249), MFCC's 512-170-512, and two small sets."""
    wl, st, nfft = FRAME_SETS[name]
    x, want = fvec["frame_%s_in" % name], fvec["frame_%s_out" % name]
    got = R.frame(x, nfft=nfft, hop=st, n_frames=want.shape[0], windowlen=wl)
    assert np.array_equal(got, want)
    assert R.frame(x, nfft=nfft, hop=st, windowlen=wl).shape[0] == (len(x) - wl) // st + 1


def test_oracle_preemph_matches_rtl(fvec):
    assert np.array_equal(R.preemph(fvec["preemph_in"]), fvec["preemph_out"])


def test_oracle_power_24_bits_matches_rtl(fvec):
    """PowerSpectrum with its class default width_output = 24 (pow2.py:44, bench at :74)."""
    x = fvec["power24_in"]
    assert np.array_equal(R.power_spectrum(x[:, 0], x[:, 1], 16, 24), fvec["power24_out"])


def test_oracle_log2_37_20_matches_rtl(fvec):
    """Log2Fix(37, 20): the widths and the value (2207315) of the reference's own bench, log.py:143-149."""
    assert np.array_equal(R.log2fix(fvec["log2_37_20_in"], 37, 20), fvec["log2_37_20_out"])
    v = int(R.log2fix(np.array([2207315]), 37, 20)[0])
    assert abs(v / 2 ** 14 - np.log2(2207315)) < 2e-3          # the bench prints res / 2**precision


def test_oracle_filterbank_other_widths_matches_rtl(fvec):
    """FilterBank(31, 32) (the reference's own bench, filterbank.py:148, 68-bit accumulators) and
    FilterBank() with the class defaults (24 / 24 / gain 0 / 16 taps)."""
    got = R.filterbank(fvec["mel_31_32_in"], width=31, width_output=32, gain=0, sample_rate=16000, nfft=512, ntap=16)
    assert np.array_equal(got, fvec["mel_31_32_out"])
    got = R.filterbank(fvec["mel_default_in"], width=24, width_output=24, gain=0, sample_rate=16000, nfft=512, ntap=16)
    assert np.array_equal(got, fvec["mel_default_out"])


@pytest.mark.gpu
def test_cuda_frame_and_preemph_match_rtl(fvec):
    import torch
    import mfcc_b200
    for name, (wl, st, nfft) in FRAME_SETS.items():
        x, want = fvec["frame_%s_in" % name], fvec["frame_%s_out" % name]
        op = mfcc_b200.Frame(16, wl, st, nfft, device="cuda:0")
        xd = torch.from_numpy(x.astype(np.int16)).to("cuda:0")
        got = op(xd, n_frames=want.shape[0]).cpu().numpy()
        assert np.array_equal(got, want), name
        full = op(xd).cpu().numpy()                            # K = full frames of the stream
        assert full.shape == ((len(x) - wl) // st + 1, nfft) and np.array_equal(full[:want.shape[0]], want)
        ext = op(xd, n_frames=full.shape[0] + 3).cpu().numpy()  # zero-extended tail
        assert np.array_equal(ext, R.frame(x, nfft=nfft, hop=st, n_frames=full.shape[0] + 3, windowlen=wl))
    assert mfcc_b200.Frame(device="cuda:0").windowlen == 400    # the reference's defaults construct
    with pytest.raises(ValueError):
        mfcc_b200.Frame(16, 513, 170, 512, device="cuda:0")     # frame.py:51
    pre = mfcc_b200.Preemph(16, device="cuda:0")
    got = pre(torch.from_numpy(fvec["preemph_in"].astype(np.int16)).to("cuda:0")).cpu().numpy()
    assert np.array_equal(got, fvec["preemph_out"])
    fb = mfcc_b200.FilterBank(width=31, width_output=32, device="cuda:0")            # filterbank.py:148
    got = fb(torch.from_numpy(fvec["mel_31_32_in"].astype(np.int32)).to("cuda:0")).cpu().numpy()
    assert np.array_equal(got.view(np.uint32).astype(np.int64), fvec["mel_31_32_out"])
    fb = mfcc_b200.FilterBank(device="cuda:0")                                       # class defaults
    got = fb(torch.from_numpy(fvec["mel_default_in"].astype(np.int32)).to("cuda:0")).cpu().numpy()
    assert np.array_equal(got.view(np.uint32).astype(np.int64), fvec["mel_default_out"])
    lg = mfcc_b200.Log2Fix(37, 20, device="cuda:0")             # log.py:143
    got = lg(torch.from_numpy(fvec["log2_37_20_in"]).to("cuda:0")).cpu().numpy()
    assert np.array_equal(got.astype(np.int64), fvec["log2_37_20_out"])
    with pytest.raises(ValueError):
        mfcc_b200.Log2Fix(16, 3, device="cuda:0")               # no fraction bits left
    pw = mfcc_b200.PowerSpectrum(device="cuda:0")               # class defaults: width 16, 24 bits out
    got = pw(torch.from_numpy(fvec["power24_in"].astype(np.int16)).to("cuda:0")).cpu().numpy()
    assert np.array_equal(got.view(np.uint32).astype(np.int64), fvec["power24_out"])


# ---- WindowHamming / FftStream / DCTStream at other sizes (rtl_size_vectors.npz) ----
FFT_SIZES = (64, 256, 1024)          # 8 and 16: see test_small_fft_hazard_of_the_reference
DCT_SIZES = (8, 64, 128)
WINDOW_SETS = ((8, 8), (64, 8), (256, 6), (1024, 8), (2048, 10))


@pytest.fixture(scope="module")
def zvec():
    return np.load(os.path.join(GOLD, "rtl_size_vectors.npz"))


def test_oracle_other_sizes_match_rtl(zvec):
    for n in FFT_SIZES:
        re, im = R.fft_stream(zvec["fft%d_in" % n], nfft=n)
        assert np.array_equal(np.stack([re, im], axis=-1), zvec["fft%d_out" % n]), n
    for n in DCT_SIZES:
        assert np.array_equal(R.dct_stream(zvec["dct%d_in" % n], nfft=n), zvec["dct%d_out" % n]), n
    for n, prec in WINDOW_SETS:
        got = R.window(zvec["window%d_%d_in" % (n, prec)], nfft=n, precision=prec)
        assert np.array_equal(got, zvec["window%d_%d_out" % (n, prec)]), (n, prec)


def test_small_fft_hazard_of_the_reference(zvec):
    """Below 32 points the reference's FFT core reads a stage's operands before the previous
    stage has written them (8-cycle butterfly, size/2 cycles per stage): what comes out is
    not a transform.  The cycle-level bank model reproduces the RTL there; the operators
    refuse those sizes instead of imitating the hazard."""
    for n in (8, 16):
        x, want = zvec["fft%d_in" % n], zvec["fft%d_out" % n]
        with pytest.raises(AssertionError):
            R.fft_rtl_banks(x[1], np.zeros_like(x[1]), n)
        for r in range(x.shape[0]):
            re, im = R.fft_rtl_banks(x[r], np.zeros_like(x[r]), n, strict=False)
            assert np.array_equal(s16(re), want[r, :, 0]) and np.array_equal(s16(im), want[r, :, 1])
        re, im = R.fft_stream(x, nfft=n)
        assert not np.array_equal(np.stack([re, im], axis=-1), want)
    x = zvec["fft64_in"][0]
    re, im = R.fft_rtl_banks(x, np.zeros_like(x), 64)           # strict: not even a same-cycle access
    assert np.array_equal(s16(re), zvec["fft64_out"][0, :, 0])
    x = (np.arange(32) * 2971) % 65536 - 32768                  # 32 points: same-cycle accesses through
    re, im = R.fft_rtl_banks(x, np.zeros_like(x), 32, strict=False)   # transparent ports, still the transform
    wr, wi = R.fft_int(x, np.zeros_like(x), 32)
    assert np.array_equal(re, wr[:16]) and np.array_equal(im, wi[:16])


def test_tables_other_sizes():
    """The curve and twiddles the general kernels are given = the oracle's."""
    from mfcc_b200 import tables
    for n, prec in WINDOW_SETS:
        assert np.array_equal(tables.window_curve(n, prec), R.window_curve(n, prec, 2)[1])
    for prec in range(1, 15):                                   # window.py:37 holds for every precision:
        tables.window_curve(64, prec)                           # the quarter table spans 0.46 * 2**(precision+1)
    for n in (8, 64, 4096):
        tw = tables.twiddle_table(n, 16)
        assert tw.shape == (n // 2, 2) and np.abs(tw).max() <= 16384
        assert all(tuple(tw[k]) == tuple(R.twiddle(n, k)) for k in range(0, n // 2, max(1, n // 16)))


@pytest.mark.gpu
def test_cuda_other_sizes_match_rtl(zvec):
    import torch
    import mfcc_b200
    dev = "cuda:0"
    up = lambda a: torch.from_numpy(np.ascontiguousarray(a).astype(np.int16)).to(dev)
    for n in FFT_SIZES:
        got = mfcc_b200.FftStream(16, n, device=dev)(up(zvec["fft%d_in" % n])).cpu().numpy()
        assert np.array_equal(got, zvec["fft%d_out" % n]), n
    for n in DCT_SIZES:
        got = mfcc_b200.DCTStream(16, n, device=dev)(up(zvec["dct%d_in" % n])).cpu().numpy()
        assert np.array_equal(got, zvec["dct%d_out" % n]), n
    for n, prec in WINDOW_SETS:
        got = mfcc_b200.WindowHamming(16, n, prec, device=dev)(up(zvec["window%d_%d_in" % (n, prec)])).cpu().numpy()
        assert np.array_equal(got, zvec["window%d_%d_out" % (n, prec)]), (n, prec)
    # the general kernels against the oracle on many random rows, every size they take
    rng = np.random.default_rng(41)
    for n in (32, 128, 512, 2048, 4096):
        x = rng.integers(-32768, 32768, (300, n))
        re, im = R.fft_stream(x, nfft=n)
        got = mfcc_b200.FftStream(16, n, device=dev)(up(x)).cpu().numpy()
        assert np.array_equal(got, np.stack([re, im], axis=-1)), n
    for n in (8, 16, 32, 256, 1024):
        x = rng.integers(-32768, 32768, (200, n))
        got = mfcc_b200.DCTStream(16, n, device=dev)(up(x)).cpu().numpy()
        assert np.array_equal(got, R.dct_stream(x, nfft=n)), n
    # 512 / 16 / 32: the specialised stage kernels and the general one agree (C ABI directly)
    import ctypes as C
    from mfcc_b200 import _lib, tables
    x = up(rng.integers(-32768, 32768, (64, 512)))
    fast = mfcc_b200.FftStream(16, 512, device=dev)(x)
    tw = up(tables.twiddle_table(512, 16))
    out = torch.empty_like(fast)
    _lib.check(_lib.lib().mfcc_b200_fft_stream(C.c_void_p(x.data_ptr()), C.c_void_p(out.data_ptr()), 64, 512,
                                               C.c_void_p(tw.data_ptr()), None))
    torch.cuda.synchronize()
    assert torch.equal(out, fast)
    with pytest.raises(ValueError):
        mfcc_b200.FftStream(16, 500, device=dev)                # fft.py:351-353
    for bad in (lambda: mfcc_b200.FftStream(16, 16, device=dev), lambda: mfcc_b200.DCTStream(16, 4, device=dev)):
        with pytest.raises(ValueError, match="hazard"):
            bad()
    with pytest.raises(ValueError):
        mfcc_b200.WindowHamming(16, 4, 8, device=dev)           # window.py:13
    with pytest.raises(ValueError):
        mfcc_b200.WindowHamming(16, 96, 8, device=dev)          # the address slices need a power of two


# ---- MFCC with other constructor arguments (rtl_param_vectors.npz) ----
PARAM_SETS = {"256_8k_8_8": (256, 8e3, 8, 8), "1024_16k_64_24": (1024, 16e3, 64, 24),
              "512_44k1_32_13": (512, 44100, 32, 13), "128_16k_16_16": (128, 16e3, 16, 16),
              "512_16k_8_8": (512, 16e3, 8, 8)}
STAGE_KEYS = ("frame", "window", "fft_r", "fft_i", "power", "filter", "log", "dct")


@pytest.fixture(scope="module")
def pvec():
    return np.load(os.path.join(GOLD, "rtl_param_vectors.npz"))


@pytest.mark.parametrize("name", sorted(PARAM_SETS))
def test_oracle_other_parameter_sets_match_rtl(pvec, name):
    nfft, sr, nf, nc = PARAM_SETS[name]
    K = pvec[name + "_out"].shape[0]
    out, t = R.mfcc(pvec[name + "_samples"], nfft=nfft, samplerate=sr, nfilters=nf, nceptrums=nc,
                    n_frames=K, taps=True)
    for k in STAGE_KEYS:
        assert np.array_equal(s16(t[k]), s16(pvec[name + "_" + k])), k
    assert np.array_equal(out, s16(pvec[name + "_out"]))


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(PARAM_SETS))
def test_cuda_other_parameter_sets_match_rtl(pvec, name):
    """MFCC(...) with arguments the fused kernel is not built for runs the chain of general
    stage kernels: every stage against the RTL, batch against the oracle, streaming against
    batch."""
    import torch
    import mfcc_b200
    nfft, sr, nf, nc = PARAM_SETS[name]
    hop = nfft // 3
    dev = "cuda:0"
    op = mfcc_b200.MFCC(nfft=nfft, samplerate=sr, nfilters=nf, nceptrums=nc, device=dev)
    x = pvec[name + "_samples"]
    K = pvec[name + "_out"].shape[0]
    xd = torch.from_numpy(x).to(dev)
    out, t = op(xd, n_frames=K, taps=True)
    got = {k: v.cpu().numpy() for k, v in t.items()}
    got["fft_r"], got["fft_i"] = got["fft"][..., 0], got["fft"][..., 1]
    for k in STAGE_KEYS:
        assert np.array_equal(s16(got[k]), s16(pvec[name + "_" + k])), k
    assert np.array_equal(out.cpu().numpy(), s16(pvec[name + "_out"]))
    # a ragged batch, zero-extended rows, host buffers
    rng = np.random.default_rng(3)
    xb = rng.integers(-32768, 32768, (3, nfft + hop * 6 + 5)).astype(np.int16)
    want = np.stack([R.mfcc(r, nfft=nfft, samplerate=sr, nfilters=nf, nceptrums=nc, n_frames=9) for r in xb])
    assert np.array_equal(op(torch.from_numpy(xb).to(dev), n_frames=9).cpu().numpy(), want)
    assert np.array_equal(op.run_host(torch.from_numpy(xb), n_frames=9).numpy(), want)
    assert op(torch.from_numpy(xb).to(dev)).shape == (3, 7, nc)
    # streaming: hop by hop from reset = the batch path on the stream with nfft - hop zeros in front
    op.reset(3)
    lead = np.zeros((3, nfft - hop), dtype=np.int16)
    full = np.concatenate([lead, xb[:, :hop * 5]], axis=1)
    wb = np.stack([R.mfcc(r, nfft=nfft, samplerate=sr, nfilters=nf, nceptrums=nc) for r in full])
    for j in range(5):
        col = op.step(torch.from_numpy(np.ascontiguousarray(xb[:, hop * j:hop * (j + 1)])).to(dev))
        assert np.array_equal(col.cpu().numpy(), wb[:, j]), j


@pytest.mark.gpu
def test_cuda_parameter_limits():
    import mfcc_b200
    with pytest.raises(ValueError, match="hazard"):
        mfcc_b200.MFCC(nfft=16, nfilters=8, nceptrums=8, device="cuda:0")
    with pytest.raises(ValueError, match="hazard"):
        mfcc_b200.MFCC(nfft=256, nfilters=4, nceptrums=4, device="cuda:0")
    with pytest.raises(ValueError):
        mfcc_b200.MFCC(nfft=500, device="cuda:0")                 # fft.py:351-353
    with pytest.raises(ValueError):
        mfcc_b200.MFCC(nfft=256, mode="fp32", device="cuda:0")
