"""CPU check of the kernel's design: tests/lane_model.py mirrors the CUDA warp
algorithm (index maps, skews, per-lane twiddles, DCT shuffle plan); here it must
reproduce the oracle bit for bit and its shared-memory accesses must be conflict free."""
import math

import numpy as np
import pytest

import lane_model as LM
from mfcc_b200 import tables as T
from oracle import ref_np as R


def test_fft512_lanes_bit_exact_and_conflict_free():
    rng = np.random.default_rng(3)
    tw = T.twiddle_table(512).astype(np.int64)
    frames = [rng.integers(-32768, 32768, 512), np.tile([-32768, 32767], 256),
              np.full(512, -32768), rng.integers(-3000, 3000, 512)]
    for w in frames:
        re, im, conf = LM.fft512_frame(w, tw)
        rr, ri = R.fft_stream(w[None, :])
        assert np.array_equal(re, rr[0]) and np.array_equal(im, ri[0])
        assert max(conf.values()) == 1, {k: v for k, v in conf.items() if v > 1}


@pytest.mark.parametrize("nf", [16, 32])
def test_epilogue_lanes(nf, wav):
    _, tp = R.mfcc(wav[:5000], nfilters=nf, nceptrums=nf, taps=True)
    pts = R.get_filter_points(0, 8000.0, nf, 512, sample_rate=16e3)
    lo = int(math.log2(pts[-1] - pts[-3])) + 60 - 34
    twd = T.twiddle_table(4 * nf).astype(np.int64)
    for k in range(tp["power"].shape[0]):
        mel = LM.mel_lanes(tp["power"][k], nf, lo_bit=lo)
        assert np.array_equal(mel, tp["filter"][k])
        lg = np.array([LM.log2_lane(m) for m in mel])
        assert np.array_equal(lg, tp["log"][k])
        assert np.array_equal(LM.dct_lanes(lg, nf, twd), tp["dct"][k])
    rng = np.random.default_rng(nf)
    L = rng.integers(0, 32768, (16, nf))
    dd = R.dct_stream(L, nf)
    for k in range(len(L)):
        assert np.array_equal(LM.dct_lanes(L[k], nf, twd), dd[k])
    tw_idx, own = LM.dct_plan(nf)
    assert own.tolist() == list(range(nf))          # lane j ends up with bin j


def test_log2_lane_all_values():
    m = np.arange(65536)
    want = R.log2fix(m[::17], 16, 15)
    got = np.array([LM.log2_lane(v) for v in m[::17]])
    assert np.array_equal(got, want)


def test_trivial_butterflies_equal_the_general_one():
    """bf1_rr / bfj_* (used for W^0 and -j in stages 0-3) are the general butterfly with
    those twiddles, for every combination of extreme and random 16-bit operands."""
    rng = np.random.default_rng(3)
    edge = np.array([-32768, -32767, -1, 0, 1, 2, 32766, 32767])
    v = np.concatenate([edge, rng.integers(-32768, 32768, 200)])
    x0r, x0i, x1r, x1i = [a.ravel() for a in np.meshgrid(v[:24], v[:24], v[:24], v[:24])]
    z = np.zeros_like(x0r)
    a, b, c, d = LM.bfly(x0r, z, x1r, z, 16384, 0)
    y0, y1 = LM.bf1_rr(x0r, x1r)
    assert np.array_equal(a, y0) and np.array_equal(c, y1) and not b.any() and not d.any()
    want = LM.bfly(x0r, x0i, x1r, x1i, 0, -16384)
    got = LM.bfj_cc(x0r, x0i, x1r, x1i)
    for w_, g_ in zip(want, got):
        assert np.array_equal(w_, LM.s16(g_)) and np.array_equal(g_, LM.s16(g_))


def test_top_half_butterfly_equals_the_rtl_butterfly():
    """bf_d / bf_s of mfcc_dev.cuh against the oracle's restatement of Butterfly
    (mfcc/misc/fft.py:93-96, :188-191): every twiddle of the 512-point ROM, operands drawn from the
    whole int16 range including the corners where the 16-bit wrap triggers, arbitrary junk in the
    low halves of the operand words."""
    from oracle import ref_np as R
    rng = np.random.default_rng(5)
    n = 20000
    corners = np.array([-32768, -32767, -1, 0, 1, 32766, 32767])
    for _ in range(4):
        x = [np.concatenate([rng.integers(-32768, 32768, n), rng.choice(corners, n)]) for _ in range(4)]
        k = rng.integers(0, 256, 2 * n)
        tw = np.array([R.twiddle(512, int(i)) for i in range(256)])
        wr, wi = tw[k, 0], tw[k, 1]
        junk = [rng.integers(0, 65536, 2 * n) for _ in range(4)]
        S = [((v & 0xFFFF) << 16) | j for v, j in zip(x, junk)]
        s0r, s0i, s1r, s1i = LM.bfly_top_half(S[0], S[1], S[2], S[3], wr, wi)
        got = [LM.s32(v) >> 16 for v in (s0r, s0i, s1r, s1i)]
        want = LM.bfly(x[0], x[1], x[2], x[3], wr, wi)
        for g, w in zip(got, want):
            assert np.array_equal(g, w)
        # and against the RTL restatement itself on a subset
        for i in range(0, 2 * n, 997):
            y = R.butterfly(int(x[0][i]), int(x[1][i]), int(x[2][i]), int(x[3][i]), int(wr[i]), int(wi[i]))
            assert tuple(int(v) for v in y) == tuple(int(g[i]) for g in got)


def test_mel_one_wide_multiply_form_equals_the_two_sum_form(nf=32, lo=31):
    """mel_warp's MFCC_MEL_ONEWIDE form (one 64-bit and one 32-bit multiply-add per bin) against the
    two-sum form that test_oracle.py pins to the FilterBank restatement: random spectra, full-scale
    spectra (every 30-bit power value at its maximum: the 16-bit output wraps), single bins."""
    rng = np.random.default_rng(8)
    top = (1 << 30) - 1
    cases = [rng.integers(0, 1 << 30, 256) for _ in range(20)] + [rng.integers(0, 5000, 256) for _ in range(5)]
    cases += [np.full(256, top), np.zeros(256, dtype=np.int64)]
    for b in (0, 1, 2, 3, 127, 128, 235, 236, 254, 255):
        e = np.zeros(256, dtype=np.int64)
        e[b] = top
        cases.append(e)
    for pw in cases:
        assert np.array_equal(LM.mel_lanes_onewide(pw, nf, lo_bit=lo), LM.mel_lanes(pw, nf, lo_bit=lo))
    # the 16-filter bank has a non-zero rising weight inside segment 0: that form does not apply (two-sum form stays)
    assert LM.mel_lanes_onewide(cases[0], 16, lo_bit=32) is None


def test_log2_table_equals_the_loop():
    t = LM.log2_table()
    for m in list(range(0, 4200)) + list(range(60000, 65536)) + [32767, 32768, 16384, 9999]:
        assert LM.log2_tab_lane(m, t) == LM.log2_lane(m)


@pytest.mark.parametrize("nf", [16, 32])
def test_mel_padded_layout_is_conflict_free(nf):
    paddr, walk, end = LM.mel_padded_layout(nf)
    maxseg = 21 if nf == 32 else 42
    assert len(set(paddr.tolist())) == 256 and end + maxseg <= 656        # injective, fits xbuf
    lanes = nf if nf == 32 else 16
    for j in range(maxseg if nf == 32 else maxseg // 2):
        assert LM.bank_conflict_degree(walk[:lanes] + j) == 1
    # the power stores (lane handles bins lane + 32 i) stay cheap
    worst = max(LM.bank_conflict_degree(paddr[np.arange(32) + 32 * i]) for i in range(8))
    assert worst <= 2


# ---- fp32 mode: index plan of power_pair_f32 (mfcc_f32.cuh) ------------------------------
def test_f32_half_warp_plan():
    """The 16 x 16 decomposition of the packed 256-point transform, the register order fft16
    leaves (k16 / j16), the transposed exchange and the mirror-lane pairing of the real-input
    split (lane l takes its k1 = 0..7 and receives registers 8..15 of lane 16 - l; lane 0 mirrors
    itself one register further round), emulated lane by lane against numpy's FFT."""
    def brev3(v):
        return ((v & 1) << 2) | (v & 2) | ((v & 4) >> 2)

    def k16(j):
        return 2 * brev3(j) if j < 8 else 2 * brev3(j - 8) + 1

    def j16(k):
        return 8 + brev3(k >> 1) if k & 1 else brev3(k >> 1)

    def fft8(v):
        X = np.fft.fft(np.array(v))
        return [X[brev3(j)] for j in range(8)]

    def fft16(v):
        e = [v[i] + v[i + 8] for i in range(8)]
        o = [(v[i] - v[i + 8]) * np.exp(-2j * np.pi * i / 16) for i in range(8)]
        return fft8(e) + fft8(o)

    assert all(k16(j16(k)) == k for k in range(16))
    rng=np.random.default_rng(0)
    y=rng.standard_normal(512)
    z=y[0::2]+1j*y[1::2]
    # pass 1
    T=np.zeros((16,16),complex)
    for l in range(16):
        v=[z[l+16*a] for a in range(16)]
        v=fft16(v)
        for j in range(16):
            k2=k16(j)
            T[k2][l]=v[j]*np.exp(-2j*np.pi*l*k2/256)
    V=[]
    for l in range(16):
        v=[T[l][n1] for n1 in range(16)]
        V.append(fft16(v))
    Z=np.fft.fft(z)
    for l in range(16):
        for k1 in range(16):
            assert abs(V[l][j16(k1)]-Z[16*k1+l])<1e-9
    # mirror + split
    X=np.fft.fft(y)[:257]
    P=np.zeros(257)
    for l in range(16):
        src=(16-l)&15
        r=[]
        for i in range(8):
            # value that lane src sends
            ls=src
            s=V[ls][j16(8+i)] if ls!=0 else V[ls][j16((9+i)&15)]
            r.append(s)
        for k1 in range(8):
            k=16*k1+l
            w=np.exp(-2j*np.pi*k/512)
            A=V[l][j16(k1)]; B=r[7-k1]
            F=0.5*(A+np.conj(B)); G=(A-np.conj(B))/(2j)
            X0=F+w*G; X1=np.conj(F-w*G)
            P[k]=abs(X0)**2; P[256-k]=abs(X1)**2
        if l==0:
            A=V[0][j16(8)]; P[128]=abs(A)**2
    assert np.abs(P - np.abs(X) ** 2).max() < 1e-9
