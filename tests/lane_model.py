"""Lane-level model of the fused CUDA kernel (mfcc_b200/csrc/mfcc_kernels.cu).

One warp (32 lanes) computes one frame.  This file mirrors the kernel's data
movement step by step in numpy (registers = arrays indexed [lane, slot], shared
memory = flat arrays, shuffles = fancy indexing) so that the index arithmetic of
the design -- bit-reversed loads, the two shared-memory exchanges and their bank
skews, per-lane twiddle indices, the shuffle plan of the DCT -- can be checked
against the oracle on a machine without a GPU (tests/test_lane_model.py).

It is a test aid, not product code.
"""
from __future__ import annotations

import numpy as np

from mfcc_b200 import tables as T

LANES = 32


def brev(v, bits):
    r = 0
    for _ in range(bits):
        r = (r << 1) | (v & 1)
        v >>= 1
    return r


def s16(v):
    return ((np.asarray(v, dtype=np.int64) + 32768) & 0xFFFF) - 32768


def s32(v):
    return ((np.asarray(v, dtype=np.int64) + (1 << 31)) & 0xFFFFFFFF) - (1 << 31)


def bfly(x0r, x0i, x1r, x1i, wr, wi):
    """The kernel's butterfly: everything doubled so that the result sits in the
    top half of a 32-bit word and the 32-bit wrap IS the 16-bit wrap of the RTL."""
    w2r, w2i = 2 * wr, 2 * wi
    qr = s32(x1r * w2r - x1i * w2i + 16382)
    qi = s32(x1i * w2r + x1r * w2i + 16382)
    br = s32(x0r * 32768)
    bi = s32(x0i * 32768)
    y0r = s32(br + qr) >> 16
    y0i = s32(bi + qi) >> 16
    y1r = s32(br - qr + 32766) >> 16
    y1i = s32(bi - qi + 32766) >> 16
    return y0r, y0i, y1r, y1i


def bfly_top_half(S0r, S0i, S1r, S1i, wr, wi):
    """The butterfly on top-half-form operands (mfcc_dev.cuh: bf_d / bf_s): the value sits in
    bits 16..31 of a 32-bit word, the low half is don't-care.  d = (x0 << 16) + 32767 is one
    mask-and-or, c = d >> 1 = (x0 << 15) + 16383, s0' = c + 2P (= s0 + 1: s0 is even, so its top
    half does not change), s1 = d - s0'.  Returns the four sums as unsigned 32-bit words (results
    in their top halves)."""
    U = 0xFFFFFFFF
    S0r, S0i, S1r, S1i = (np.asarray(v, dtype=np.int64) & U for v in (S0r, S0i, S1r, S1i))
    x1r, x1i = s32(S1r) >> 16, s32(S1i) >> 16
    dr, di = (S0r & 0xFFFF0000) | 0x7FFF, (S0i & 0xFFFF0000) | 0x7FFF
    cr, ci = s32(dr) >> 1, s32(di) >> 1
    w2r, w2i = 2 * wr, 2 * wi
    s0r = (cr + x1r * w2r - x1i * w2i) & U
    s0i = (ci + x1i * w2r + x1r * w2i) & U
    s1r = (dr - s0r) & U
    s1i = (di - s0i) & U
    return s0r, s0i, s1r, s1i


# ---- exchange address maps (32-bit words inside the per-warp buffer) -------------
def xa_addr(row, col):
    """exchange 1: row = lane of pass A (16 positions each), col = slot."""
    return row * 20 + col + 16 * (row >> 4)


def xb_addr(p):
    """exchange 2: natural position p in 0..511."""
    return p + 16 * (p >> 8)


def bank_conflict_degree(addrs):
    """max number of distinct words hitting one bank for a warp-wide 32-bit access."""
    addrs = np.asarray(addrs)
    worst = 1
    for b in range(32):
        words = np.unique(addrs[(addrs % 32) == b])
        worst = max(worst, len(words))
    return worst


def fft512_frame(w, tw):
    """w: 512 windowed samples (int).  tw: decoded twiddles [256, 2].
    Returns (re, im) of bins 0..255 and the bank-conflict report."""
    lane = np.arange(LANES)
    conflicts = {}
    # ---- load: slot l of lane t holds position p = 16 t + l = sample brev9(p)
    xr = np.zeros((LANES, 16), dtype=np.int64)
    xi = np.zeros((LANES, 16), dtype=np.int64)
    for l in range(16):
        n = np.array([brev(int(t), 5) for t in lane]) + 32 * brev(l, 4)
        assert all(n[t] == brev(16 * t + l, 9) for t in lane)
        xr[:, l] = w[n]
        conflicts["load%d" % l] = bank_conflict_degree(n // 2)
    # ---- pass A: stages 0..3 inside the lane, compile-time twiddles
    for s in range(4):
        half = 1 << s
        for l in range(16):
            if l & half:
                continue
            k = l % half
            wr, wi = tw[k * (256 >> s)]
            a, b, c, d = bfly(xr[:, l], xi[:, l], xr[:, l + half], xi[:, l + half], wr, wi)
            xr[:, l], xi[:, l], xr[:, l + half], xi[:, l + half] = a, b, c, d
    # ---- exchange 1
    buf_r = np.zeros(700, dtype=np.int64)
    buf_i = np.zeros(700, dtype=np.int64)
    for l in range(16):
        a = xa_addr(lane, l)
        buf_r[a], buf_i[a] = xr[:, l], xi[:, l]
    for q in range(4):   # 128-bit stores are issued per quarter warp
        for c in range(4):
            a = xa_addr(lane[8 * c: 8 * c + 8], 4 * q)
            banks = np.concatenate([(a + i) % 32 for i in range(4)])
            assert len(np.unique(banks)) == 32, "STS.128 conflict"
    h = lane >> 4
    l4 = lane & 15
    for m in range(16):
        a = xa_addr(16 * h + m, l4)
        conflicts["xa_read%d" % m] = bank_conflict_degree(a)
        xr[:, m], xi[:, m] = buf_r[a], buf_i[a]
    # lane (h, l4) now holds p = 256 h + 16 m + l4 in slot m
    # ---- pass B: stages 4..7, twiddle index (p mod 2^s) * (256 >> s)
    for s in range(4, 8):
        hm = 1 << (s - 4)
        for m in range(16):
            if m & hm:
                continue
            k = (m & (hm - 1)) * 16 + l4
            idx = k * (256 >> s)
            wr, wi = tw[idx, 0], tw[idx, 1]
            a, b, c, d = bfly(xr[:, m], xi[:, m], xr[:, m + hm], xi[:, m + hm], wr, wi)
            xr[:, m], xi[:, m], xr[:, m + hm], xi[:, m + hm] = a, b, c, d
    # ---- exchange 2
    for m in range(16):
        p = 256 * h + 16 * m + l4
        a = xb_addr(p)
        conflicts["xb_write%d" % m] = bank_conflict_degree(a)
        buf_r[a], buf_i[a] = xr[:, m], xi[:, m]
    re = np.zeros(256, dtype=np.int64)
    im = np.zeros(256, dtype=np.int64)
    for i in range(8):
        b = lane + 32 * i
        a0, a1 = xb_addr(b), xb_addr(b + 256)
        conflicts["xb_read%d" % i] = max(bank_conflict_degree(a0), bank_conflict_degree(a1))
        wr, wi = tw[b, 0], tw[b, 1]
        y0r, y0i, _, _ = bfly(buf_r[a0], buf_i[a0], buf_r[a1], buf_i[a1], wr, wi)
        re[b], im[b] = y0r, y0i
    return re, im, conflicts


def bf1_rr(x0, x1):
    """kernel's W^0 butterfly on real operands (mfcc_dev.cuh: bf1_rr)."""
    return (x0 + x1) >> 1, (x0 - x1) >> 1


def bfj_cc(x0r, x0i, x1r, x1i):
    """kernel's -j butterfly (mfcc_dev.cuh: bfj_cc / bfj_rr with x0i = x1i = 0)."""
    tr, ti = x1i, -x1r
    return (x0r + tr) >> 1, (x0i + ti) >> 1, (x0r - tr) >> 1, (x0i - ti) >> 1


def mel_tables(nfilters, samplerate=16e3, nfft=512):
    """What mfcc_b200_create() builds for mel_warp(): per lane (segment lane+1) and step
    the weight pair (wa, wb), zero past the segment's end; segment 0 separately."""
    w, seg, close = T.mel_weights(nfilters, nfft, samplerate)
    bounds = T.mel_segment_bounds(nfilters, nfft, samplerate)
    maxseg = 21 if nfilters == 32 else 42
    assert len(bounds) == nfilters + 1
    ab = np.zeros((maxseg, nfilters, 2), dtype=np.int64)
    len0 = bounds[0][1] - bounds[0][0] + 1
    for g, (first, cl) in enumerate(bounds):
        n = cl - first + 1
        j0 = len0 if g == 1 else 0
        assert j0 + n <= maxseg
        for j in range(n):
            closing = first + j == cl
            wa = (1 << 30) if closing else int(w[first + j])
            wb = 0 if closing else (1 << 30) - int(w[first + j])
            if g == 0:
                ab[j, 0] = (0, wa)           # Abar_0 feeds filter 0: kept in lane 0's wb slot
            else:
                ab[j0 + j, g - 1] = (wa, wb)
    return ab, bounds


def mel_padded_layout(nfilters, samplerate=16e3, nfft=512):
    """DevTables::paddr / seg_walk as mfcc_b200_create() builds them: segment g starts at
    the first free word whose bank no earlier segment start took (segment 1 follows
    segment 0 directly: lane 0 walks both)."""
    bounds = T.mel_segment_bounds(nfilters, nfft, samplerate)
    nl = 32 if nfilters == 32 else 16
    used, nxt = set(), 0
    paddr = np.zeros(nfft // 2, dtype=np.int64)
    walk = np.zeros(nfilters, dtype=np.int64)
    for g, (first, cl) in enumerate(bounds):
        base = nxt
        if g != 1:
            while base % nl in used:
                base += 1
            used.add(base % nl)
        n = cl - first + 1
        paddr[first:first + n] = base + np.arange(n)
        nxt = base + n
        if g == 0:
            walk[0] = base
        elif g > 1:
            walk[g - 1] = base
    return paddr, walk, nxt


def mel_lanes(pw, nfilters, samplerate=16e3, nfft=512, lo_bit=31):
    """Lane g-1 walks segment g (lane 0: segments 0 and 1) with two 64-bit multiply-adds
    per step (fixed trip count, zero-padded weights); filter g-1 = Abar_{g-1} (from the
    neighbour) + Bbar_g."""
    ab, bounds = mel_tables(nfilters, samplerate, nfft)
    maxseg = ab.shape[0]
    M64 = (1 << 64) - 1
    paddr, walk, end = mel_padded_layout(nfilters, samplerate, nfft)
    pwx = np.full(end + maxseg + 8, 12345, dtype=np.int64)      # gaps and overrun hold garbage
    pwx[paddr] = np.asarray(pw, dtype=np.int64)
    a = [0] * nfilters
    b = [0] * nfilters
    for li in range(nfilters):
        first = int(walk[li])
        for j in range(maxseg):
            d = int(pwx[first + j])
            a[li] = (a[li] + d * int(ab[j, li, 0])) & M64
            b[li] = (b[li] + d * int(ab[j, li, 1])) & M64
    out = np.zeros(nfilters, dtype=np.int64)
    for li in range(nfilters):
        pa = 0 if li == 0 else a[li - 1]                   # shuffle up
        out[li] = (((pa + b[li]) & M64) >> lo_bit) & 0xFFFF
    return out


def mel_lanes_onewide(pw, nfilters, samplerate=16e3, nfft=512, lo_bit=31):
    """mel_warp() in its one-wide-multiply form (32 filters): one 64-bit multiply-add and one 32-bit
    one per step.  With wa the weight of a bin (2^30 on the closing bin) and nc = 1 on every bin of the
    segment (0 on the zero padding), a = sum d*wa (64-bit), D = sum d*nc (32-bit, wraps):
        Bbar_g = sum_{non-closing} d (2^30 - w) = ((D_nc + PK) << 30) - (A' + PK << 30) = (D << 30) - a
    so the register the hardware slices is  Abar_{g-1} + Bbar_g = a_{g-1} + (D_g << 30) - a_g, of which
    only bits lo_bit .. lo_bit+15 are kept: (D mod 2^32) << 30 is exact modulo 2^62.  Lane 0 walks
    segment 0 first: usable when every non-closing weight of segment 0 is zero (true for the 32-filter
    bank: segment 0 is bins 0..1 and bin 0 has weight 0), because then Abar_0 = PK_0 << 30 and the closing
    bin of segment 0 simply counts into lane 0's D.  Returns None when the bank does not allow it."""
    ab, bounds = mel_tables(nfilters, samplerate, nfft)
    maxseg = ab.shape[0]
    M64, M32 = (1 << 64) - 1, (1 << 32) - 1
    paddr, walk, end = mel_padded_layout(nfilters, samplerate, nfft)
    pwx = np.full(end + maxseg + 8, 12345, dtype=np.int64)
    pwx[paddr] = np.asarray(pw, dtype=np.int64)
    len0 = bounds[0][1] - bounds[0][0] + 1
    if any(int(ab[j, 0, 1]) != 0 for j in range(len0 - 1)) or int(ab[len0 - 1, 0, 1]) != 1 << 30:
        return None
    a, D = [0] * nfilters, [0] * nfilters
    for li in range(nfilters):
        first = int(walk[li])
        seglen = bounds[li + 1][1] - bounds[li + 1][0] + 1
        for j in range(maxseg):
            d = int(pwx[first + j])
            if li == 0 and j < len0:
                wa, nc = 0, (1 if j == len0 - 1 else 0)       # segment 0: only its closing bin counts, as d << 30
            else:
                jj = j - (len0 if li == 0 else 0)
                inside = 0 <= jj < seglen
                wa = int(ab[j, li, 0]) if inside else 0
                nc = 1 if inside else 0
            a[li] = (a[li] + d * wa) & M64
            D[li] = (D[li] + d * nc) & M32
    out = np.zeros(nfilters, dtype=np.int64)
    for li in range(nfilters):
        pa = 0 if li == 0 else a[li - 1]
        out[li] = (((pa + (D[li] << 30) - a[li]) & M64) >> lo_bit) & 0xFFFF
    return out


def log2_table():
    """logfrac[] of DevTables: the ten squarings for every 12-bit mantissa."""
    t = np.zeros(2048, dtype=np.int64)
    for z0 in range(2048, 4096):
        z, frac = z0, 0
        for _ in range(10):
            c = z * z
            bit = (c >> 23) & 1
            z = c >> (11 + bit)
            frac = (frac << 1) | bit
        t[z0 - 2048] = frac
    return t


def log2_tab_lane(m, table):
    mm = max(int(m), 1)
    sh = mm.bit_length() - 1
    z = (mm << 11) >> sh
    return ((sh << 11) + (int(table[z - 2048]) << 1)) & 0x7FFF


def log2_lane(m, prec=11):
    """clz-based integer part, then `prec-1` squarings."""
    m = int(m)
    mm = max(m, 1)
    sh = mm.bit_length() - 1
    z = (mm << prec) >> sh
    frac = 0
    for _ in range(prec - 1):
        c = z * z
        bit = (c >> (2 * prec + 1)) & 1
        z = c >> (prec + bit)
        frac = (frac << 1) | bit
    return ((sh << prec) + (frac << 1)) & 0x7FFF


def dct_plan(nf):
    """Ownership of positions by lanes through the 2*nf-point sub-FFT (see
    DESIGN.md, "DCT").  Returns per stage s>=1 the twiddle ROM index of each lane's
    butterfly and, at the end, the bin each lane owns in slot 0."""
    S = (2 * nf).bit_length() - 1          # stages of the 2nf-point FFT
    lanes = np.arange(nf)
    A = 2 * lanes                          # slot0 position; slot1 = A + 1
    tw_idx = np.zeros((S, nf), dtype=np.int64)
    for s in range(1, S):
        bit = (lanes >> (s - 1)) & 1
        partner = lanes ^ (1 << (s - 1))
        # lane with bit 0 takes (A, A + 2^s); lane with bit 1 takes (A_p + 2^(s-1), .. + 2^s)
        assert np.all(A[partner][bit == 0] == A[bit == 0] + (1 << s))
        x0pos = np.where(bit == 0, A, A[partner] + (1 << (s - 1)))
        k = x0pos % (1 << s)
        tw_idx[s] = k * ((2 * nf) >> s)    # ROM of the 4nf-point FFT has 2nf entries
        A = x0pos
    return tw_idx, A


def dct_lanes(L, nf, tw):
    """L: nf log values (lane g holds L[g]).  tw: decoded ROM [2nf, 2] of the 4nf FFT."""
    S = (2 * nf).bit_length() - 1
    nb = nf.bit_length() - 1
    lanes = np.arange(nf)
    tw_idx, own = dct_plan(nf)
    src = np.array([brev(int(j), nb) for j in lanes])
    a = np.asarray(L, dtype=np.int64)[src]                # shuffle
    b = np.asarray(L, dtype=np.int64)[nf - 1 - src]       # shuffle
    z = np.zeros(nf, dtype=np.int64)
    r0, i0, r1, i1 = bfly(a, z, b, z, 16384, 0)           # stage 0
    for s in range(1, S):
        bit = (lanes >> (s - 1)) & 1
        partner = lanes ^ (1 << (s - 1))
        send_r = np.where(bit == 0, r1, r0)
        send_i = np.where(bit == 0, i1, i0)
        rc_r, rc_i = send_r[partner], send_i[partner]     # shuffle xor
        x0r = np.where(bit == 0, r0, rc_r)
        x0i = np.where(bit == 0, i0, rc_i)
        x1r = np.where(bit == 0, rc_r, r1)
        x1i = np.where(bit == 0, rc_i, i1)
        wr, wi = tw[tw_idx[s], 0], tw[tw_idx[s], 1]
        r0, i0, r1, i1 = bfly(x0r, x0i, x1r, x1i, wr, wi)
    # final stage of the 4nf FFT: even half is all zero, bins own[j] < nf
    assert sorted(own.tolist()) == list(range(nf))
    wr, wi = tw[own, 0], tw[own, 1]
    y, _, _, _ = bfly(z, z, r0, i0, wr, wi)
    out = np.zeros(nf, dtype=np.int64)
    out[own] = y
    return out
