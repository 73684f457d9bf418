// Device-side building blocks of the B200 MFCC path (integer mode).
//
// One warp computes one frame.  The 512-point FFT runs as 4 + 4 + 1 radix-2 stages:
// stages 0-3 and 4-7 entirely in registers (16 complex values per lane) with two
// shared-memory exchanges in between, stage 8 fused with the power spectrum.
// Every radix-2 stage keeps the reference's rounding: the butterfly below yields,
// bit for bit, Butterfly(width=16, bias_width=14, scale_bit=1) of
// /root/reference/mfcc/misc/fft.py:64-194.
//
// tests/lane_model.py mirrors this file step by step in numpy.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace mfccb {

constexpr int NFFT   = 512;
constexpr int HOP    = 170;      // nfft // 3, mfcc/core/mfcc.py:43
constexpr int NBINS  = 256;      // FftStream emits bins 0..255, fft_stream.py:24,28
constexpr int XW     = 656;      // 32-bit words of the per-warp exchange buffer
constexpr unsigned FULL = 0xffffffffu;

// Tables every CTA copies from global to shared memory once (16-byte multiples).
struct alignas(16) DevTables {
    int2     twB[128];       // stages 4..7, doubled (2re, 2im): [16 | 16 | 32 | 64] entries
    int2     tw8[256];       // stage 8, doubled, natural order
    uint2    melab[704];     // FilterBank weights, padded per segment: see mel_warp()
    uint32_t pad0[4];
    alignas(16) uint16_t logfrac[2048];  // Log2Fix fraction bits 10..1 for every 12-bit mantissa: see log2fix_tab()
    int2     dct_tw[6 * 32]; // DCT sub-FFT, stage s lane j at [s*32+j], doubled
    int2     dct_twf[32];    // final DCT stage, lane j = bin j, doubled
    // window multipliers (9 bits, unsigned), window.py:59-115, as each lane of frame_warp needs them:
    // [q][lane] = the multipliers of positions brev5(lane) + 32 brev4(l), l = 4q..4q+3 (curve_at() for position n)
    uint32_t curveL[4 * 32 * 4];
    int16_t  seg_walk[40];   // word address (padded power layout) where lane's walk starts:
                             // segment lane+1, lane 0: segments 0 and 1
    int16_t  seg_close[40];  // closing bin of mel segment g
    uint16_t paddr[256];     // word address of bin b in the padded power layout: every segment
                             // starts in a bank no other segment start uses, so the 32 lanes of
                             // mel_warp() read 32 different banks at every step
    uint4    pstore[2 * 32]; // [i / 4][lane]: paddr[lane + 32 i], i = 0..7: where the lane's power values go
    int32_t  mel_lo_bit;     // filterbank.py:122
    int32_t  nseg;
    int32_t  mel_onewide;    // 32 filters: melab holds (wa, nc) pairs, see mel_warp()
    int32_t  pad[1];
};
static_assert(sizeof(DevTables) % 16 == 0, "DevTables must be int4-copyable");

// W512^(32k), k = 0..7: the only twiddles stages 0..3 need (fft.py:28-36, :43-59).
// mfcc_b200_create() checks them against the ROM the host passes in.
__device__ constexpr int TWA_RE[8] = {16384, 15137, 11585, 6270, 0, -6270, -11585, -15137};
__device__ constexpr int TWA_IM[8] = {0, -6270, -11585, -15137, -16384, -15137, -11585, -6270};
constexpr int H_TWA_RE[8] = {16384, 15137, 11585, 6270, 0, -6270, -11585, -15137};
constexpr int H_TWA_IM[8] = {0, -6270, -11585, -15137, -16384, -15137, -11585, -6270};

__host__ __device__ constexpr int brev4(int v) {
    return ((v & 1) << 3) | ((v & 2) << 1) | ((v & 4) >> 1) | ((v & 8) >> 3);
}
__host__ __device__ constexpr int brev5(int v) {
    return ((v & 1) << 4) | ((v & 2) << 2) | (v & 4) | ((v & 8) >> 2) | ((v & 16) >> 4);
}
// window multiplier of sample position n (0..511) in either table layout
__host__ __device__ __forceinline__ int curve_index(int n) {       // index into DevTables::curveL
    const int lane = brev5(n & 31), l = brev4(n >> 5);
    return ((l >> 2) * 32 + lane) * 4 + (l & 3);
}

// ---------------------------------------------------------------------------------
// Butterfly.  With P = x1*w (complex, 30-bit products) the RTL computes
//   t  = (P + 8191) >> 14
//   y0 = wrap16((x0 + t) >> 1),  y1 = wrap16((x0 - t) >> 1)        fft.py:93-96, :188-191
// (x0 + t) >> 1 = ((x0 << 14) + P + 8191) >> 15, so y0 is bits 15..30 of that sum.
// Doubling everything puts those bits in the top half of a 32-bit register:
//   s0 = (x0 << 15) + 2P + 16382,   y0 = s0 >> 16 (arithmetic)
// and the natural 32-bit wrap of s0 *is* the 16-bit wrap of the hardware.  Likewise
//   x0 - t = ((x0 << 14) - P + 8192) >> 14   (floor of a negated floor)
//   s1 = (x0 << 15) - 2P + 16384 = (x0 << 15) - q + 32766 with q = 2P + 16382.
// w2r, w2i are the doubled twiddle components.
struct Bf { unsigned s0r, s0i, s1r, s1i; };

// nw2i = -w2i is passed in so that the subtraction of the imaginary product is a
// multiply-add too (twiddles are table entries or compile-time constants: free).
__device__ __forceinline__ Bf bf_core(int x0r, int x0i, int x1r, int x1i, int w2r, int w2i, int nw2i) {
    unsigned qr = (unsigned)(x1r * w2r) + 16382u;
    qr += (unsigned)(x1i * nw2i);
    unsigned qi = (unsigned)(x1i * w2r) + 16382u;
    qi += (unsigned)(x1r * w2i);
    Bf o;
    o.s0r = (unsigned)(x0r * 32768) + qr;
    o.s0i = (unsigned)(x0i * 32768) + qi;
    o.s1r = (unsigned)(x0r * 32768) + (32766u - qr);
    o.s1i = (unsigned)(x0i * 32768) + (32766u - qi);
    return o;
}

// in-place, unpacked (sign-extended 16-bit values in 32-bit registers)
__device__ __forceinline__ void bf(int &x0r, int &x0i, int &x1r, int &x1i, int w2r, int w2i, int nw2i) {
    Bf o = bf_core(x0r, x0i, x1r, x1i, w2r, w2i, nw2i);
    x0r = (int)o.s0r >> 16; x0i = (int)o.s0i >> 16;
    x1r = (int)o.s1r >> 16; x1i = (int)o.s1i >> 16;
}
__device__ __forceinline__ void bf(int &x0r, int &x0i, int &x1r, int &x1i, int w2r, int w2i) {
    bf(x0r, x0i, x1r, x1i, w2r, w2i, -w2i);
}

// packed outputs (re in the low half, im in the high half): one PRMT per value
__device__ __forceinline__ void bf_pk(int x0r, int x0i, int x1r, int x1i, int w2r, int w2i, int nw2i,
                                      unsigned &p0, unsigned &p1) {
    Bf o = bf_core(x0r, x0i, x1r, x1i, w2r, w2i, nw2i);
    p0 = __byte_perm(o.s0r, o.s0i, 0x7632);
    p1 = __byte_perm(o.s1r, o.s1i, 0x7632);
}
__device__ __forceinline__ void bf_pk(int x0r, int x0i, int x1r, int x1i, int w2r, int w2i,
                                      unsigned &p0, unsigned &p1) {
    bf_pk(x0r, x0i, x1r, x1i, w2r, w2i, -w2i, p0, p1);
}

// ---------------------------------------------------------------------------------
// The same butterfly on operands in "top-half form": the 16-bit value sits in bits 16..31 of a
// 32-bit word and the low half is don't-care -- which is what the sums s0, s1 above already are,
// and what the imaginary part of a packed (re | im << 16) exchange word is.
// (x0 << 15) + 2P is even, so the bias may be 16383 instead of 16382 without changing bit 16 and
// up of s0 (an even number plus one never carries out of its low 16 bits).  With
//   d = (x0 << 16) + 32767 = (S0 & 0xffff0000) | 0x7fff        (one LOP3; from a packed word's
//                                                               real half: one PRMT)
//   c = d >> 1             = (x0 << 15) + 16383                (one shift)
// the sums are  s0' = c + 2P = s0 + 1  (the two multiply-adds, chained onto c)  and, because
// s0 + s1 = (x0 << 16) + 32766,  s1 = d - s0'  exactly (one subtraction, written as a multiply-add
// by -1 so that it runs on the multiplier pipe: the ALU pipe is the busier one in this kernel).
// Twelve instructions with the two shifts that turn x1 into a multiplicand, against fourteen for
// the form above; no value is unpacked into a sign-extended integer unless it is multiplied.
// (S & 0xffff0000) | 0x7fff as ONE lop3 (LUT 0xEA = (a & b) | c).  Written in PTX because the
// optimiser otherwise turns the OR into an add and moves the constant down the chain, which costs
// an extra add per component.
__device__ __forceinline__ unsigned d_of_top(unsigned S) {
    unsigned d;
    asm("lop3.b32 %0, %1, 0xffff0000, 0x7fff, 0xEA;" : "=r"(d) : "r"(S));
    return d;
}
// -v kept as a value of its own (the optimiser would rather negate a multiplicand per butterfly)
__device__ __forceinline__ int neg_opaque(int v) {
    int n;
    asm("neg.s32 %0, %1;" : "=r"(n) : "r"(v));
    return n;
}
__device__ __forceinline__ unsigned d_of_lo16(unsigned w) { return __byte_perm(w, 0x7fffu, 0x1054); }   // real half of a packed word
__device__ __forceinline__ unsigned sub_on_fma(unsigned d, unsigned s) {      // d - s as mad.lo(s, -1, d)
    unsigned r;
    asm("mad.lo.u32 %0, %1, 0xffffffff, %2;" : "=r"(r) : "r"(s), "r"(d));
    return r;
}
__device__ __forceinline__ unsigned c_of_d(unsigned d) { return (unsigned)((int)d >> 1); }

// x0 given as (dr, di), x1 as sign-extended integers; returns the four sums
__device__ __forceinline__ Bf bf_d(unsigned dr, unsigned di, int x1r, int x1i, int w2r, int w2i, int nw2i) {
    Bf o;
    o.s0r = c_of_d(dr) + (unsigned)(x1r * w2r);
    o.s0r += (unsigned)(x1i * nw2i);
    o.s0i = c_of_d(di) + (unsigned)(x1i * w2r);
    o.s0i += (unsigned)(x1r * w2i);
    o.s1r = sub_on_fma(dr, o.s0r);
    o.s1i = sub_on_fma(di, o.s0i);
    return o;
}
// in place on top-half-form operands
__device__ __forceinline__ void bf_s(unsigned &S0r, unsigned &S0i, unsigned &S1r, unsigned &S1i, int w2r, int w2i, int nw2i) {
    const Bf o = bf_d(d_of_top(S0r), d_of_top(S0i), (int)S1r >> 16, (int)S1i >> 16, w2r, w2i, nw2i);
    S0r = o.s0r; S0i = o.s0i; S1r = o.s1r; S1i = o.s1i;
}

// Trivial twiddles.  For tw = (16384, 0) the RTL's t = (x1 * 16384 + 8191) >> 14 is x1
// itself, for tw = (0, -16384) it is (x1.im, -x1.re) (fft.py:93-96, :152-180; SURVEY.md
// a5), so the butterfly is y = (x0 +- t) >> 1.  The sum of two 16-bit values halved
// fits 16 bits: the 16-bit wrap of :188-191 cannot trigger.  Real operands (the frame
// is real, fft_stream.py:33) stay real through W^0 butterflies.
__device__ __forceinline__ void bf1_rr(int &x0, int &x1) {                 // W^0, real (+) real
    const int s = x0 + x1, d = x0 - x1;
    x0 = s >> 1; x1 = d >> 1;
}
__device__ __forceinline__ void bfj_rr(int &x0r, int &x0i, int &x1r, int &x1i) {   // -j, real (+) real
    const int a = x0r >> 1, b = x1r;                                       // t = (0, -x1)
    x0r = a; x0i = (-b) >> 1;
    x1r = a; x1i = b >> 1;
}
__device__ __forceinline__ void bfj_cc(int &x0r, int &x0i, int &x1r, int &x1i) {   // -j, complex
    const int tr = x1i, ti = -x1r;
    const int a = x0r, b = x0i;
    x0r = (a + tr) >> 1; x0i = (b + ti) >> 1;
    x1r = (a - tr) >> 1; x1i = (b - ti) >> 1;
}

__device__ __forceinline__ int lo16(unsigned w) { return (int)(short)(w & 0xffffu); }
__device__ __forceinline__ int hi16(unsigned w) { return (int)w >> 16; }
__device__ __forceinline__ unsigned pk16(int re, int im) { return __byte_perm((unsigned)re, (unsigned)im, 0x5410); }

// exchange 1: pass-A lane `row` owns positions 16*row + col.  The skew makes both
// the 128-bit row stores and the column reads of pass B conflict free.
__device__ __forceinline__ int xa_addr(int row, int col) { return row * 20 + col + ((row >> 4) << 4); }

// ---------------------------------------------------------------------------------
// FFT-512 of one frame by one warp.
//   xr[l]  in: windowed sample for position 16*lane + l, i.e. sample index
//              brev5(lane) + 32*brev4(l)   (bit-reversed load, fft.py:413-418)
//   out: re[i], im[i] = bin lane + 32*i, i = 0..7
__device__ __forceinline__ void fft512_warp(int (&xr)[16], unsigned *xbuf, const DevTables *tab,
                                            int lane, int (&re)[8], int (&im)[8])
{
    int xi[16];
    // ---- stage 0: W^0 everywhere, real operands
#pragma unroll
    for (int l = 0; l < 16; l += 2) bf1_rr(xr[l], xr[l + 1]);
    // ---- stage 1: W^0 (real stays real) and -j (real pair -> complex pair)
#pragma unroll
    for (int l = 0; l < 16; l += 4) {
        bf1_rr(xr[l], xr[l + 2]);
        bfj_rr(xr[l + 1], xi[l + 1], xr[l + 3], xi[l + 3]);
    }
    // ---- stage 2: k = 0 W^0 (real), 1 W^64, 2 -j (real -> complex), 3 W^192
#pragma unroll
    for (int l = 0; l < 16; l += 8) {
        bf1_rr(xr[l], xr[l + 4]);
        bf(xr[l + 1], xi[l + 1], xr[l + 5], xi[l + 5], 2 * TWA_RE[2], 2 * TWA_IM[2], -2 * TWA_IM[2]);
        bfj_rr(xr[l + 2], xi[l + 2], xr[l + 6], xi[l + 6]);
        bf(xr[l + 3], xi[l + 3], xr[l + 7], xi[l + 7], 2 * TWA_RE[6], 2 * TWA_IM[6], -2 * TWA_IM[6]);
    }
    // ---- stage 3, packed straight into the exchange buffer: slots 0 and 8 are still
    //      real, 4 and 12 turn complex through -j, the other six are general
    {
        unsigned pk[16];
        bf1_rr(xr[0], xr[8]);
        pk[0] = (unsigned)xr[0] & 0xffffu; pk[8] = (unsigned)xr[8] & 0xffffu;
        bfj_rr(xr[4], xi[4], xr[12], xi[12]);
        pk[4] = pk16(xr[4], xi[4]); pk[12] = pk16(xr[12], xi[12]);
#pragma unroll
        for (int l = 1; l < 8; ++l) {
            if (l == 4) continue;
            bf_pk(xr[l], xi[l], xr[l + 8], xi[l + 8], 2 * TWA_RE[l], 2 * TWA_IM[l], -2 * TWA_IM[l], pk[l], pk[l + 8]);
        }
        uint4 *row = reinterpret_cast<uint4 *>(xbuf + xa_addr(lane, 0));
#pragma unroll
        for (int q = 0; q < 4; ++q) row[q] = make_uint4(pk[4 * q], pk[4 * q + 1], pk[4 * q + 2], pk[4 * q + 3]);
    }
    __syncwarp();
    // ---- pass B: lane (h, l4) takes positions 256h + 16m + l4, m = 0..15
    const int h = lane >> 4, l4 = lane & 15;
    // stage s pairs m, m + 2^(s-4); twiddle index ((m mod 2^(s-4))*16 + l4) * (256 >> s).
    // Indices >= 128 are the stored entry turned by -j: (re, im) -> (im, -re)  (fft.py:55-59)
    unsigned sr[16], si[16];                 // top-half form
    {   // stage 4: one twiddle per lane, operands straight from the packed exchange words
        const int2 w = tab->twB[l4];
        const int nwy = neg_opaque(w.y);
#pragma unroll
        for (int m = 0; m < 16; m += 2) {
            const unsigned w0 = xbuf[xa_addr(16 * h + m, l4)], w1 = xbuf[xa_addr(16 * h + m + 1, l4)];
            const Bf o = bf_d(d_of_lo16(w0), d_of_top(w0), lo16(w1), hi16(w1), w.x, w.y, nwy);
            sr[m] = o.s0r; si[m] = o.s0i; sr[m + 1] = o.s1r; si[m + 1] = o.s1i;
        }
    }
    __syncwarp();
    {   // stage 5
        const int2 w = tab->twB[16 + l4];
        const int nwy = neg_opaque(w.y), nwx = neg_opaque(w.x);
#pragma unroll
        for (int m = 0; m < 16; m += 4) {
            bf_s(sr[m], si[m], sr[m + 2], si[m + 2], w.x, w.y, nwy);
            bf_s(sr[m + 1], si[m + 1], sr[m + 3], si[m + 3], w.y, nwx, w.x);    // W * (-j)
        }
    }
    {   // stage 6
#pragma unroll
        for (int mm = 0; mm < 2; ++mm) {
            const int2 w = tab->twB[32 + mm * 16 + l4];
            const int nwy = neg_opaque(w.y), nwx = neg_opaque(w.x);
#pragma unroll
            for (int m = 0; m < 16; m += 8) {
                bf_s(sr[m + mm], si[m + mm], sr[m + mm + 4], si[m + mm + 4], w.x, w.y, nwy);
                bf_s(sr[m + mm + 2], si[m + mm + 2], sr[m + mm + 6], si[m + mm + 6], w.y, nwx, w.x);
            }
        }
    }
    {   // stage 7, packed into the exchange buffer at natural position + 16*(p >> 8)
        unsigned *dst = xbuf + 272 * h + l4;
#pragma unroll
        for (int mm = 0; mm < 4; ++mm) {
            const int2 w = tab->twB[64 + mm * 16 + l4];
            bf_s(sr[mm], si[mm], sr[mm + 8], si[mm + 8], w.x, w.y, neg_opaque(w.y));
            bf_s(sr[mm + 4], si[mm + 4], sr[mm + 12], si[mm + 12], w.y, neg_opaque(w.x), w.x);
            dst[16 * mm] = __byte_perm(sr[mm], si[mm], 0x7632);
            dst[16 * (mm + 8)] = __byte_perm(sr[mm + 8], si[mm + 8], 0x7632);
            dst[16 * (mm + 4)] = __byte_perm(sr[mm + 4], si[mm + 4], 0x7632);
            dst[16 * (mm + 12)] = __byte_perm(sr[mm + 12], si[mm + 12], 0x7632);
        }
    }
    __syncwarp();
    // ---- stage 8: only y0 is ever read (bins 0..255, fft_stream.py:24)
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int b = lane + 32 * i;
        const int2 w = tab->tw8[b];
        const unsigned u0 = xbuf[b], u1 = xbuf[272 + b];
        const int x1r = lo16(u1), x1i = hi16(u1);
        unsigned s0r = c_of_d(d_of_lo16(u0)) + (unsigned)(x1r * w.x);
        s0r -= (unsigned)(x1i * w.y);
        unsigned s0i = c_of_d(d_of_top(u0)) + (unsigned)(x1i * w.x);
        s0i += (unsigned)(x1r * w.y);
        re[i] = (int)s0r >> 16;
        im[i] = (int)s0i >> 16;
    }
    __syncwarp();   // xbuf may be reused by the caller
}

// ---------------------------------------------------------------------------------
// FilterBank.  pw[]: the power spectrum in shared memory in the padded layout
// DevTables::paddr (readable MAXSEG words past its end).  Returns the 16-bit mel value of filter (lane mod NF).
// Segment g = bins seg_first[g] .. seg_close[g].  With A_g = sum d*w over its non-closing
// bins, D_g = sum d over them and PK_g = d at the closing bin, the register the hardware
// reads at the close of segment g is (filterbank.py:119-142)
//     A_{g-1} + (PK_{g-1} << 30)  +  (D_g << 30) - A_g
// of which bits lo_bit .. lo_bit+15 are kept (:122): arithmetic modulo 2^64 is exact.
// Both halves are plain weighted sums over one segment:
//     Abar_g = sum d * wa,  wa = w on non-closing bins, 2^30 on the closing bin
//     Bbar_g = sum d * wb,  wb = 2^30 - w on non-closing bins, 0 on the closing bin
// so lane g-1 walks segment g once with two 64-bit multiply-adds per bin and takes
// Abar_{g-1} from its neighbour.  melab[j * NF + lane] = (wa, wb) of the j-th bin of the
// lane's segment, zero past its end: every lane runs the same MAXSEG steps, no branches
// (32 filters: the weights of steps 2jj, 2jj+1 sit side by side, melab[(jj * NF + lane) * 2 + {0, 1}],
// one 128-bit load per two steps, 22 steps).
template <int NF>
__device__ __forceinline__ unsigned mel_warp(const unsigned *pw, const DevTables *tab, int lane)
{
    constexpr int MAXSEG = (NF == 32) ? 21 : 42;
    const int li = lane & (NF - 1);
    const unsigned *p = pw + tab->seg_walk[li];
    unsigned long long a = 0, b = 0;
    if (NF == 32 && tab->mel_onewide) {
        // One 64-bit and one 32-bit multiply-add per bin instead of two 64-bit ones (IMAD.WIDE issues at
        // half the rate of IMAD).  With nc = 1 on every bin of the segment and 0 on the padding,
        // D = sum d nc (32 bits, wraps):  Bbar_g = ((D_nc + PK) << 30) - (A' + (PK << 30)) = (D << 30) - a,
        // so the slice is taken from  a_{g-1} + (D_g << 30) - a_g;  only bits lo_bit .. lo_bit+15 < 62 are
        // kept, for which (D mod 2^32) << 30 is exact.  Lane 0 walks segment 0 too: mfcc_b200_create()
        // selects this form only when the non-closing weights of segment 0 are all zero (32-filter bank:
        // bins 0..1, bin 0 weighs 0), so Abar_0 = PK_0 << 30 and that bin just counts into lane 0's D.
        // tests/lane_model.py: mel_lanes_onewide.
        const uint4 *wt = reinterpret_cast<const uint4 *>(tab->melab) + li;
        unsigned D = 0;
#pragma unroll
        for (int jj = 0; jj < (MAXSEG + 1) / 2; ++jj) {
            const unsigned d0 = p[2 * jj], d1 = p[2 * jj + 1];
            const uint4 w = wt[jj * NF];
            a += (unsigned long long)d0 * w.x;
            D += d0 * w.y;
            a += (unsigned long long)d1 * w.z;
            D += d1 * w.w;
        }
        unsigned long long pa = __shfl_up_sync(FULL, a, 1);
        if (li == 0) pa = 0;
        const unsigned long long x = pa - a + ((unsigned long long)D << 30);
        return (unsigned)(x >> tab->mel_lo_bit) & 0xffffu;
    } else if (NF == 32) {
        // two steps per 128-bit weight load: melab as uint4 [step / 2][lane]
        const uint4 *wt = reinterpret_cast<const uint4 *>(tab->melab) + li;
#pragma unroll
        for (int jj = 0; jj < (MAXSEG + 1) / 2; ++jj) {
            const unsigned d0 = p[2 * jj], d1 = p[2 * jj + 1];
            const uint4 w = wt[jj * NF];
            a += (unsigned long long)d0 * w.x;
            b += (unsigned long long)d0 * w.y;
            a += (unsigned long long)d1 * w.z;
            b += (unsigned long long)d1 * w.w;
        }
    } else {
        // 16 filters: the two half-warps split the steps of each segment
        const int j0 = (lane >> 4) * (MAXSEG / 2);
#pragma unroll
        for (int j = 0; j < MAXSEG / 2; ++j) {
            const unsigned d = p[j0 + j];
            const uint2 w = tab->melab[(j0 + j) * NF + li];
            a += (unsigned long long)d * w.x;
            b += (unsigned long long)d * w.y;
        }
        a += __shfl_xor_sync(FULL, a, 16);
        b += __shfl_xor_sync(FULL, b, 16);
    }
    // Lane 0 walks segment 0 too (its own output is suppressed, filterbank.py:139, but
    // Abar_0 feeds filter 0): those bins carry wa in the wb slot, so b already holds
    // Abar_0 + Bbar_1 and nothing comes from a neighbour.
    unsigned long long pa = __shfl_up_sync(FULL, a, 1);
    if (li == 0) pa = 0;
    const unsigned long long x = pa + b;
    return (unsigned)(x >> tab->mel_lo_bit) & 0xffffu;
}

// Log2Fix(16, 15) -> Log2FixCalc(27, 11), log.py:123-126, :57-62, :80-102, :133.
// The SHIFT-RIGHT loop runs floor(log2 m) times; the ten squarings set fraction
// bits 10..1 (bit 0 is never reached: cnt stops at 1).
__host__ __device__ __forceinline__ unsigned log2fix_frac(unsigned z)
{
    unsigned frac = 0;
    for (int it = 0; it < 10; ++it) {
        const unsigned c = z * z;
        const unsigned bit = (c >> 23) & 1u;
        z = c >> (11 + bit);
        frac = (frac << 1) | bit;
    }
    return frac;
}

// Same function through the table: the ten squarings depend only on the normalised
// 12-bit mantissa z in [2048, 4096), so their result is tabulated (logfrac[z - 2048],
// filled by mfcc_b200_create() with log2fix_frac).
__device__ __forceinline__ unsigned log2fix_tab(unsigned m, const uint16_t *logfrac)
{
    const unsigned mm = m ? m : 1u;
    const int sh = 31 - __clz(mm);
    const unsigned z = (mm << 11) >> sh;
    return (((unsigned)sh << 11) + ((unsigned)logfrac[z - 2048] << 1)) & 0x7fffu;
}

__device__ __forceinline__ unsigned log2fix_lane(unsigned m)
{
    const unsigned mm = m ? m : 1u;
    const int sh = 31 - __clz(mm);
    unsigned z = (mm << 11) >> sh;
    unsigned frac = 0;
#pragma unroll
    for (int it = 0; it < 10; ++it) {
        const unsigned c = z * z;
        const unsigned bit = (c >> 23) & 1u;
        z = c >> (11 + bit);
        frac = (frac << 1) | bit;
    }
    return (((unsigned)sh << 11) + (frac << 1)) & 0x7fffu;
}

// DCTStream(nfft=NF): a 4NF-point FFT of [0,a,0,b,...,0,b,0,a], real part of bins
// 0..NF-1 (dct_stream.py:29-37).  The even-index half of the first DIT split is all
// zero and yields exact zeros, so only the 2NF-point sub-FFT of the odd samples
// u[m] = L[m], u[2NF-1-m] = L[m] is computed: stage 0 inside the lane, stages
// 1..S-1 by exchanging one value with lane ^ 2^(s-1), then the last stage of the
// 4NF FFT with x0 = 0.  Lane j ends up owning bin j (tests/lane_model.py: dct_plan).
template <int NF>
__device__ __forceinline__ int dct_warp(unsigned L, const DevTables *tab, int lane)
{
    constexpr int S  = (NF == 32) ? 6 : 5;
    constexpr int NB = (NF == 32) ? 5 : 4;
    const int j = lane & (NF - 1);
    const int src = (int)(__brev((unsigned)j) >> (32 - NB));
    const int base = lane & ~(NF - 1);
    int a = (int)__shfl_sync(FULL, L, base + src);
    int b = (int)__shfl_sync(FULL, L, base + NF - 1 - src);
    // stage 0: real inputs < 2^15, twiddle 1: no wrap can occur
    unsigned p0 = (unsigned)((a + b) >> 1) & 0xffffu;       // im = 0
    unsigned p1 = (unsigned)((a - b) >> 1) & 0xffffu;
#pragma unroll
    for (int s = 1; s < S; ++s) {
        const bool up = (lane >> (s - 1)) & 1;
        const unsigned send = up ? p0 : p1;
        const unsigned rc = __shfl_xor_sync(FULL, send, 1 << (s - 1));
        const unsigned x0 = up ? rc : p0;
        const unsigned x1 = up ? p1 : rc;
        const int2 w = tab->dct_tw[s * 32 + j];
        const Bf o = bf_d(d_of_lo16(x0), d_of_top(x0), lo16(x1), hi16(x1), w.x, w.y, -w.y);
        p0 = __byte_perm(o.s0r, o.s0i, 0x7632);
        p1 = __byte_perm(o.s1r, o.s1i, 0x7632);
    }
    const int2 w = tab->dct_twf[j];
    const unsigned q = (unsigned)(lo16(p0) * w.x) - (unsigned)(hi16(p0) * w.y) + 16382u;
    return (int)q >> 16;
}

}  // namespace mfccb
