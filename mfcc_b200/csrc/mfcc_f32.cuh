// Device-side building blocks of the fp32 mode of the B200 MFCC path.
//
// Same chain as the integer mode, in float32, following the reference's float
// prototype notebook/MFCC-INT.ipynb cells 2-10 (pre-emphasis 0.96875, periodic Hamming,
// fft/512 with 257 bins, |X|^2, dense triangular mel filters, log2, orthonormal DCT-II).
//
// Half a warp = one frame through the spectrum, a whole warp = one frame through mel/log/DCT;
// a warp carries two frames at a time.  The 512 real samples are packed into 256 complex
// values and transformed as 16 x 16: one 16-point DFT inside each of 16 lanes, twiddles, ONE
// transpose through shared memory, a second 16-point DFT inside the lanes; the split of the
// packed spectrum into bins 0..256 (fused with the power spectrum) takes its conjugate
// partners from the mirror lane by shuffles.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>


namespace mfccf {

constexpr int NFFT = 512;
constexpr int HOP = 170;
constexpr int NBINS1 = 257;            // the notebook keeps bins 0..256 (cell 5)
constexpr int XWF = 592;               // 32-bit words of one frame's exchange buffer (= 16 mod 32: the buffers of the
                                       // two half-warps lie 16 banks apart)
constexpr int T_PITCH = 18;            // float2 per row of the transpose: 64-bit row stores and 128-bit column
                                       // loads are both conflict free at this pitch
constexpr int PW_OFF = 0;              // the power spectrum overwrites the transpose area (16 * T_PITCH * 2 = 576 words)
constexpr int PW_WORDS = 464;          // padded power-spectrum area: see DevTablesF::paddr
constexpr int LM_OFF = 480;            // log-mel row of the frame (NF words), behind the power area
static_assert(16 * T_PITCH * 2 <= XWF && LM_OFF + 32 <= XWF && LM_OFF >= PW_WORDS, "exchange buffer layout");
constexpr int MEL_J0 = 7;              // 32 filters: segments 1..16 have at most 7 bins (checked by create)
constexpr unsigned FULL = 0xffffffffu;

struct alignas(16) DevTablesF {
    float  window[NFFT];               // hamming(512, periodic)
    float2 tw16[16 * 16];              // [k2][n1]    exp(-2 pi i n1 k2 / 256)
    uint32_t pst[8 * 16];              // [k1][l]     paddr[16 k1 + l] | paddr[256 - (16 k1 + l)] << 16
    float2 tw512[132];                 // exp(-2 pi i k / 512), k = 0..128
    float  dct[16 * 32];               // dct[n * 32 + k] = D[k][n], n < nfilters / 2 (the rest by symmetry)
    float2 melw[672];                  // [step][lane] (rise weight, fall weight) of the lane's segment
    uint16_t paddr[264];               // word address of bin b in the padded power area
    uint16_t seg_base[32];             // address of the first bin of lane's segment (segment lane+1)
    float  mel_floor;                  // energies are clamped to this before log2 (0 = off)
    int32_t mel_tail_lane;             // lanes below this have no bins at steps >= MEL_J0 (32 filters): 16, or 0
    int32_t pad[2];
};
static_assert(sizeof(DevTablesF) % 16 == 0, "DevTablesF must be int4-copyable");

__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ float2 cmul(float2 a, float2 w) {
    return make_float2(fmaf(a.x, w.x, -a.y * w.y), fmaf(a.x, w.y, a.y * w.x));
}
__device__ __forceinline__ float2 mul_mj(float2 a) { return make_float2(a.y, -a.x); }    // * (-j)

// 8-point forward DFT in registers (decimation in frequency); on return v[j] holds
// bin brev3(j).
__device__ __forceinline__ void fft8(float2 (&v)[8])
{
    constexpr float S = 0.70710678118654752f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float2 a = cadd(v[i], v[i + 4]), b = csub(v[i], v[i + 4]);
        v[i] = a;
        if (i == 0) v[4] = b;
        else if (i == 1) v[5] = make_float2((b.x + b.y) * S, (b.y - b.x) * S);     // * W8^1
        else if (i == 2) v[6] = mul_mj(b);                                         // * W8^2
        else v[7] = make_float2((b.y - b.x) * S, -(b.x + b.y) * S);                // * W8^3
    }
#pragma unroll
    for (int h = 0; h < 8; h += 4) {
        const float2 a0 = cadd(v[h], v[h + 2]), b0 = csub(v[h], v[h + 2]);
        const float2 a1 = cadd(v[h + 1], v[h + 3]), b1 = mul_mj(csub(v[h + 1], v[h + 3]));
        v[h] = a0; v[h + 1] = a1; v[h + 2] = b0; v[h + 3] = b1;
    }
#pragma unroll
    for (int h = 0; h < 8; h += 2) {
        const float2 a = cadd(v[h], v[h + 1]), b = csub(v[h], v[h + 1]);
        v[h] = a; v[h + 1] = b;
    }
}
__host__ __device__ constexpr int brev3(int v) { return ((v & 1) << 2) | (v & 2) | ((v & 4) >> 2); }

// 16-point forward DFT in registers (decimation in frequency: one radix-2 stage with the
// W16 twiddles, then two 8-point DFTs); on return v[j] holds bin k16(j).
__host__ __device__ constexpr int k16(int j) { return j < 8 ? 2 * brev3(j) : 2 * brev3(j - 8) + 1; }
__host__ __device__ constexpr int j16(int k) { return (k & 1) ? 8 + brev3(k >> 1) : brev3(k >> 1); }   // inverse of k16
__device__ __forceinline__ void fft16(float2 (&v)[16])
{
    constexpr float C1 = 0.92387953251128674f, S1 = 0.38268343236508977f, R = 0.70710678118654752f;
    float2 e[8], o[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        e[i] = cadd(v[i], v[i + 8]);
        const float2 b = csub(v[i], v[i + 8]);
        if (i == 0) o[0] = b;
        else if (i == 1) o[1] = make_float2(fmaf(b.y, S1, b.x * C1), fmaf(b.x, -S1, b.y * C1));       // * (C1, -S1)
        else if (i == 2) o[2] = make_float2((b.x + b.y) * R, (b.y - b.x) * R);                        // * (R, -R)
        else if (i == 3) o[3] = make_float2(fmaf(b.y, C1, b.x * S1), fmaf(b.x, -C1, b.y * S1));       // * (S1, -C1)
        else if (i == 4) o[4] = mul_mj(b);
        else if (i == 5) o[5] = make_float2(fmaf(b.y, C1, b.x * -S1), fmaf(b.x, -C1, b.y * -S1));     // * (-S1, -C1)
        else if (i == 6) o[6] = make_float2((b.y - b.x) * R, -(b.x + b.y) * R);                       // * (-R, -R)
        else o[7] = make_float2(fmaf(b.y, S1, b.x * -C1), fmaf(b.x, -S1, b.y * -C1));                 // * (-C1, -S1)
    }
    fft8(e);
    fft8(o);
#pragma unroll
    for (int j = 0; j < 8; ++j) { v[j] = e[j]; v[8 + j] = o[j]; }
}

// Pre-emphasis + window for the two samples of one 32-bit word (samples 2n, 2n+1), as power_pair_f32 does it.
// int16 -> float without I2F: with its sign bit flipped a sample is the unsigned number
// x + 32768; dropped into the low mantissa bits of 1.5 * 2^23 it reads 12582912 + 32768 + x as a
// float, exactly.  y = x - 0.96875 xp = (32 x - 31 xp) / 32 is exactly representable
// (|32 x - 31 xp| < 2^22), so the single rounding of the FFMA returns it exactly.

// ---------------------------------------------------------------------------------
// Power spectrum |X[b]|^2 / 512^2, b = 0..256, of TWO frames into shared memory: the lower
// half-warp works on one frame, the upper on the other, with the same instruction stream.
//   raw: this lane's frame: its int16 samples in shared memory (even halfword offset), raw[-1]
//        is the sample before the frame (0 at the start of a stream)
//   xb:  this lane's frame's exchange buffer (XWF words); the buffers of the two frames lie
//        16 banks apart, and so do their samples when the frames are 16 apart
// Index plan (tests/lane_model.py: f32_plan): z[n] = (y[2n], y[2n+1]) w, n = n1 + 16 n2;
// bin k = 16 k1 + k2:  Z[k] = sum_n1 W16^(n1 k1) W256^(n1 k2) sum_n2 W16^(n2 k2) z[n1 + 16 n2].
// Pass 1: lane n1 holds n2 = 0..15 in registers; pass 2: lane k2 holds n1 = 0..15 and ends with
// Z[16 k1 + k2], k1 = 0..15.  The real-input split pairs bin k with bin 256 - k =
// 16 (15 - k1) + (16 - k2): lane 16 - k2, so a lane takes the pairs of its k1 = 0..7 and
// receives the upper eight registers of its mirror lane (lane 0 mirrors itself, one
// register further round: 256 - 16 k1 = 16 (16 - k1)).
// NP frames per half-warp go through every phase together, so that each table entry (window,
// twiddles, layout addresses) is loaded once for all of them: a shared-memory wavefront delivers
// 128 bytes to the lanes whether or not the two half-warps ask for the same addresses, so a
// table costs per frame what it costs per lane.
template <int NP>
__device__ __forceinline__ void power_pair_f32(const int16_t *const (&raw)[NP], float *const (&xb)[NP],
                                               const DevTablesF *tab, int lane)
{
    const int l = lane & 15;
    const float2 *win2 = reinterpret_cast<const float2 *>(tab->window);
    float2 v[NP][16];
    // The predecessor of a lane's even sample is the odd sample of the word one lane down (word n - 1
    // = lane l - 1, same a; for l = 0 lane 15's word of a - 1), so it comes by one shuffle of the
    // converted value instead of a second load + conversion; lane 15 sends its previous word's.
    constexpr float BIAS = 12582912.0f + 32768.0f;
    const int srcl = (lane & 16) | ((l + 15) & 15);
    float x1prev[NP];
#pragma unroll
    for (int f = 0; f < NP; ++f) {
        // sample before the frame, into lane 15's "previous word" slot (it reaches lane 0 at a = 0)
        const unsigned up = reinterpret_cast<const unsigned *>(raw[f])[-1] ^ 0x80008000u;
        x1prev[f] = __uint_as_float(__byte_perm(up, 0x4B400000u, 0x7632)) - BIAS;
    }
#pragma unroll
    for (int a = 0; a < 16; ++a) {
        const int n = l + 16 * a;
        const float2 ww = win2[n];
#pragma unroll
        for (int f = 0; f < NP; ++f) {
            const unsigned u = reinterpret_cast<const unsigned *>(raw[f])[n] ^ 0x80008000u;
            const float x0 = __uint_as_float(__byte_perm(u, 0x4B400000u, 0x7610)) - BIAS;
            const float x1 = __uint_as_float(__byte_perm(u, 0x4B400000u, 0x7632)) - BIAS;
            const float xp = __shfl_sync(FULL, l == 15 ? x1prev[f] : x1, srcl);
            x1prev[f] = x1;
            v[f][a] = make_float2(fmaf(xp, -0.96875f, x0) * ww.x, fmaf(x0, -0.96875f, x1) * ww.y);
        }
    }
    // ---- pass 1: DFT-16 over n2, twiddle W256^(n1 k2), transposed store T[k2][n1]
#pragma unroll
    for (int f = 0; f < NP; ++f) fft16(v[f]);
#pragma unroll
    for (int j = 0; j < 16; ++j) {
        const int k2 = k16(j);
        float2 w = make_float2(1.0f, 0.0f);
        if (k2) w = tab->tw16[k2 * 16 + l];
#pragma unroll
        for (int f = 0; f < NP; ++f) {
            float2 y = v[f][j];
            if (k2) y = cmul(y, w);
            reinterpret_cast<float2 *>(xb[f])[k2 * T_PITCH + l] = y;
        }
    }
    __syncwarp();
    // lane k2 = l: row l of T, two entries per 128-bit load
#pragma unroll
    for (int f = 0; f < NP; ++f)
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            const float4 t = reinterpret_cast<const float4 *>(xb[f])[l * (T_PITCH / 2) + q];
            v[f][2 * q] = make_float2(t.x, t.y);
            v[f][2 * q + 1] = make_float2(t.z, t.w);
        }
    __syncwarp();
    // ---- pass 2: DFT-16 over n1: v[j16(k1)] = Z[16 k1 + l]
#pragma unroll
    for (int f = 0; f < NP; ++f) fft16(v[f]);
    // ---- split: X[k] = F + W512^k G, X[256-k] = conj(F - W512^k G), with
    //      F = (Z[k] + conj Z[256-k]) / 2, G = (Z[k] - conj Z[256-k]) / 2j
    // Z[256 - (16 k1 + l)] is register 15 - k1 of the mirror lane (lane 0: its own register 16 - k1)
    constexpr float SC = 1.0f / (512.0f * 512.0f);          // fft / 512 (cell 5), squared
    const int src = (lane & 16) | ((16 - l) & 15);
#pragma unroll
    for (int k1 = 0; k1 < 8; ++k1) {
        const int k = 16 * k1 + l;
        const float2 w = tab->tw512[k];
        const unsigned pa = tab->pst[k1 * 16 + l];                      // where bins k and 256 - k go
#pragma unroll
        for (int f = 0; f < NP; ++f) {
            float2 s = v[f][j16(15 - k1)];
            if (l == 0) s = v[f][j16((16 - k1) & 15)];
            float2 B;
            B.x = __shfl_sync(FULL, s.x, src);
            B.y = __shfl_sync(FULL, s.y, src);
            const float2 A = v[f][j16(k1)];
            const float2 F = make_float2(0.5f * (A.x + B.x), 0.5f * (A.y - B.y));
            const float2 G = make_float2(0.5f * (A.y + B.y), -0.5f * (A.x - B.x));
            const float2 T = cmul(G, w);
            const float2 X0 = cadd(F, T), X1 = csub(F, T);
            float *pw = xb[f] + PW_OFF;
            pw[pa & 0xffffu] = (X0.x * X0.x + X0.y * X0.y) * SC;
            pw[pa >> 16] = (X1.x * X1.x + X1.y * X1.y) * SC;
        }
    }
    if (l == 0) {                                            // bin 128: W512^128 = -j
        const int pa = tab->paddr[128];
#pragma unroll
        for (int f = 0; f < NP; ++f) {
            const float2 A = v[f][j16(8)];
            xb[f][PW_OFF + pa] = (A.x * A.x + A.y * A.y) * SC;
        }
    }
    __syncwarp();
}

// mel -> log2 -> DCT-II of the power spectra in xb[f] + PW_OFF (padded layout); lane k gets
// coefficient k of frame f in out[f]; xb[f][LM_OFF..LM_OFF+NF) is scratch.  The filters are triangles
// (cell 7): bin b of segment g (the bins
// between two filter points) carries the rising weight of filter g and the falling
// weight of filter g-1, so lane g-1 walks segment g once with two multiply-adds per
// bin (weights zero-padded to MAXSEG steps: no branches) and filter m is the rising sum
// of segment m, taken from the neighbouring lane, plus the falling sum of segment m+1.
// The padded layout (DevTablesF::paddr) starts every segment in a bank of its own, so the
// 32 lanes read 32 different banks at every step.
template <int NF, int NFR>
__device__ __forceinline__ void cepstrum_warp_f32(float *const (&xb)[NFR], const DevTablesF *tab, int lane,
                                                  float (&out)[NFR])
{
    constexpr int MAXSEG = (NF == 32) ? 21 : 42;
    const int li = lane & (NF - 1);
    const int sb = tab->seg_base[li];
    float u[NFR], v[NFR];
#pragma unroll
    for (int f = 0; f < NFR; ++f) u[f] = v[f] = 0.0f;
    if (NF == 32) {
#pragma unroll
        for (int j = 0; j < MEL_J0; ++j) {
            const float2 w = tab->melw[j * NF + li];
#pragma unroll
            for (int f = 0; f < NFR; ++f) {
                const float d = xb[f][PW_OFF + sb + j];
                u[f] = fmaf(d, w.x, u[f]);
                v[f] = fmaf(d, w.y, v[f]);
            }
        }
        // only the long segments of the upper filters go on: with the lower half-warp
        // predicated off a 64-bit weight load is one shared-memory wavefront, not two
        if (lane >= tab->mel_tail_lane) {
#pragma unroll
            for (int j = MEL_J0; j < MAXSEG; ++j) {
                const float2 w = tab->melw[j * NF + li];
#pragma unroll
                for (int f = 0; f < NFR; ++f) {
                    const float d = xb[f][PW_OFF + sb + j];
                    u[f] = fmaf(d, w.x, u[f]);
                    v[f] = fmaf(d, w.y, v[f]);
                }
            }
        }
    } else {
        const int j0 = (lane >> 4) * (MAXSEG / 2);
#pragma unroll
        for (int j = 0; j < MAXSEG / 2; ++j) {
            const float2 w = tab->melw[(j0 + j) * NF + li];
#pragma unroll
            for (int f = 0; f < NFR; ++f) {
                const float d = xb[f][PW_OFF + sb + j0 + j];
                u[f] = fmaf(d, w.x, u[f]);
                v[f] = fmaf(d, w.y, v[f]);
            }
        }
#pragma unroll
        for (int f = 0; f < NFR; ++f) {
            u[f] += __shfl_xor_sync(FULL, u[f], 16);
            v[f] += __shfl_xor_sync(FULL, v[f], 16);
        }
    }
    // lane 0 walks segment 0 as well (its rise weights sit in the fall slot: both feed
    // filter 0), so it takes nothing from a neighbour
#pragma unroll
    for (int f = 0; f < NFR; ++f) {
        float up = __shfl_up_sync(FULL, u[f], 1);
        if (li == 0) up = 0.0f;
        xb[f][LM_OFF + lane] = __log2f(fmaxf(up + v[f], tab->mel_floor));
    }
    // orthonormal DCT-II with D[k][NF-1-n] = (-1)^k D[k][n]: NF/2 multiply-adds per lane
    __syncwarp();
    const float sg = (li & 1) ? -1.0f : 1.0f;
#pragma unroll
    for (int f = 0; f < NFR; ++f) out[f] = 0.0f;
#pragma unroll
    for (int q = 0; q < NF / 8; ++q) {
        float4 a[NFR], b[NFR];
#pragma unroll
        for (int f = 0; f < NFR; ++f) {
            a[f] = reinterpret_cast<const float4 *>(xb[f] + LM_OFF)[q];                       // L[4q .. 4q+3]
            b[f] = reinterpret_cast<const float4 *>(xb[f] + LM_OFF)[NF / 4 - 1 - q];          // L[NF-4-4q .. NF-1-4q]
        }
        const float d0 = tab->dct[(4 * q + 0) * 32 + li], d1 = tab->dct[(4 * q + 1) * 32 + li];
        const float d2 = tab->dct[(4 * q + 2) * 32 + li], d3 = tab->dct[(4 * q + 3) * 32 + li];
#pragma unroll
        for (int f = 0; f < NFR; ++f) {
            out[f] = fmaf(fmaf(sg, b[f].w, a[f].x), d0, out[f]);
            out[f] = fmaf(fmaf(sg, b[f].z, a[f].y), d1, out[f]);
            out[f] = fmaf(fmaf(sg, b[f].y, a[f].z), d2, out[f]);
            out[f] = fmaf(fmaf(sg, b[f].x, a[f].w), d3, out[f]);
        }
    }
    __syncwarp();
}

}  // namespace mfccf
