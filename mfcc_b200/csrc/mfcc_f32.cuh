// Device-side building blocks of the fp32 mode of the B200 MFCC path.
//
// Same chain as the integer mode, in float32, following the reference's float
// prototype notebook/MFCC-INT.ipynb cells 2-10 (pre-emphasis 0.96875, periodic Hamming,
// fft/512 with 257 bins, |X|^2, dense triangular mel filters, log2, orthonormal DCT-II).
//
// One warp = one frame.  The 512 real samples are packed into 256 complex values and
// transformed as 8 x 8 x 4: two radix-8 passes inside the lanes with one transpose
// through shared memory, a radix-4 pass across groups of four lanes by shuffles, then
// the split of the packed spectrum into bins 0..256 fused with the power spectrum.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace mfccf {

constexpr int NFFT = 512;
constexpr int HOP = 170;
constexpr int NBINS1 = 257;            // the notebook keeps bins 0..256 (cell 5)
constexpr int XWF = 1040;              // 32-bit words of the per-warp exchange buffer
constexpr int T1_STRIDE = 36;          // float2 per row of the first transpose
constexpr int PW_OFF = 576;            // power spectrum lives behind the transpose area
constexpr int PW_WORDS = XWF - PW_OFF; // padded power-spectrum area: see DevTablesF::paddr
constexpr int MEL_J0 = 7;              // 32 filters: segments 1..16 have at most 7 bins (checked by create)
constexpr unsigned FULL = 0xffffffffu;

struct alignas(16) DevTablesF {
    float  window[NFFT];               // hamming(512, periodic) / 32: see emph_win()
    float2 twA[8 * 32];                // [k1][lane]  exp(-2 pi i lane k1 / 256)
    float2 twB[8 * 4];                 // [p][c]      exp(-2 pi i c p / 32)
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

// Pre-emphasis + window for the two samples of one 32-bit word (samples 2n, 2n+1).
//   y = x - 0.96875 * xp = (32 x - 31 xp) / 32 is an exact integer / 32 (|.| < 2^21), so
// it is formed in integer arithmetic and turned into a float by the 1.5 * 2^23 bias
// trick (no I2F); the 1/32 is folded into the window table.
__device__ __forceinline__ float2 emph_win(unsigned w, unsigned wprev, float w0, float w1)
{
    const int x0 = (int)(short)(w & 0xffffu), x1 = (int)w >> 16, xp = (int)wprev >> 16;
    constexpr int MAGIC = 0x4B400000;                      // 1.5 * 2^23 = 12582912.0f
    const float f0 = __int_as_float(MAGIC + (32 * x0 - 31 * xp)) - 12582912.0f;
    const float f1 = __int_as_float(MAGIC + (32 * x1 - 31 * x0)) - 12582912.0f;
    return make_float2(f0 * w0, f1 * w1);
}

// index of packed-spectrum bin k in the exchange buffer (4 float2 of padding per 64)
__device__ __forceinline__ int zaddr(int k) { return k + ((k >> 6) << 2); }

// ---------------------------------------------------------------------------------
// Power spectrum |X[b]|^2 / 512^2, b = 0..256, of NFR frames into pw[] (shared memory).
//   raw[f]: frame f's int16 samples in shared memory (even halfword offset), raw[f][-1] is
//           the sample before the frame (0 at the start of a stream)
//   xb[f]:  frame f's exchange buffer (XWF words)
// The kernel is bound by shared-memory bandwidth, and more than half of a frame's
// wavefronts are loads of tables that do not depend on the frame (window, twiddles, mel
// weights, DCT matrix, layout addresses).  With NFR = 2 a warp carries two frames through
// every phase together and loads each table entry once for both.
template <int NFR>
__device__ __forceinline__ void power_warp_f32(const int16_t *const (&raw)[NFR], float *const (&xb)[NFR],
                                               const DevTablesF *tab, int lane)
{
    const float2 *win2 = reinterpret_cast<const float2 *>(tab->window);
    // ---- z[n] = (y[2n] w[2n], y[2n+1] w[2n+1]), n = lane + 32 a
    float2 v[NFR][8];
#pragma unroll
    for (int a = 0; a < 8; ++a) {
        const int n = lane + 32 * a;
        const float2 ww = win2[n];
#pragma unroll
        for (int f = 0; f < NFR; ++f) {
            const unsigned *r32 = reinterpret_cast<const unsigned *>(raw[f]);
            v[f][a] = emph_win(r32[n], r32[n - 1], ww.x, ww.y);
        }
    }
    // ---- pass 1: DFT-8 over a, twiddle W256^(lane * k1), transpose
#pragma unroll
    for (int f = 0; f < NFR; ++f) fft8(v[f]);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int k1 = brev3(j);
        float2 w = make_float2(1.0f, 0.0f);
        if (k1) w = tab->twA[k1 * 32 + lane];
#pragma unroll
        for (int f = 0; f < NFR; ++f) {
            float2 y = v[f][j];
            if (k1) y = cmul(y, w);
            reinterpret_cast<float2 *>(xb[f])[k1 * T1_STRIDE + lane] = y;
        }
    }
    __syncwarp();
    // lane = (k1, c): the 8 values l = c + 4 b of row k1
    const int k1 = lane >> 2, c = lane & 3;
#pragma unroll
    for (int f = 0; f < NFR; ++f)
#pragma unroll
        for (int b = 0; b < 8; ++b) v[f][b] = reinterpret_cast<const float2 *>(xb[f])[k1 * T1_STRIDE + c + 4 * b];
    __syncwarp();
    // ---- pass 2: DFT-8 over b, twiddle W32^(c p) = W256^(8 c p)
#pragma unroll
    for (int f = 0; f < NFR; ++f) fft8(v[f]);
    // ---- pass 3: DFT-4 over c across the four lanes of a group; lane c ends with
    //      q = {0, 2, 1, 3}[c] of Z[k1 + 8 p + 64 q]
    const float sg = (c & 2) ? -1.0f : 1.0f, sg1 = (c & 1) ? -1.0f : 1.0f;
    const int q = ((c & 1) << 1) | (c >> 1);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int p = brev3(j);
        float2 w = make_float2(1.0f, 0.0f);
        if (p) w = tab->twB[p * 4 + c];
#pragma unroll
        for (int f = 0; f < NFR; ++f) {
            float2 y = v[f][j];
            if (p) y = cmul(y, w);
            float2 o;
            o.x = __shfl_xor_sync(FULL, y.x, 2);
            o.y = __shfl_xor_sync(FULL, y.y, 2);
            y = make_float2(fmaf(sg, y.x, o.x), fmaf(sg, y.y, o.y));     // c<2: mine+other, else other-mine
            if (c == 3) y = mul_mj(y);
            o.x = __shfl_xor_sync(FULL, y.x, 1);
            o.y = __shfl_xor_sync(FULL, y.y, 1);
            y = make_float2(fmaf(sg1, y.x, o.x), fmaf(sg1, y.y, o.y));
            reinterpret_cast<float2 *>(xb[f])[zaddr(k1 + 8 * p + 64 * q)] = y;
        }
    }
    __syncwarp();
    // ---- split: X[k] = F + W512^k G, X[256-k] = conj(F - W512^k G), with
    //      F = (Z[k] + conj Z[256-k]) / 2, G = (Z[k] - conj Z[256-k]) / 2j
    constexpr float SC = 1.0f / (512.0f * 512.0f);          // fft / 512 (cell 5), squared
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int k = lane + 32 * i;
        const float2 w = tab->tw512[k];
        const int pa = tab->paddr[k], pb = tab->paddr[256 - k];          // k = 0 gives bin 256
#pragma unroll
        for (int f = 0; f < NFR; ++f) {
            const float2 *xc = reinterpret_cast<const float2 *>(xb[f]);
            float *pw = xb[f] + PW_OFF;
            const float2 A = xc[zaddr(k)];
            const float2 B = xc[zaddr((256 - k) & 255)];
            const float2 F = make_float2(0.5f * (A.x + B.x), 0.5f * (A.y - B.y));
            const float2 G = make_float2(0.5f * (A.y + B.y), -0.5f * (A.x - B.x));
            const float2 T = cmul(G, w);
            const float2 X0 = cadd(F, T), X1 = csub(F, T);
            pw[pa] = (X0.x * X0.x + X0.y * X0.y) * SC;
            pw[pb] = (X1.x * X1.x + X1.y * X1.y) * SC;
        }
    }
    if (lane == 0) {                                         // bin 128: W512^128 = -j
        const int pa = tab->paddr[128];
#pragma unroll
        for (int f = 0; f < NFR; ++f) {
            const float2 A = reinterpret_cast<const float2 *>(xb[f])[zaddr(128)];
            xb[f][PW_OFF + pa] = (A.x * A.x + A.y * A.y) * SC;
        }
    }
    __syncwarp();
}

// mel -> log2 -> DCT-II of the power spectra in xb[f] + PW_OFF (padded layout); lane k gets
// coefficient k of frame f in out[f]; xb[f][0..NF) is scratch.  The filters are triangles
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
        xb[f][lane] = __log2f(fmaxf(up + v[f], tab->mel_floor));
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
            a[f] = reinterpret_cast<const float4 *>(xb[f])[q];                       // L[4q .. 4q+3]
            b[f] = reinterpret_cast<const float4 *>(xb[f])[NF / 4 - 1 - q];          // L[NF-4-4q .. NF-1-4q]
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
