// The librosa parameter set of the reference's comparison script, on the GPU
// (/root/reference/software/genlibrosa.py:21-24:
//  librosa.feature.mfcc(samples, sr=16000, hop_length=170, n_mfcc=32), i.e. centred
//  reflect-padded 2048-point frames, Hann window, power spectrum, 128 Slaney mel filters,
//  10*log10 with an 80 dB floor under the clip's maximum, orthonormal DCT-II).
//
// Two kernels, because the dB floor depends on the maximum over the whole clip:
//   lr_logmel  ONE WARP = ONE FRAME (persistent CTAs of 8 warps).  The 2048 real samples
//              are packed into 1024 complex values and transformed as 32 x 32: every lane
//              runs a 32-point FFT in registers, the warp transposes once through shared
//              memory (with the W1024 twiddles applied on the way), every lane runs a
//              second 32-point FFT.  Real-FFT split + |X|^2, sparse mel filters (four per
//              lane), 10*log10 -> workspace; clip maximum by atomicMax on an
//              order-preserving key.  No block-wide barrier in the frame loop.
//   lr_dct     one warp per four frames: clamp to max - top_db, n_mfcc dot products of n_mels
// Float32 throughout (librosa computes the STFT in complex64 for float32 audio).
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>
#include <algorithm>
#include <cmath>
#include <cuda_runtime.h>

#include "../../include/mfcc_b200.h"

// Scratch of mfcc_b200_librosa_run_static (the entry point that allocates nothing): per-clip
// maxima + log-mel rows of one call.  Every GPU that loads the library holds these bytes until
// the process exits.  In host code this name is a stand-in, not a device address: the address
// comes from cudaGetSymbolAddress (mfcc_b200_librosa_create, once per handle, with the handle's
// device current).
static __device__ __align__(256) unsigned char lr_static_scratch[MFCC_B200_LIBROSA_STATIC_BYTES];

namespace mfccb_host {
int fail_ext(int code, const char *msg);
void count_launch();
}
using mfccb_host::fail_ext;

namespace {

constexpr int LR_NFFT = 2048;
constexpr int LR_HALF = LR_NFFT / 2;          // length of the packed complex transform
constexpr int LR_BINS = LR_HALF + 1;          // 1025 real-spectrum bins
constexpr int LR_WARPS = 8;                   // frames in flight per CTA
constexpr int LR_THREADS = 32 * LR_WARPS;
constexpr int LR_ROW = 34;                    // row pitch of the transposition buffer (float2): 64-bit row stores and
                                              // 128-bit column loads are both conflict free
constexpr int LR_WBUF = 32 * LR_ROW;          // float2 per warp: transposition, spectrum, power
constexpr int LR_MAX_MELS = 128;
constexpr int LR_MAX_MFCC = 64;
constexpr int LR_SLOTS = 5;                   // mel segments per lane (129 segments for 128 filters)
constexpr int LR_SCR_A = 1400, LR_SCR_B = 1568; // per-warp scratch (floats, inside the warp buffer) for the segment sums

struct LrDev {                                 // device-resident tables of one handle
    float *window;                             // [2048], includes the 1/32768 input scale
    float2 *tw32x32;                           // [k1][lane] = exp(-2 pi i k1 lane / 1024)
    float2 *tw2048;                            // [513]  exp(-2 pi i k / 2048)
    int *mel_seg, *mel_base;                   // [5][32] segment of lane l in slot s (-1: none), padded address of its first bin
    uint16_t *paddr;                           // [1025 (+pad)] bin -> word of the bank-padded power layout
    float2 *melT;                              // [steps][32] (rise, fall) weight of each lane at each step
    int mel_steps[LR_SLOTS];                   // steps of each slot (longest segment in the slot)
    float *dct;                                // [n_mels][n_mfcc] (transposed for the lanes)
};

__device__ __forceinline__ float2 cmul(float2 x, float2 w) {
    return make_float2(x.x * w.x - x.y * w.y, x.x * w.y + x.y * w.x);
}

// low half of a word as a signed value: one PRMT in sign-replicate mode (the intrinsic
// __byte_perm masks that selector bit off)
__device__ __forceinline__ int sext16(int w) {
    int r;
    asm("prmt.b32 %0, %1, 0, 0x9910;" : "=r"(r) : "r"(w));
    return r;
}

// numpy.pad(mode="reflect") index map, any number of folds
__device__ __forceinline__ long long reflect_index(long long s, long long n) {
    if (n == 1) return 0;
    const long long p = 2 * (n - 1);
    s %= p;
    if (s < 0) s += p;
    return s < n ? s : p - s;
}

// order-preserving float -> int key (so that atomicMax on ints is a float max)
__device__ __forceinline__ int float_key(float v) {
    const int b = __float_as_int(v);
    return b >= 0 ? b : b ^ 0x7fffffff;
}
__device__ __forceinline__ float key_float(int k) {
    return __int_as_float(k >= 0 ? k : k ^ 0x7fffffff);
}

// cos(pi k / 16), -8 <= k <= 16; no recursion and forced inlining, so that with the loops
// below unrolled every use folds to a literal operand
__device__ __forceinline__ constexpr float cos16(int k) {
    const int a = k < 0 ? -k : k;
    const int r = a > 8 ? 16 - a : a;
    const float c = r == 0 ? 1.0f
                  : r == 1 ? 0.98078528040323044913f
                  : r == 2 ? 0.92387953251128675613f
                  : r == 3 ? 0.83146961230254523708f
                  : r == 4 ? 0.70710678118654752440f
                  : r == 5 ? 0.55557023301960222474f
                  : r == 6 ? 0.38268343236508977173f
                  : r == 7 ? 0.19509032201612826785f
                  : 0.0f;
    return a > 8 ? -c : c;
}
__host__ __device__ constexpr int brev5(int v) {
    return ((v & 1) << 4) | ((v & 2) << 2) | (v & 4) | ((v & 8) >> 2) | ((v & 16) >> 4);
}

// 32-point forward DFT in registers, radix-2 decimation in frequency: on return v[j] holds
// bin brev5(j).  Twiddles are literals; W^0 and W^8 = -j cost nothing.
__device__ __forceinline__ void fft32(float2 (&v)[32])
{
#pragma unroll
    for (int h = 16; h >= 1; h >>= 1) {
#pragma unroll
        for (int g = 0; g < 32; g += 2 * h) {
#pragma unroll
            for (int k = 0; k < h; k++) {
                const float2 a = v[g + k], b = v[g + k + h];
                v[g + k] = make_float2(a.x + b.x, a.y + b.y);
                const float2 d = make_float2(a.x - b.x, a.y - b.y);
                const int e = k * (16 / h);                    // exponent of W32
                if (e == 0) v[g + k + h] = d;
                else if (e == 8) v[g + k + h] = make_float2(d.y, -d.x);
                else v[g + k + h] = cmul(d, make_float2(cos16(e), -cos16(8 - e)));
            }
        }
    }
}

// Shared memory of lr_logmel (dynamic): [tw32x32 1024][tw2048 520] float2, the bin -> padded
// address table (1032 uint16), then the mel weight table [steps][32] float2, then one
// LR_WBUF float2 buffer per warp.
constexpr int LR_TW2048_PAD = 520;
constexpr int LR_PADDR_PAD = 1032;            // uint16 entries, a multiple of 4 (keeps float2 alignment)
__host__ __device__ constexpr size_t lr_smem_bytes(int steps) {
    return (size_t)(1024 + LR_TW2048_PAD) * sizeof(float2) + LR_PADDR_PAD * sizeof(uint16_t) +
           (size_t)steps * 32 * sizeof(float2) +
           (size_t)LR_WARPS * LR_WBUF * sizeof(float2);
}

// d_x [n_clips][clip_stride] int16; logmel [n_clips][n_frames][n_mels]; clipmax [n_clips] keys
// FIXED: the mel step counts are the default bank's (128 Slaney filters at 16 kHz), known at
// compile time and fully unrolled; otherwise they come from the tables.
__host__ __device__ constexpr int lr_def_steps(int s) { return s == 0 ? 3 : s == 1 ? 3 : s == 2 ? 5 : s == 3 ? 11 : 25; }
template <bool FIXED>
__global__ void __launch_bounds__(LR_THREADS, 2)
lr_logmel(const int16_t *__restrict__ d_x, long long clip_stride, long long n_samples, int n_clips,
          long long n_frames, int hop, int center, int n_mels, float amin, LrDev t,
          float *__restrict__ logmel, int *__restrict__ clipmax, float *__restrict__ power_tap)
{
    extern __shared__ __align__(16) unsigned char lr_smem_raw[];
    float2 *tw = reinterpret_cast<float2 *>(lr_smem_raw);              // [k1][lane]
    float2 *tw2 = tw + 1024;                                           // exp(-2 pi i k / 2048)
    uint16_t *pa = reinterpret_cast<uint16_t *>(tw2 + LR_TW2048_PAD);  // bin -> padded power address
    float2 *wT = reinterpret_cast<float2 *>(pa + LR_PADDR_PAD);        // [step][lane] (rise, fall)
    const int steps = t.mel_steps[0] + t.mel_steps[1] + t.mel_steps[2] + t.mel_steps[3] + t.mel_steps[4];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float2 *buf = wT + steps * 32 + warp * LR_WBUF;
    float *pw = reinterpret_cast<float *>(buf);                        // [1025] after the split

    for (int i = tid; i < 1024; i += LR_THREADS) tw[i] = t.tw32x32[i];
    for (int i = tid; i < 513; i += LR_THREADS) tw2[i] = t.tw2048[i];
    for (int i = tid; i < LR_BINS; i += LR_THREADS) pa[i] = t.paddr[i];
    for (int i = tid; i < steps * 32; i += LR_THREADS) wT[i] = t.melT[i];
    // the mel walk reads a few zero-weighted words past bin 1024: they must be finite
    for (int i = lane; i < LR_WBUF; i += 32) buf[i] = make_float2(0.0f, 0.0f);
    __syncthreads();

    // frame-invariant per-lane values: the mel segments this lane walks
    int mseg[LR_SLOTS], mlo[LR_SLOTS];
#pragma unroll
    for (int s = 0; s < LR_SLOTS; s++) {
        mseg[s] = __ldg(t.mel_seg + 32 * s + lane);
        mlo[s] = __ldg(t.mel_base + 32 * s + lane);
    }
    const float2 *win2 = reinterpret_cast<const float2 *>(t.window) + lane;
    // rows whose samples can be fetched as aligned pairs (frame starts are even then)
    const bool pairs = ((reinterpret_cast<uintptr_t>(d_x) & 3) == 0) && ((clip_stride & 1) == 0) &&
                       ((hop & 1) == 0);

    const long long total = (long long)n_clips * n_frames;
    // (clip, frame) of the warp's item, stepped by the grid's stride without a division per frame
    const long long stride = (long long)gridDim.x * LR_WARPS;
    const long long dq = stride / n_frames, dr = stride - dq * n_frames;
    long long item = (long long)blockIdx.x * LR_WARPS + warp;
    long long clip_l = item / n_frames, f = item - clip_l * n_frames;
    for (; item < total; item += stride, clip_l += dq, f += dr) {
        if (f >= n_frames) { f -= n_frames; ++clip_l; }
        const int clip = (int)clip_l;
        const int16_t *x = d_x + clip_l * clip_stride;
        const long long s0 = f * hop - (center ? LR_HALF : 0);

        // lane L takes z[L + 32 i], z[n] = (y[2n] w[2n], y[2n+1] w[2n+1]): coalesced 128-byte rows
        float2 v[32];
        const bool inside = s0 >= 0 && s0 + LR_NFFT <= n_samples;
        if (inside && pairs) {
            const int *xp = reinterpret_cast<const int *>(x + s0) + lane;
#pragma unroll
            for (int i = 0; i < 32; i++) {
                const int s = __ldg(xp + 32 * i);
                const float2 w = __ldg(win2 + 32 * i);
                // low half sign-extended by one PRMT (an I2F.S16 would go through the slow conversion pipe)
                v[i] = make_float2((float)sext16(s) * w.x, (float)(s >> 16) * w.y);
                if ((i & 7) == 7) __syncwarp();                 // keeps at most 8 rows of loads in flight (registers)
            }
        } else {
            // frames that touch the padding (or unaligned rows): staged through the warp's
            // buffer so that v[] keeps compile-time indices
            for (int i = 0; i < 32; i++) {
                const long long i0 = s0 + 2 * (lane + 32 * i), i1 = i0 + 1;
                float y0, y1;
                if (inside) {
                    y0 = (float)x[i0];
                    y1 = (float)x[i1];
                } else if (center) {
                    y0 = (float)x[reflect_index(i0, n_samples)];
                    y1 = (float)x[reflect_index(i1, n_samples)];
                } else {
                    y0 = i0 < n_samples ? (float)x[i0] : 0.0f;
                    y1 = i1 < n_samples ? (float)x[i1] : 0.0f;
                }
                const float2 w = __ldg(win2 + 32 * i);
                buf[lane + 32 * i] = make_float2(y0 * w.x, y1 * w.y);
            }
#pragma unroll
            for (int i = 0; i < 32; i++) v[i] = buf[lane + 32 * i];
            __syncwarp();
        }

        // Z[k1 + 32 k2] = sum_L W32^(L k2) [ W1024^(L k1) sum_i z[L + 32 i] W32^(i k1) ]
        fft32(v);                                               // over i: v[j] = Y_L[k1 = brev5(j)]
#pragma unroll
        for (int j = 0; j < 32; j++) {
            const int k1 = brev5(j);
            buf[k1 * LR_ROW + lane] = k1 ? cmul(v[j], tw[k1 * 32 + lane]) : v[j];
            if ((j & 7) == 7) __syncwarp();                     // bounds the twiddle loads in flight
        }
        __syncwarp();
#pragma unroll
        for (int L = 0; L < 32; L += 2) {                       // lane = k1, values over L, two per 128-bit load
            const float4 q = *reinterpret_cast<const float4 *>(buf + lane * LR_ROW + L);
            v[L] = make_float2(q.x, q.y);
            v[L + 1] = make_float2(q.z, q.w);
        }
        __syncwarp();
        fft32(v);                                               // over L: v[j] = Z[lane + 32 brev5(j)]
        // real-FFT split: X[k] = F + W^k G, X[1024-k] = conj(F - W^k G), k = lane + 32 k2, k2 < 16.
        // Z[1024 - k] = Z[(32 - lane) + 32 (31 - k2)] is register 31 - k2 of the mirror lane 32 - lane
        // (lane 0 mirrors itself one register further round: Z[32 (32 - k2)]), so the partners come
        // by shuffles and the spectrum never goes back to shared memory.
        const int src = (32 - lane) & 31;
        const float2 zc = v[brev5(16)];                         // lane 0: Z[512]
#pragma unroll
        for (int k2 = 0; k2 < 16; k2++) {
            const int k = lane + 32 * k2;
            float2 sd = v[brev5(31 - k2)];
            if (lane == 0) sd = v[brev5((32 - k2) & 31)];
            float2 zb;
            zb.x = __shfl_sync(0xffffffffu, sd.x, src);
            zb.y = __shfl_sync(0xffffffffu, sd.y, src);
            const float2 za = v[brev5(k2)];
            const float2 F = make_float2(0.5f * (za.x + zb.x), 0.5f * (za.y - zb.y));
            const float2 G = make_float2(0.5f * (za.y + zb.y), -0.5f * (za.x - zb.x));
            const float2 T = cmul(G, tw2[k]);
            const float pr = F.x + T.x, pi = F.y + T.y, qr = F.x - T.x, qi = F.y - T.y;
            pw[pa[k]] = pr * pr + pi * pi;                      // bank-padded layout: see the mel walk
            pw[pa[LR_HALF - k]] = qr * qr + qi * qi;
        }
        if (lane == 0) pw[pa[512]] = zc.x * zc.x + zc.y * zc.y;     // W^512 = -i
        __syncwarp();
        if (power_tap) {
            float *dst = power_tap + item * LR_BINS;
            for (int k = lane; k < LR_BINS; k += 32) dst[k] = pw[pa[k]];
        }

        // Mel filters by segments, the way the reference's FilterBank does it
        // (mfcc/core/filterbank.py:119-134): the bins between two filter centres feed the
        // rising side of one filter and the falling side of the previous one, so one pass
        // over a segment with a (rise, fall) weight pair per bin yields both partial sums and
        // every power value is read once.  Lane l walks one segment per slot; weights sit in
        // shared memory as [step][lane], zero beyond a segment's end.  Segments are disjoint,
        // so the power values are laid out with gaps that start every segment of a slot in a
        // bank of its own: the 32 lanes' walks never collide.
        const float2 *w = wT + lane;
#pragma unroll
        for (int s = 0; s < LR_SLOTS; s++) {
            const int n = FIXED ? lr_def_steps(s) : t.mel_steps[s];
            const float *p = pw + mlo[s];
            float ra = 0.0f, fa = 0.0f;
            if (FIXED) {
#pragma unroll
                for (int i = 0; i < lr_def_steps(s); i++) {
                    const float2 ww = w[i * 32];
                    const float d = p[i];
                    ra = fmaf(ww.x, d, ra);
                    fa = fmaf(ww.y, d, fa);
                }
            } else {
#pragma unroll 2
                for (int i = 0; i < n; i++) {
                    const float2 ww = w[i * 32];
                    const float d = p[i];
                    ra = fmaf(ww.x, d, ra);
                    fa = fmaf(ww.y, d, fa);
                }
            }
            w += n * 32;
            if (mseg[s] >= 0) {
                pw[LR_SCR_A + mseg[s]] = ra;
                pw[LR_SCR_B + mseg[s]] = fa;
            }
        }
        __syncwarp();
        float db = -3.0e38f;
        float *lm = logmel + item * n_mels;
#pragma unroll
        for (int j = 0; j < LR_MAX_MELS / 32; j++) {
            const int m = lane + 32 * j;                        // filter m = rise of segment m + fall of segment m + 1
            if (m < n_mels) {
                const float e = pw[LR_SCR_A + m] + pw[LR_SCR_B + m + 1];
                const float d = 3.0102999566398120f * __log2f(fmaxf(e, amin));           // 10 log10
                lm[m] = d;
                db = fmaxf(db, d);
            }
        }
        for (int o = 16; o; o >>= 1) db = fmaxf(db, __shfl_xor_sync(0xffffffffu, db, o));
        if (lane == 0) atomicMax(clipmax + clip, float_key(db));
        __syncwarp();                                           // pw is free again
    }
}

// logmel [n_clips][n_frames][n_mels] -> out [n_clips][n_frames][n_mfcc] (row pitch out_clip_stride).
// One warp takes four consecutive frames at a time: their clamped log-mel rows are interleaved
// in shared memory ([n][4], one 16-byte broadcast load per mel band), lane k accumulates
// coefficient k of all four, so every DCT weight is loaded once per four frames.
__global__ void __launch_bounds__(256)
lr_dct(const float *__restrict__ logmel, const int *__restrict__ clipmax, int n_clips, long long n_frames,
       int n_mels, int n_mfcc, float top_db, LrDev t, float *__restrict__ out, long long out_clip_stride)
{
    extern __shared__ __align__(16) float dsm[];
    float *D = dsm;                                            // [n_mels][n_mfcc]
    float4 *vbuf = reinterpret_cast<float4 *>(dsm + ((n_mels * n_mfcc + 3) & ~3));    // [8 warps][n_mels]
    for (int i = threadIdx.x; i < n_mels * n_mfcc; i += blockDim.x) D[i] = t.dct[i];
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float4 *v4 = vbuf + warp * n_mels;
    float *v = reinterpret_cast<float *>(v4);
    const long long total = (long long)n_clips * n_frames;
    const long long groups = (total + 3) / 4;
    for (long long g = (long long)blockIdx.x * 8 + warp; g < groups; g += (long long)gridDim.x * 8) {
        const long long item0 = g * 4;
#pragma unroll
        for (int q = 0; q < 4; q++) {
            const long long item = item0 + q;
            if (item < total) {
                const int clip = (int)(item / n_frames);
                const float floor_db = top_db >= 0.0f ? key_float(clipmax[clip]) - top_db : -3.0e38f;
                for (int n = lane; n < n_mels; n += 32) v[4 * n + q] = fmaxf(logmel[item * n_mels + n], floor_db);
            } else {
                for (int n = lane; n < n_mels; n += 32) v[4 * n + q] = 0.0f;
            }
        }
        __syncwarp();
        for (int k = lane; k < n_mfcc; k += 32) {
            float4 acc = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
#pragma unroll 8
            for (int n = 0; n < n_mels; n++) {
                const float d = D[n * n_mfcc + k];
                const float4 x = v4[n];
                acc.x = fmaf(d, x.x, acc.x);
                acc.y = fmaf(d, x.y, acc.y);
                acc.z = fmaf(d, x.z, acc.z);
                acc.w = fmaf(d, x.w, acc.w);
            }
            const float r[4] = {acc.x, acc.y, acc.z, acc.w};
#pragma unroll
            for (int q = 0; q < 4; q++) {
                const long long item = item0 + q;
                if (item < total) {
                    const int clip = (int)(item / n_frames);
                    const long long f = item - (long long)clip * n_frames;
                    out[(long long)clip * out_clip_stride + f * n_mfcc + k] = r[q];
                }
            }
        }
        __syncwarp();
    }
}

// The same product for the default shape (128 mel bands, 32 coefficients) with the DCT matrix
// in the kernel's parameter space (16 KB of constant bank 0): lane = frame, 32 accumulators per
// lane, the weights are constant-bank operands of the multiply-adds, so the inner loop has one
// shared-memory load per 32 multiply-adds.  A warp takes 32 consecutive frames: their clamped
// rows are transposed through shared memory ([band][frame], pitch 33).  Summation order per
// coefficient is the same as lr_dct's (ascending band), so the results are identical.
constexpr int DC_MELS = 128, DC_MFCC = 32, DC_WARPS = 4, DC_PITCH = 33;
struct alignas(16) LrDctParam { float4 D[DC_MELS * DC_MFCC / 4]; };   // [band][coefficient], four coefficients per entry
__global__ void __launch_bounds__(DC_WARPS * 32)
lr_dct_c(const float *__restrict__ logmel, const int *__restrict__ clipmax, long long total, long long n_frames,
         float top_db, const __grid_constant__ LrDctParam prm, float *__restrict__ out, long long out_clip_stride)
{
    extern __shared__ __align__(16) float dsm[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float *xT = dsm + warp * (DC_MELS * DC_PITCH);
    const long long groups = (total + 31) / 32;
    for (long long g = (long long)blockIdx.x * DC_WARPS + warp; g < groups; g += (long long)gridDim.x * DC_WARPS) {
        const long long item0 = g * 32;
        // lane j's frame: clip and frame index, and the dB floor of its clip
        const long long my_item = item0 + lane;
        const long long my_clip = (my_item < total ? my_item : total - 1) / n_frames;
        const long long my_f = my_item - my_clip * n_frames;
        const float my_floor = top_db >= 0.0f ? key_float(clipmax[my_clip]) - top_db : -3.0e38f;
        // eight rows at a time: 32 independent global loads in flight per lane
#pragma unroll 1
        for (int jb = 0; jb < 32; jb += 8) {
            float r[8][DC_MELS / 32];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const long long item = item0 + jb + j < total ? item0 + jb + j : total - 1;
                const float *row = logmel + item * DC_MELS;
#pragma unroll
                for (int c = 0; c < DC_MELS / 32; ++c) r[j][c] = row[lane + 32 * c];
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float floor_db = __shfl_sync(0xffffffffu, my_floor, jb + j);
#pragma unroll
                for (int c = 0; c < DC_MELS / 32; ++c) xT[(lane + 32 * c) * DC_PITCH + jb + j] = fmaxf(r[j][c], floor_db);
            }
        }
        __syncwarp();
        float acc[DC_MFCC];
#pragma unroll
        for (int k = 0; k < DC_MFCC; ++k) acc[k] = 0.0f;
#pragma unroll 4
        for (int n = 0; n < DC_MELS; ++n) {
            const float x = xT[n * DC_PITCH + lane];
#pragma unroll
            for (int k = 0; k < DC_MFCC; k += 4) {
                const float4 d = prm.D[n * (DC_MFCC / 4) + (k >> 2)];       // uniform 128-bit constant load
                acc[k] = fmaf(d.x, x, acc[k]);
                acc[k + 1] = fmaf(d.y, x, acc[k + 1]);
                acc[k + 2] = fmaf(d.z, x, acc[k + 2]);
                acc[k + 3] = fmaf(d.w, x, acc[k + 3]);
            }
        }
        if (item0 + lane < total) {
            float4 *o = reinterpret_cast<float4 *>(out + my_clip * out_clip_stride + my_f * DC_MFCC);
#pragma unroll
            for (int k = 0; k < DC_MFCC; k += 4) o[k >> 2] = make_float4(acc[k], acc[k + 1], acc[k + 2], acc[k + 3]);
        }
        __syncwarp();
    }
}

// The same product on the tensor cores (default shape: 128 mel bands, 32 coefficients):
//   out[frames x 32] = max(logmel, floor)[frames x 128] . D[128 x 32]
// as warp-level mma.sync.m16n8k8 TF32 with every operand split into a TF32 head and an fp32 tail
// (x = hi + lo, hi = x with its 13 low mantissa bits cleared): hi.hi + lo.hi + hi.lo leaves a
// relative error near 2^-21 per product, fp32 accumulation -- float32-class, checked against the
// float64 restatement like the FMA kernels (tests/test_librosa_mode.py).  A warp takes 32 consecutive
// frames (two 16-row tiles) and all 32 coefficients (four 8-column tiles): 16 k-steps x 4 x 2 x 3 =
// 384 MMAs per 32 frames, 12 per frame (0.46 MMA per clock per SM measured,
// profiles/micro/mma_bench.cu: 26 SM cycles per frame, about what the rows' 640 bytes take at the
// HBM rate) against the 128 FFMA + 32 constant loads per frame of lr_dct_c.
// A operands come straight from global memory, one 128-bit load per row and pair of k-steps: the k
// index of an MMA is only summed over, so slot t of k-step 2S+e stands for band 16S + 4t + 2e and
// slot t+4 for the next band, and the four lanes of a row read 64 contiguous bytes.  The B fragments
// (the DCT matrix in that band order, head and tail) are built once per CTA in shared memory,
// [k-step][column tile][lane], one 128-bit load per MMA triple.
constexpr int DM_WARPS = 8;
__device__ __forceinline__ float tf32_head(float x) { return __uint_as_float(__float_as_uint(x) & 0xffffe000u); }
__device__ __forceinline__ void mma_tf32(float (&c)[4], const float (&a)[4], float b0, float b1) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(__float_as_uint(a[0])), "r"(__float_as_uint(a[1])), "r"(__float_as_uint(a[2])), "r"(__float_as_uint(a[3])),
                   "r"(__float_as_uint(b0)), "r"(__float_as_uint(b1)));
}

__global__ void __launch_bounds__(DM_WARPS * 32)
lr_dct_mma(const float *__restrict__ logmel, const int *__restrict__ clipmax, long long total, long long n_frames,
           float top_db, const float *__restrict__ D /* [128][32] */, float *__restrict__ out, long long out_clip_stride)
{
    extern __shared__ __align__(16) float4 bfrag[];            // [16 k-steps][4 column tiles][32 lanes]: (hi0, hi1, lo0, lo1)
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, g = lane >> 2, t = lane & 3;
    for (int i = threadIdx.x; i < 16 * 4 * 32; i += DM_WARPS * 32) {
        const int ln = i & 31, j = (i >> 5) & 3, s = i >> 7;
        const int band = 16 * (s >> 1) + 4 * (ln & 3) + 2 * (s & 1), col = 8 * j + (ln >> 2);
        const float v0 = D[band * DC_MFCC + col], v1 = D[(band + 1) * DC_MFCC + col];
        const float h0 = tf32_head(v0), h1 = tf32_head(v1);
        bfrag[i] = make_float4(h0, h1, v0 - h0, v1 - h1);
    }
    __syncthreads();
    const long long groups = (total + 31) / 32;
    for (long long grp = (long long)blockIdx.x * DM_WARPS + warp; grp < groups; grp += (long long)gridDim.x * DM_WARPS) {
        const long long item0 = grp * 32;
        // lane j looks after frame j of the group: its clip, frame index and the dB floor of its clip
        const long long my_item = item0 + lane < total ? item0 + lane : total - 1;
        const long long my_clip = my_item / n_frames;
        const long long my_f = my_item - my_clip * n_frames;
        const float my_floor = top_db >= 0.0f ? key_float(clipmax[my_clip]) - top_db : -3.0e38f;
        const float *rowp[4];                                   // rows g, g + 8, g + 16, g + 24 of the group
        float rfloor[4];
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const int row = g + 8 * r;
            const long long item = item0 + row < total ? item0 + row : total - 1;
            rowp[r] = logmel + item * DC_MELS + 4 * t;
            rfloor[r] = __shfl_sync(0xffffffffu, my_floor, row);
        }
        float acc[2][4][4];
#pragma unroll
        for (int m = 0; m < 2; ++m)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[m][j][0] = acc[m][j][1] = acc[m][j][2] = acc[m][j][3] = 0.0f;
#pragma unroll 1
        for (int S = 0; S < 8; ++S) {
            float4 x[4];
            float part[2][4][4];
#pragma unroll
            for (int m = 0; m < 2; ++m)
#pragma unroll
                for (int j = 0; j < 4; ++j) part[m][j][0] = part[m][j][1] = part[m][j][2] = part[m][j][3] = 0.0f;
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                x[r] = *reinterpret_cast<const float4 *>(rowp[r] + 16 * S);
                x[r].x = fmaxf(x[r].x, rfloor[r]); x[r].y = fmaxf(x[r].y, rfloor[r]);
                x[r].z = fmaxf(x[r].z, rfloor[r]); x[r].w = fmaxf(x[r].w, rfloor[r]);
            }
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                float ahi[2][4], alo[2][4];
#pragma unroll
                for (int m = 0; m < 2; ++m) {
                    // a0 (row g, slot t), a1 (row g + 8, slot t), a2 (row g, slot t + 4), a3 (row g + 8, slot t + 4)
                    const float a[4] = {e ? x[2 * m].z : x[2 * m].x, e ? x[2 * m + 1].z : x[2 * m + 1].x,
                                        e ? x[2 * m].w : x[2 * m].y, e ? x[2 * m + 1].w : x[2 * m + 1].y};
#pragma unroll
                    for (int q = 0; q < 4; ++q) { ahi[m][q] = tf32_head(a[q]); alo[m][q] = a[q] - ahi[m][q]; }
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float4 b = bfrag[((2 * S + e) * 4 + j) * 32 + lane];
#pragma unroll
                    for (int m = 0; m < 2; ++m) {
                        // small terms first; the tensor core's own accumulation truncates, so it only ever
                        // sums the 16 bands of one S step from zero (part), and the running sums are
                        // ordinary round-to-nearest float32 additions below
                        mma_tf32(part[m][j], alo[m], b.x, b.y);
                        mma_tf32(part[m][j], ahi[m], b.z, b.w);
                        mma_tf32(part[m][j], ahi[m], b.x, b.y);
                    }
                }
            }
#pragma unroll
            for (int m = 0; m < 2; ++m)
#pragma unroll
                for (int j = 0; j < 4; ++j)
#pragma unroll
                    for (int q = 0; q < 4; ++q) acc[m][j][q] += part[m][j][q];
        }
        // c0, c1: row g, columns 8j + 2t, + 1; c2, c3: row g + 8
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            const int row = g + 8 * r;
            const long long clip = __shfl_sync(0xffffffffu, my_clip, row), f = __shfl_sync(0xffffffffu, my_f, row);
            if (item0 + row < total) {
                float *o = out + clip * out_clip_stride + f * DC_MFCC + 2 * t;
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    *reinterpret_cast<float2 *>(o + 8 * j) = make_float2(acc[r >> 1][j][2 * (r & 1)], acc[r >> 1][j][2 * (r & 1) + 1]);
            }
        }
    }
}

__global__ void lr_init_max(int *clipmax, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) clipmax[i] = (int)0x80000000;
}

using LrLogmelFn = void (*)(const int16_t *, long long, long long, int, long long, int, int, int, float, LrDev, float *,
                            int *, float *);

struct LrHandle {
    int hop, n_mels, n_mfcc, center, sms, ctas1;
    LrLogmelFn logmel;
    size_t smem1;
    float amin, top_db;
    LrDev dev;
    std::vector<void *> allocs;
    LrDctParam *dct_param = nullptr;            // host copy of the DCT matrix for lr_dct_c (default shape only)
    long long chunk_bytes = 1LL << 60;          // log-mel rows kept between the two passes (default: the whole batch)
    int dct_tensor = 1;                         // default shape: DCT product by lr_dct_mma (0: lr_dct_c, FMA pipes)
    void *static_scratch = nullptr;             // device address of lr_static_scratch on the handle's device
    int device = -1;                            // the device the handle was created on
    std::string static_err;                     // why static_scratch is null, if it is
    ~LrHandle() { delete dct_param; }
    // clips per (log-mel, DCT) launch pair.  With chunks of ~48 MiB the rows written by lr_logmel are read
    // back by the DCT kernel while they are still in L2 and the next chunk overwrites the same lines:
    // measured 1499 -> 696 bytes of DRAM traffic per frame, but 8 % fewer frames/s (one persistent launch
    // pair per 22 one-minute clips, each with its own tail), on a path that uses 7 % of the HBM
    // bandwidth either way (profiles/r02_lr_chunks.txt).  So the default is one chunk.
    long long chunk_clips(long long n_clips, long long n_frames) const {
        long long per = chunk_bytes / (n_frames * n_mels * (long long)sizeof(float));
        if (per < 1) per = 1;
        return per < n_clips ? per : n_clips;
    }
};

}  // namespace

// =================================================================== C ABI
extern "C" {

int mfcc_b200_librosa_create(const mfcc_b200_librosa_cfg *cfg, void **handle)
{
    if (!cfg || !handle) return fail_ext(MFCC_B200_EINVAL, "null argument");
    *handle = nullptr;
    if (cfg->struct_size != (int32_t)sizeof(mfcc_b200_librosa_cfg))
        return fail_ext(MFCC_B200_EINVAL, "mfcc_b200_librosa_cfg size mismatch");
    if (cfg->n_fft != LR_NFFT) return fail_ext(MFCC_B200_EUNSUPPORTED, "the librosa-set kernels cover n_fft = 2048");
    if (cfg->hop < 1) return fail_ext(MFCC_B200_EINVAL, "hop must be >= 1");
    if (cfg->n_mels < 1 || cfg->n_mels > LR_MAX_MELS) return fail_ext(MFCC_B200_EUNSUPPORTED, "n_mels must be 1..128");
    if (cfg->n_mfcc < 1 || cfg->n_mfcc > LR_MAX_MFCC || cfg->n_mfcc > cfg->n_mels)
        return fail_ext(MFCC_B200_EUNSUPPORTED, "n_mfcc must be 1..min(64, n_mels)");
    if (!cfg->window || !cfg->mel || !cfg->dct) return fail_ext(MFCC_B200_EINVAL, "window, mel and dct tables are required");
    if (!(cfg->amin > 0.0f)) return fail_ext(MFCC_B200_EINVAL, "amin must be > 0");

    LrHandle *h = new LrHandle();
    h->hop = cfg->hop; h->n_mels = cfg->n_mels; h->n_mfcc = cfg->n_mfcc; h->center = cfg->center ? 1 : 0;
    h->amin = cfg->amin; h->top_db = cfg->top_db;
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&h->sms, cudaDevAttrMultiProcessorCount, dev);

    // host-side tables
    const double PI = 3.14159265358979323846;
    std::vector<float2> tw1(1024), tw2(513);
    for (int k1 = 0; k1 < 32; k1++)
        for (int L = 0; L < 32; L++) {
            const double ang = 2 * PI * (k1 * L) / 1024.0;
            tw1[k1 * 32 + L] = make_float2((float)cos(ang), (float)-sin(ang));
        }
    for (int k = 0; k < 513; k++) tw2[k] = make_float2((float)cos(2 * PI * k / 2048), (float)-sin(2 * PI * k / 2048));
    // Segments: every bin may feed two consecutive filters.  Segment g holds the bins whose
    // filters are {g-1, g}; its bins carry the rise weight of filter g and the fall weight
    // of filter g-1 (the structure of any bank of ordered, half-overlapping triangles).
    const int nm = cfg->n_mels, nseg = nm + 1;
    std::vector<int> seg(LR_BINS), first(nseg, -1), len(nseg, 0);
    {
        int cur = 0;
        for (int b = 0; b < LR_BINS; b++) {
            int lo = -1, hi = -1;
            for (int m = 0; m < nm; m++)
                if (cfg->mel[(size_t)m * LR_BINS + b] != 0.0f) { if (lo < 0) lo = m; hi = m; }
            int g = cur;
            if (lo >= 0) {
                if (hi - lo > 1) { delete h; return fail_ext(MFCC_B200_EUNSUPPORTED, "mel: a bin feeds more than two consecutive filters"); }
                if (hi > lo) g = hi;
                else {
                    // one filter only: on its falling side (weight going down, or already in the
                    // next segment) the bin belongs to segment lo+1, else to segment lo
                    const bool falling = b > 0 && cfg->mel[(size_t)lo * LR_BINS + b] < cfg->mel[(size_t)lo * LR_BINS + b - 1];
                    g = (falling || cur == lo + 1) ? lo + 1 : (cur > lo ? cur : lo);
                }
                if (g < cur || g > lo + 1 || g > nm) { delete h; return fail_ext(MFCC_B200_EUNSUPPORTED, "mel: filters are not ordered triangles"); }
            }
            seg[b] = cur = g;
        }
        for (int b = 0; b < LR_BINS; b++) { if (first[seg[b]] < 0) first[seg[b]] = b; ++len[seg[b]]; }
        // every non-zero weight must be accounted for by the (rise, fall) pair of its bin
        for (int m = 0; m < nm; m++)
            for (int b = 0; b < LR_BINS; b++)
                if (cfg->mel[(size_t)m * LR_BINS + b] != 0.0f && seg[b] != m && seg[b] != m + 1) {
                    delete h;
                    return fail_ext(MFCC_B200_EUNSUPPORTED, "mel: filters are not ordered triangles");
                }
    }
    // lane l walks one segment per slot; the first slot takes what is left over after full
    // blocks of 32 (the shortest segments), every slot runs as many steps as its longest segment
    const int nslots = (nseg + 31) / 32, head = nseg - 32 * (nslots - 1);
    std::vector<int> segs(LR_SLOTS * 32, -1), base(LR_SLOTS * 32, 0), pbase(nseg, 0);
    {
        // padded layout: the segments of one slot start in 32 different banks
        bool used[LR_SLOTS][32] = {};
        int next = 0;
        for (int g = 0; g < nseg; g++) {
            const int s = g < head ? 0 : 1 + (g - head) / 32;
            int a = next;
            while (used[s][a & 31]) a++;
            used[s][a & 31] = true;
            pbase[g] = a;
            next = a + len[g];
        }
    }
    int steps = 0, soff[LR_SLOTS];
    for (int s = 0; s < LR_SLOTS; s++) {
        int n = 0;
        for (int l = 0; l < 32 && s < nslots; l++) {
            const int g = s == 0 ? (l < head ? l : -1) : head + 32 * (s - 1) + l;
            if (g < 0 || g >= nseg) continue;
            segs[32 * s + l] = g;
            base[32 * s + l] = pbase[g];
            n = std::max(n, len[g]);
        }
        h->dev.mel_steps[s] = n;
        soff[s] = steps;
        steps += n;
    }
    std::vector<uint16_t> paddr(LR_PADDR_PAD, 0);
    int max_read = 0;
    for (int b = 0; b < LR_BINS; b++) paddr[b] = (uint16_t)(pbase[seg[b]] + (b - first[seg[b]]));
    for (int s = 0; s < LR_SLOTS; s++)
        for (int l = 0; l < 32; l++)
            if (segs[32 * s + l] >= 0) max_read = std::max(max_read, base[32 * s + l] + h->dev.mel_steps[s]);
    if (max_read > LR_SCR_A || LR_SCR_B + nseg > 2 * LR_WBUF) {
        delete h;
        return fail_ext(MFCC_B200_EUNSUPPORTED, "mel: padded power layout does not fit the warp buffer");
    }
    std::vector<float2> wT((size_t)std::max(1, steps) * 32, make_float2(0.0f, 0.0f));
    for (int s = 0; s < LR_SLOTS; s++)
        for (int l = 0; l < 32; l++) {
            const int g = segs[32 * s + l];
            if (g < 0) continue;
            for (int i = 0; i < len[g]; i++) {
                const int b = first[g] + i;
                const float rise = g < nm ? cfg->mel[(size_t)g * LR_BINS + b] : 0.0f;
                const float fall = g >= 1 ? cfg->mel[(size_t)(g - 1) * LR_BINS + b] : 0.0f;
                wT[(size_t)(soff[s] + i) * 32 + l] = make_float2(rise, fall);
            }
        }
    h->smem1 = lr_smem_bytes(steps);
    std::vector<float> dT((size_t)cfg->n_mels * cfg->n_mfcc);
    for (int k = 0; k < cfg->n_mfcc; k++)
        for (int n = 0; n < cfg->n_mels; n++) dT[(size_t)n * cfg->n_mfcc + k] = cfg->dct[(size_t)k * cfg->n_mels + n];

    cudaError_t e = cudaSuccess;
    auto upload = [&](const void *src, size_t bytes) -> void * {
        void *p = nullptr;
        if (e != cudaSuccess) return nullptr;
        e = cudaMalloc(&p, bytes);
        if (e != cudaSuccess) return nullptr;
        h->allocs.push_back(p);
        e = cudaMemcpy(p, src, bytes, cudaMemcpyHostToDevice);
        return p;
    };
    h->dev.window = (float *)upload(cfg->window, LR_NFFT * sizeof(float));
    h->dev.tw32x32 = (float2 *)upload(tw1.data(), tw1.size() * sizeof(float2));
    h->dev.tw2048 = (float2 *)upload(tw2.data(), tw2.size() * sizeof(float2));
    h->dev.mel_seg = (int *)upload(segs.data(), segs.size() * sizeof(int));
    h->dev.mel_base = (int *)upload(base.data(), base.size() * sizeof(int));
    h->dev.paddr = (uint16_t *)upload(paddr.data(), paddr.size() * sizeof(uint16_t));
    h->dev.melT = (float2 *)upload(wT.data(), wT.size() * sizeof(float2));
    if (e == cudaSuccess && h->smem1 > 200 * 1024) {
        for (void *p : h->allocs) cudaFree(p);
        delete h;
        return fail_ext(MFCC_B200_EUNSUPPORTED, "mel filters too wide for the shared-memory weight table");
    }
    bool def = true;
    for (int s = 0; s < LR_SLOTS; s++) def = def && h->dev.mel_steps[s] == lr_def_steps(s);
    h->logmel = def ? lr_logmel<true> : lr_logmel<false>;
    if (e == cudaSuccess) e = cudaFuncSetAttribute(h->logmel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->smem1);
    if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&h->ctas1, h->logmel, LR_THREADS, h->smem1);
    if (e == cudaSuccess && h->ctas1 < 1) h->ctas1 = 1;
    h->dev.dct = (float *)upload(dT.data(), dT.size() * sizeof(float));
    if (h->n_mels == DC_MELS && h->n_mfcc == DC_MFCC) {
        h->dct_param = new LrDctParam;
        memcpy(h->dct_param->D, dT.data(), sizeof(LrDctParam));
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(lr_dct_c, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)(DC_WARPS * DC_MELS * DC_PITCH * sizeof(float)));
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(lr_dct_mma, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(16 * 4 * 32 * sizeof(float4)));
    }
    if (e != cudaSuccess) {
        for (void *p : h->allocs) cudaFree(p);
        delete h;
        return fail_ext(MFCC_B200_ECUDA, cudaGetErrorString(e));
    }
    // set-up of mfcc_b200_librosa_run_static: the device address of the file-scope scratch array on
    // THIS device (the current one).  A failure here does not fail create; the entry point reports it.
    {
        void *sp = nullptr;
        cudaError_t es = cudaGetDevice(&h->device);
        if (es == cudaSuccess) es = cudaGetSymbolAddress(&sp, lr_static_scratch);
        if (es == cudaSuccess) h->static_scratch = sp;
        else { h->static_err = cudaGetErrorString(es); cudaGetLastError(); }
    }
    *handle = h;
    return MFCC_B200_OK;
}

void mfcc_b200_librosa_destroy(void *handle)
{
    LrHandle *h = static_cast<LrHandle *>(handle);
    if (!h) return;
    for (void *p : h->allocs) cudaFree(p);
    delete h;
}

int64_t mfcc_b200_librosa_frames(void *handle, int64_t n_samples)
{
    LrHandle *h = static_cast<LrHandle *>(handle);
    if (!h || n_samples < 0) return -1;
    if (h->center) return n_samples >= 1 ? 1 + n_samples / h->hop : 0;
    return n_samples >= LR_NFFT ? 1 + (n_samples - LR_NFFT) / h->hop : 0;
}

static int64_t lr_max_bytes(int32_t n_clips) { return (((int64_t)n_clips * 4) + 255) & ~(int64_t)255; }

int mfcc_b200_librosa_config(void *handle, int64_t chunk_bytes)
{
    LrHandle *h = static_cast<LrHandle *>(handle);
    if (!h || chunk_bytes < 0) return fail_ext(MFCC_B200_EINVAL, "bad argument");
    if (chunk_bytes) h->chunk_bytes = chunk_bytes;
    return MFCC_B200_OK;
}

int mfcc_b200_librosa_dct_path(void *handle, int32_t tensor_cores)
{
    LrHandle *h = static_cast<LrHandle *>(handle);
    if (!h) return fail_ext(MFCC_B200_EINVAL, "null handle");
    h->dct_tensor = tensor_cores ? 1 : 0;
    return MFCC_B200_OK;
}

int mfcc_b200_librosa_workspace(void *handle, int32_t n_clips, int64_t n_frames, int64_t *bytes)
{
    LrHandle *h = static_cast<LrHandle *>(handle);
    if (!h || !bytes || n_clips < 0 || n_frames < 0) return fail_ext(MFCC_B200_EINVAL, "bad argument");
    const long long cc = n_frames ? h->chunk_clips(n_clips, n_frames) : 0;
    *bytes = lr_max_bytes(n_clips) + cc * n_frames * h->n_mels * (int64_t)sizeof(float);
    return MFCC_B200_OK;
}

int mfcc_b200_librosa_run(void *handle, const int16_t *d_x, int64_t clip_stride, int64_t n_samples,
                          int32_t n_clips, int64_t n_frames, void *d_work, float *d_out,
                          int64_t out_clip_stride, const mfcc_b200_librosa_taps *taps, void *stream)
{
    LrHandle *h = static_cast<LrHandle *>(handle);
    if (!h) return fail_ext(MFCC_B200_EINVAL, "null handle");
    if (n_clips < 0 || n_frames < 0 || n_samples < 0) return fail_ext(MFCC_B200_EINVAL, "negative size");
    if (n_clips == 0 || n_frames == 0) return MFCC_B200_OK;
    if (!d_x || !d_work || !d_out) return fail_ext(MFCC_B200_EINVAL, "null buffer");
    if (n_samples < 1) return fail_ext(MFCC_B200_EINVAL, "a clip needs at least one sample");
    if (clip_stride < n_samples) return fail_ext(MFCC_B200_EINVAL, "clip_stride < n_samples");
    if (out_clip_stride < n_frames * h->n_mfcc) return fail_ext(MFCC_B200_EINVAL, "out_clip_stride too small");
    if (h->center && n_frames > 1 + n_samples / h->hop)
        return fail_ext(MFCC_B200_EINVAL, "n_frames exceeds 1 + n_samples / hop (librosa's frame count)");
    if (taps && n_clips != 1) return fail_ext(MFCC_B200_EINVAL, "taps require n_clips == 1");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int *clipmax = static_cast<int *>(d_work);
    float *logmel = reinterpret_cast<float *>(static_cast<char *>(d_work) + lr_max_bytes(n_clips));

    lr_init_max<<<(n_clips + 255) / 256, 256, 0, st>>>(clipmax, n_clips);
    mfccb_host::count_launch();
    // float4 stores of whole rows: the output rows must be 16-byte aligned for lr_dct_c
    const bool rows16 = (reinterpret_cast<uintptr_t>(d_out) & 15) == 0 && (out_clip_stride % 4) == 0;
    const long long cc = h->chunk_clips(n_clips, n_frames);
    for (long long c0 = 0; c0 < n_clips; c0 += cc) {
        const int nc = (int)((n_clips - c0) < cc ? (n_clips - c0) : cc);
        const long long total = (long long)nc * n_frames;
        const int16_t *x = d_x + c0 * clip_stride;
        float *out = d_out + c0 * out_clip_stride;
        long long g1 = (long long)h->sms * h->ctas1;
        if (g1 > (total + LR_WARPS - 1) / LR_WARPS) g1 = (total + LR_WARPS - 1) / LR_WARPS;
        h->logmel<<<(unsigned)g1, LR_THREADS, h->smem1, st>>>(x, clip_stride, n_samples, nc, n_frames, h->hop, h->center,
                                                      h->n_mels, h->amin, h->dev, logmel, clipmax + c0,
                                                      taps ? taps->d_power : nullptr);
        mfccb_host::count_launch();
        if (h->dct_param && rows16 && h->dct_tensor) {
            long long g2 = (total + 32 * DM_WARPS - 1) / (32 * DM_WARPS);
            if (g2 > (long long)h->sms * 2) g2 = (long long)h->sms * 2;
            lr_dct_mma<<<(unsigned)g2, DM_WARPS * 32, 16 * 4 * 32 * sizeof(float4), st>>>(
                logmel, clipmax + c0, total, n_frames, h->top_db, h->dev.dct, out, out_clip_stride);
        } else if (h->dct_param && rows16) {
            long long g2 = (total + 32 * DC_WARPS - 1) / (32 * DC_WARPS);
            if (g2 > (long long)h->sms * 3) g2 = (long long)h->sms * 3;
            lr_dct_c<<<(unsigned)g2, DC_WARPS * 32, DC_WARPS * DC_MELS * DC_PITCH * sizeof(float), st>>>(
                logmel, clipmax + c0, total, n_frames, h->top_db, *h->dct_param, out, out_clip_stride);
        } else {
            long long g2 = (total + 31) / 32;
            if (g2 > (long long)h->sms * 8) g2 = (long long)h->sms * 8;
            const size_t dsm = ((((size_t)h->n_mels * h->n_mfcc + 3) & ~(size_t)3) + 8 * 4 * (size_t)h->n_mels) * sizeof(float);
            lr_dct<<<(unsigned)g2, 256, dsm, st>>>(logmel, clipmax + c0, nc, n_frames, h->n_mels, h->n_mfcc, h->top_db, h->dev,
                                                   out, out_clip_stride);
        }
        mfccb_host::count_launch();
    }
    if (taps && taps->d_logmel)
        cudaMemcpyAsync(taps->d_logmel, logmel, (size_t)n_frames * h->n_mels * sizeof(float), cudaMemcpyDeviceToDevice, st);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail_ext(MFCC_B200_ECUDA, cudaGetErrorString(e));
    return MFCC_B200_OK;
}

int mfcc_b200_librosa_run_static(void *handle, const int16_t *d_x, int64_t clip_stride, int64_t n_samples,
                                 int32_t n_clips, int64_t n_frames, float *d_out, int64_t out_clip_stride, void *stream)
{
    LrHandle *h = static_cast<LrHandle *>(handle);
    if (!h) return fail_ext(MFCC_B200_EINVAL, "null handle");
    if (n_clips < 0 || n_frames < 0 || n_samples < 0) return fail_ext(MFCC_B200_EINVAL, "negative size");
    if (!h->static_scratch) {
        static thread_local std::string msg;
        msg = "mfcc_b200_librosa_run_static: cudaGetSymbolAddress failed at create: " + h->static_err;
        return fail_ext(MFCC_B200_ECUDA, msg.c_str());
    }
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev != h->device)
        return fail_ext(MFCC_B200_EINVAL, "mfcc_b200_librosa_run_static: the handle's device must be current");
    // the whole batch in one chunk: maxima + rows must fit the static array (no overflow: sizes are
    // bounded before they are multiplied)
    const int64_t row_bytes = (int64_t)h->n_mels * (int64_t)sizeof(float);
    const int64_t cap_rows = ((int64_t)MFCC_B200_LIBROSA_STATIC_BYTES - lr_max_bytes(n_clips)) / row_bytes;
    if (n_clips > (int32_t)(MFCC_B200_LIBROSA_STATIC_BYTES / 8) || cap_rows < 0 ||
        (n_clips > 0 && n_frames > cap_rows / n_clips)) {
        static thread_local std::string msg;
        msg = "mfcc_b200_librosa_run_static: " + std::to_string(n_clips) + " clips x " + std::to_string(n_frames) +
              " frames need more than the " + std::to_string((long long)MFCC_B200_LIBROSA_STATIC_BYTES) +
              " bytes of static scratch (n_clips * n_frames * n_mels * 4 + 256-byte-rounded 4 * n_clips); "
              "use mfcc_b200_librosa_run with a caller-owned workspace";
        return fail_ext(MFCC_B200_EINVAL, msg.c_str());
    }
    // same kernels as mfcc_b200_librosa_run, with the chunk covering the whole batch; lr_init_max clears
    // the maxima (the only part of the scratch that must not carry over), every row is written before it is read
    const long long keep = h->chunk_bytes;
    h->chunk_bytes = 1LL << 60;
    const int rc = mfcc_b200_librosa_run(handle, d_x, clip_stride, n_samples, n_clips, n_frames, h->static_scratch, d_out,
                                         out_clip_stride, nullptr, stream);
    h->chunk_bytes = keep;
    return rc;
}

}  // extern "C"
