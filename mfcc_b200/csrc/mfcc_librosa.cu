// The librosa parameter set of the reference's comparison script, on the GPU
// (/root/reference/software/genlibrosa.py:21-24:
//  librosa.feature.mfcc(samples, sr=16000, hop_length=170, n_mfcc=32), i.e. centred
//  reflect-padded 2048-point frames, Hann window, power spectrum, 128 Slaney mel filters,
//  10*log10 with an 80 dB floor under the clip's maximum, orthonormal DCT-II).
//
// Two kernels, because the dB floor depends on the maximum over the whole clip:
//   lr_logmel  one CTA per frame at a time (persistent, grid-stride over clip*frame):
//              2048 real samples -> packed 1024-point complex FFT (5 radix-4 Stockham
//              passes in shared memory) -> real-FFT split + |X|^2 -> sparse mel filters ->
//              10*log10 -> workspace, clip maximum by atomicMax on an order-preserving key
//   lr_dct     one warp per frame: clamp to max - top_db, n_mfcc dot products of n_mels
// Float32 throughout (librosa computes the STFT in complex64 for float32 audio).
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>
#include <algorithm>
#include <cmath>
#include <cuda_runtime.h>

#include "../../include/mfcc_b200.h"

namespace mfccb_host {
int fail_ext(int code, const char *msg);
void count_launch();
}
using mfccb_host::fail_ext;

namespace {

constexpr int LR_NFFT = 2048;
constexpr int LR_HALF = LR_NFFT / 2;          // length of the packed complex transform
constexpr int LR_BINS = LR_HALF + 1;          // 1025 real-spectrum bins
constexpr int LR_THREADS = 256;               // one radix-4 butterfly per thread and pass
constexpr int LR_MAX_MELS = 128;
constexpr int LR_MAX_MFCC = 64;

struct LrDev {                                 // device-resident tables of one handle
    float *window;                             // [2048], includes the 1/32768 input scale
    float2 *twp;                               // [3][340] per-pass twiddles, see LrSmem
    float2 *tw2048;                            // [513]  exp(-2 pi i k / 2048)
    int *mel_id, *mel_base;                    // [2][64] filter of thread group g in round 0 / 1 (-1: none), its first bin
    float *melT;                               // [mel_r0 + mel_r1][256] weight of each thread at each step
    int mel_r0, mel_r1;                        // steps of round 0 / 1 (longest span / 4, rounded up)
    float *dct;                                // [n_mels][n_mfcc] (transposed for the lanes)
};

constexpr int LR_TWP = 340;                    // 4 + 16 + 64 + 256 twiddles per multiple r = 1..3
constexpr int LR_BUF = 1280;                   // 1024 values + room for the padded layouts

struct LrSmem {                                // followed by the mel weight table (dynamic)
    float2 a[LR_BUF];
    float2 b[LR_BUF];
    float2 twp[3 * LR_TWP];                    // per pass, contiguous in k: W^(k r 256/NS)
    float red[8];
};

__device__ __forceinline__ float2 cmul(float2 x, float2 w) {
    return make_float2(x.x * w.x - x.y * w.y, x.x * w.y + x.y * w.x);
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

// Buffer layouts: value i lives at i + C*(i >> 4).  Any C keeps the contiguous reads of
// the next pass conflict-free (aligned runs of 16 stay contiguous); C = 1 spreads the
// stride-4 stores of the first pass over all banks, C = 4 those of the second pass.
template <int C>
__device__ __forceinline__ int padi(int i) { return C ? i + C * (i >> 4) : i; }

__host__ __device__ constexpr int lr_tw_off(int ns) { return ns == 4 ? 0 : ns == 16 ? 4 : ns == 64 ? 20 : 84; }

// One radix-4 Stockham pass: thread j combines in[j + 256 r], r = 0..3, with twiddles
// exp(-2 pi i k r / (4 NS)), k = j mod NS, into out[4 (j - k) + k + r NS].  All eight
// addresses are one base per side plus constants (the padded layouts keep that true).
template <int NS, int CIN, int COUT>
__device__ __forceinline__ void stockham_pass(const float2 *__restrict__ in, float2 *__restrict__ out,
                                              const float2 *__restrict__ twp, int j) {
    const int k = j & (NS - 1);
    const float2 *ip = in + padi<CIN>(j);
    constexpr int IS = 256 + 16 * CIN;                       // (j + 256 r) >> 4 = (j >> 4) + 16 r
    float2 u0 = ip[0], u1 = ip[IS], u2 = ip[2 * IS], u3 = ip[3 * IS];
    if (NS > 1) {
        const float2 *t = twp + lr_tw_off(NS) + k;
        u1 = cmul(u1, t[0]);
        u2 = cmul(u2, t[LR_TWP]);
        u3 = cmul(u3, t[2 * LR_TWP]);
    }
    const float2 v0 = make_float2(u0.x + u2.x, u0.y + u2.y);
    const float2 v1 = make_float2(u0.x - u2.x, u0.y - u2.y);
    const float2 v2 = make_float2(u1.x + u3.x, u1.y + u3.y);
    const float2 v3 = make_float2(u1.y - u3.y, u3.x - u1.x);       // -i (u1 - u3)
    const int j0 = ((j - k) << 2) + k;
    // NS < 16: the four outputs share one 16-group, so the pad term is that of j0
    float2 *op = out + (NS >= 16 ? padi<COUT>(j0) : j0 + COUT * (j0 >> 4));
    constexpr int OS = NS >= 16 ? NS + COUT * (NS / 16) : NS;
    op[0] = make_float2(v0.x + v2.x, v0.y + v2.y);
    op[OS] = make_float2(v1.x + v3.x, v1.y + v3.y);
    op[2 * OS] = make_float2(v0.x - v2.x, v0.y - v2.y);
    op[3 * OS] = make_float2(v1.x - v3.x, v1.y - v3.y);
}

// d_x [n_clips][clip_stride] int16; logmel [n_clips][n_frames][n_mels]; clipmax [n_clips] keys
// TR0/TR1: compile-time mel step counts (0 = take them from the tables at run time)
template <int TR0, int TR1>
__global__ void __launch_bounds__(LR_THREADS, 4)
lr_logmel(const int16_t *__restrict__ d_x, long long clip_stride, long long n_samples, int n_clips,
          long long n_frames, int hop, int center, int n_mels, float amin, LrDev t,
          float *__restrict__ logmel, int *__restrict__ clipmax, float *__restrict__ power_tap)
{
    extern __shared__ __align__(16) unsigned char lr_smem_raw[];
    LrSmem &sm = *reinterpret_cast<LrSmem *>(lr_smem_raw);
    float *wT = reinterpret_cast<float *>(lr_smem_raw + sizeof(LrSmem));    // [R0 + R1][256]
    const int tid = threadIdx.x;
    const int R0 = TR0 ? TR0 : t.mel_r0, R1 = TR1 ? TR1 : t.mel_r1;
    for (int i = tid; i < 3 * LR_TWP; i += LR_THREADS) sm.twp[i] = t.twp[i];
    for (int i = tid; i < (R0 + R1) * LR_THREADS; i += LR_THREADS) wT[i] = t.melT[i];
    // the mel walk reads a few zero-weighted words past bin 1024: they must be finite
    for (int i = tid; i < LR_BUF; i += LR_THREADS) sm.a[i] = sm.b[i] = make_float2(0.0f, 0.0f);
    __syncthreads();
    float *pw = reinterpret_cast<float *>(sm.a);               // [1025] after the transform

    // frame-invariant per-thread values: window pairs, split twiddles, mel assignment
    float2 win[4], tws[2];
#pragma unroll
    for (int r = 0; r < 4; r++) win[r] = __ldg(reinterpret_cast<const float2 *>(t.window) + tid + 256 * r);
#pragma unroll
    for (int r = 0; r < 2; r++) tws[r] = __ldg(t.tw2048 + tid + 256 * r);
    const int grp = tid >> 2, q = tid & 3;
    const int m0 = __ldg(t.mel_id + grp), m1 = __ldg(t.mel_id + 64 + grp);       // -1: none
    const float *pw0 = pw + __ldg(t.mel_base + grp) + q, *pw1 = pw + __ldg(t.mel_base + 64 + grp) + q;
    const float *wT0 = wT + tid, *wT1 = wT + R0 * LR_THREADS + tid;

    // rows whose samples can be fetched as aligned pairs (frame starts are even then)
    const bool pairs = ((reinterpret_cast<uintptr_t>(d_x) & 3) == 0) && ((clip_stride & 1) == 0) &&
                       ((hop & 1) == 0);

    int clip = (int)((long long)blockIdx.x / n_frames);
    long long f = (long long)blockIdx.x - (long long)clip * n_frames;
    while (clip < n_clips) {
        const long long item = (long long)clip * n_frames + f;
        const int16_t *x = d_x + (long long)clip * clip_stride;
        const long long s0 = f * hop - (center ? LR_HALF : 0);

        // windowed samples, even/odd packed: z[n] = (y[2n] w[2n], y[2n+1] w[2n+1])
        const bool inside = s0 >= 0 && s0 + LR_NFFT <= n_samples;
        if (inside && pairs) {
            const int *xp = reinterpret_cast<const int *>(x + s0) + tid;
#pragma unroll
            for (int r = 0; r < 4; r++) {
                const int v = __ldg(xp + 256 * r);
                sm.a[tid + 256 * r] = make_float2((float)(short)(v & 0xffff) * win[r].x, (float)(v >> 16) * win[r].y);
            }
        } else {
#pragma unroll
            for (int r = 0; r < 4; r++) {
                const int n = tid + 256 * r;
                const long long i0 = s0 + 2 * n, i1 = i0 + 1;
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
                sm.a[n] = make_float2(y0 * win[r].x, y1 * win[r].y);
            }
        }
        __syncthreads();
        stockham_pass<1, 0, 1>(sm.a, sm.b, sm.twp, tid);
        __syncthreads();
        stockham_pass<4, 1, 4>(sm.b, sm.a, sm.twp, tid);
        __syncthreads();
        stockham_pass<16, 4, 0>(sm.a, sm.b, sm.twp, tid);
        __syncthreads();
        stockham_pass<64, 0, 0>(sm.b, sm.a, sm.twp, tid);
        __syncthreads();
        stockham_pass<256, 0, 0>(sm.a, sm.b, sm.twp, tid);
        __syncthreads();

        // real-FFT split: X[k] = F + W^k G, X[1024-k] = conj(F - W^k G); power into pw (= sm.a)
#pragma unroll
        for (int r = 0; r < 2; r++) {
            const int k = tid + 256 * r;
            const float2 za = sm.b[k], zb = sm.b[(LR_HALF - k) & (LR_HALF - 1)];
            const float2 F = make_float2(0.5f * (za.x + zb.x), 0.5f * (za.y - zb.y));
            const float2 G = make_float2(0.5f * (za.y + zb.y), -0.5f * (za.x - zb.x));
            const float2 T = cmul(G, tws[r]);
            const float pr = F.x + T.x, pi = F.y + T.y, qr = F.x - T.x, qi = F.y - T.y;
            pw[k] = pr * pr + pi * pi;
            pw[LR_HALF - k] = qr * qr + qi * qi;
        }
        if (tid == 0) {                                         // k = 512: W^512 = -i
            const float2 z = sm.b[512];
            pw[512] = z.x * z.x + z.y * z.y;
        }
        __syncthreads();
        if (power_tap) {
            float *dst = power_tap + item * LR_BINS;
            for (int k = tid; k < LR_BINS; k += LR_THREADS) dst[k] = pw[k];
        }

        // mel filters: four threads share a filter (bins lo+q, lo+q+4, ...); group g takes
        // filter g and then filter n_mels-1-g, so narrow low filters pair with wide high
        // ones; weights sit in shared memory as [step][thread], zero beyond a filter's span
        float a0 = 0.0f, a1 = 0.0f;
#pragma unroll(TR0 ? TR0 : 4)
        for (int i = 0; i < R0; i++) a0 = fmaf(wT0[i * LR_THREADS], pw0[4 * i], a0);
#pragma unroll(TR1 ? TR1 : 4)
        for (int i = 0; i < R1; i++) a1 = fmaf(wT1[i * LR_THREADS], pw1[4 * i], a1);
        a0 += __shfl_xor_sync(0xffffffffu, a0, 1);
        a1 += __shfl_xor_sync(0xffffffffu, a1, 1);
        a0 += __shfl_xor_sync(0xffffffffu, a0, 2);
        a1 += __shfl_xor_sync(0xffffffffu, a1, 2);
        float db = -3.0e38f;
        {
            float *lm = logmel + item * n_mels;
            const float v0 = 3.0102999566398120f * __log2f(fmaxf(a0, amin));    // 10 log10
            const float v1 = 3.0102999566398120f * __log2f(fmaxf(a1, amin));
            if (m0 >= 0) db = v0;
            if (m1 >= 0) db = fmaxf(db, v1);
            if (q == 0 && m0 >= 0) lm[m0] = v0;
            if (q == 1 && m1 >= 0) lm[m1] = v1;
        }
        for (int o = 16; o; o >>= 1) db = fmaxf(db, __shfl_xor_sync(0xffffffffu, db, o));
        if ((tid & 31) == 0) sm.red[tid >> 5] = db;
        __syncthreads();                                        // also: pw is free again
        if (tid == 0) {
            float v = sm.red[0];
            for (int i = 1; i < 8; i++) v = fmaxf(v, sm.red[i]);
            atomicMax(clipmax + clip, float_key(v));
        }
        f += gridDim.x;
        if (f >= n_frames) {
            clip += (int)(f / n_frames);
            f %= n_frames;
        }
    }
}

// logmel [n_clips][n_frames][n_mels] -> out [n_clips][n_frames][n_mfcc] (row pitch out_clip_stride)
__global__ void __launch_bounds__(256)
lr_dct(const float *__restrict__ logmel, const int *__restrict__ clipmax, int n_clips, long long n_frames,
       int n_mels, int n_mfcc, float top_db, LrDev t, float *__restrict__ out, long long out_clip_stride)
{
    extern __shared__ float dsm[];
    float *D = dsm;                                            // [n_mels][n_mfcc]
    float *vbuf = dsm + n_mels * n_mfcc;                       // [8 warps][n_mels]
    for (int i = threadIdx.x; i < n_mels * n_mfcc; i += blockDim.x) D[i] = t.dct[i];
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float *v = vbuf + warp * n_mels;
    const long long total = (long long)n_clips * n_frames;
    for (long long item = (long long)blockIdx.x * 8 + warp; item < total; item += (long long)gridDim.x * 8) {
        const int clip = (int)(item / n_frames);
        const long long f = item - (long long)clip * n_frames;
        const float floor_db = top_db >= 0.0f ? key_float(clipmax[clip]) - top_db : -3.0e38f;
        for (int n = lane; n < n_mels; n += 32) v[n] = fmaxf(logmel[item * n_mels + n], floor_db);
        __syncwarp();
        for (int k = lane; k < n_mfcc; k += 32) {
            float acc = 0.0f;
            for (int n = 0; n < n_mels; n++) acc = fmaf(D[n * n_mfcc + k], v[n], acc);
            out[(long long)clip * out_clip_stride + f * n_mfcc + k] = acc;
        }
        __syncwarp();
    }
}

__global__ void lr_init_max(int *clipmax, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) clipmax[i] = (int)0x80000000;
}

struct LrHandle {
    int hop, n_mels, n_mfcc, center, sms, ctas1;
    void (*logmel)(const int16_t *, long long, long long, int, long long, int, int, int, float, LrDev, float *, int *,
                   float *);
    size_t smem1;
    float amin, top_db;
    LrDev dev;
    std::vector<void *> allocs;
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
    std::vector<float2> tw1(3 * LR_TWP), tw2(513);
    for (int r = 1; r <= 3; r++)
        for (int ns = 4; ns <= 256; ns *= 4)
            for (int k = 0; k < ns; k++) {
                const double ang = 2 * PI * k * r / (4.0 * ns);
                tw1[(r - 1) * LR_TWP + lr_tw_off(ns) + k] = make_float2((float)cos(ang), (float)-sin(ang));
            }
    for (int k = 0; k < 513; k++) tw2[k] = make_float2((float)cos(2 * PI * k / 2048), (float)-sin(2 * PI * k / 2048));
    // thread group g (four threads) owns filter g in round 0 and filter n_mels-1-g in round 1
    std::vector<int> lo(cfg->n_mels), cnt(cfg->n_mels), ids(128, -1), base(128, 0);
    for (int m = 0; m < cfg->n_mels; m++) {
        const float *row = cfg->mel + (size_t)m * LR_BINS;
        int a = 0, b = LR_BINS;
        while (a < LR_BINS && row[a] == 0.0f) a++;
        while (b > a && row[b - 1] == 0.0f) b--;
        lo[m] = a < LR_BINS ? a : 0;
        cnt[m] = b - a;
    }
    const int half_up = (cfg->n_mels + 1) / 2, half_dn = cfg->n_mels / 2;
    int r0 = 0, r1 = 0;
    for (int g = 0; g < 64; g++) {
        if (g < half_up) { ids[g] = g; base[g] = lo[g]; r0 = std::max(r0, (cnt[g] + 3) / 4); }
        if (g < half_dn) {
            const int m = cfg->n_mels - 1 - g;
            ids[64 + g] = m; base[64 + g] = lo[m]; r1 = std::max(r1, (cnt[m] + 3) / 4);
        }
    }
    h->dev.mel_r0 = r0; h->dev.mel_r1 = r1;
    std::vector<float> wT((size_t)std::max(1, r0 + r1) * LR_THREADS, 0.0f);
    for (int tid = 0; tid < LR_THREADS; tid++)
        for (int round = 0; round < 2; round++) {
            const int m = ids[64 * round + (tid >> 2)];
            if (m < 0) continue;
            const float *row = cfg->mel + (size_t)m * LR_BINS;
            const int steps = round ? r1 : r0;
            for (int i = 0; i < steps; i++) {
                const int bin = lo[m] + (tid & 3) + 4 * i;
                if (bin < lo[m] + cnt[m]) wT[(size_t)((round ? r0 : 0) + i) * LR_THREADS + tid] = row[bin];
            }
        }
    h->smem1 = sizeof(LrSmem) + (size_t)(r0 + r1) * LR_THREADS * sizeof(float);
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
    h->dev.twp = (float2 *)upload(tw1.data(), tw1.size() * sizeof(float2));
    h->dev.tw2048 = (float2 *)upload(tw2.data(), tw2.size() * sizeof(float2));
    h->dev.mel_id = (int *)upload(ids.data(), ids.size() * sizeof(int));
    h->dev.mel_base = (int *)upload(base.data(), base.size() * sizeof(int));
    h->dev.melT = (float *)upload(wT.data(), wT.size() * sizeof(float));
    if (e == cudaSuccess && h->smem1 > 200 * 1024) {
        for (void *p : h->allocs) cudaFree(p);
        delete h;
        return fail_ext(MFCC_B200_EUNSUPPORTED, "mel filters too wide for the shared-memory weight table");
    }
    // the default filter bank (128 Slaney filters at 16 kHz) has 3 + 12 steps: unrolled instantiation
    h->logmel = (r0 == 3 && r1 == 12) ? lr_logmel<3, 12> : lr_logmel<0, 0>;
    if (e == cudaSuccess) e = cudaFuncSetAttribute(h->logmel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h->smem1);
    if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&h->ctas1, h->logmel, LR_THREADS, h->smem1);
    if (e == cudaSuccess && h->ctas1 < 1) h->ctas1 = 1;
    h->dev.dct = (float *)upload(dT.data(), dT.size() * sizeof(float));
    if (e != cudaSuccess) {
        for (void *p : h->allocs) cudaFree(p);
        delete h;
        return fail_ext(MFCC_B200_ECUDA, cudaGetErrorString(e));
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

int mfcc_b200_librosa_workspace(void *handle, int32_t n_clips, int64_t n_frames, int64_t *bytes)
{
    LrHandle *h = static_cast<LrHandle *>(handle);
    if (!h || !bytes || n_clips < 0 || n_frames < 0) return fail_ext(MFCC_B200_EINVAL, "bad argument");
    *bytes = lr_max_bytes(n_clips) + (int64_t)n_clips * n_frames * h->n_mels * (int64_t)sizeof(float);
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
    const long long total = (long long)n_clips * n_frames;

    lr_init_max<<<(n_clips + 255) / 256, 256, 0, st>>>(clipmax, n_clips);
    mfccb_host::count_launch();
    long long g1 = (long long)h->sms * h->ctas1;
    if (g1 > total) g1 = total;
    h->logmel<<<(unsigned)g1, LR_THREADS, h->smem1, st>>>(d_x, clip_stride, n_samples, n_clips, n_frames, h->hop, h->center,
                                                  h->n_mels, h->amin, h->dev, logmel, clipmax,
                                                  taps ? taps->d_power : nullptr);
    mfccb_host::count_launch();
    long long g2 = (total + 7) / 8;
    if (g2 > (long long)h->sms * 8) g2 = (long long)h->sms * 8;
    const size_t dsm = ((size_t)h->n_mels * h->n_mfcc + 8 * (size_t)h->n_mels) * sizeof(float);
    lr_dct<<<(unsigned)g2, 256, dsm, st>>>(logmel, clipmax, n_clips, n_frames, h->n_mels, h->n_mfcc, h->top_db, h->dev,
                                           d_out, out_clip_stride);
    mfccb_host::count_launch();
    if (taps && taps->d_logmel)
        cudaMemcpyAsync(taps->d_logmel, logmel, (size_t)total * h->n_mels * sizeof(float), cudaMemcpyDeviceToDevice, st);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail_ext(MFCC_B200_ECUDA, cudaGetErrorString(e));
    return MFCC_B200_OK;
}

}  // extern "C"
