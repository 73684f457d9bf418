// Kernels and C ABI of the B200 MFCC path (include/mfcc_b200.h).
//
// mfcc_int_fused: the whole chain Preemph -> Frame -> WindowHamming -> FftStream ->
// PowerSpectrum -> FilterBank -> Log2Fix -> DCTStream -> Discard of
// /root/reference/mfcc/core/mfcc.py:90-104 in one launch.  A persistent CTA (12 warps on
// 72-frame tiles for large batches, 8 warps on 24-frame tiles otherwise: IntShape below) walks
// tiles of consecutive frames of one clip: the tile's int16 sample run arrives in
// shared memory through one TMA bulk copy (prefetched while the previous tile is
// computed), is pre-emphasised once, and each warp then turns frames into cepstra;
// only the cepstra go back to HBM.
// mfcc_f32_fused: the float prototype of the same chain, two frames per warp at a time.
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <cmath>
#include <cuda_runtime.h>
#include <atomic>
#include <initializer_list>
#include <string>
#include <vector>

#include "../../include/mfcc_b200.h"
#include "mfcc_dev.cuh"
#include "mfcc_f32.cuh"

using namespace mfccb;

namespace {

constexpr int WARPS   = 8;
constexpr int THREADS = WARPS * 32;
// CTA shapes of the fused integer kernel.  Large batches take 72-frame tiles on 2 CTAs x 12 warps
// (72 registers per thread instead of 64 and six frames per warp between block barriers: fewer
// instructions per frame and a smaller tile prologue per frame); small inputs keep 24-frame tiles
// on 4 CTAs x 8 warps, which cut a short clip into enough tiles to fill the machine.
template <int W, int F, int C>
struct IntShape {
    static constexpr int WARPS = W, THREADS = W * 32, FT = F, CTAS = C;
    static constexpr int RAWN = ((HOP * (F - 1) + NFFT + 15) + 7) & ~7;   // samples per tile buffer
};
#ifndef MFCC_MEL_ONEWIDE
#define MFCC_MEL_ONEWIDE 1
#endif
#ifndef MFCC_INT_BIG_W
#define MFCC_INT_BIG_W 12
#endif
#ifndef MFCC_INT_BIG_FT
#define MFCC_INT_BIG_FT 72
#endif
using ShapeBig = IntShape<MFCC_INT_BIG_W, MFCC_INT_BIG_FT, 2>;
using ShapeSmall = IntShape<8, 24, 4>;
// the fp32 kernel has its own CTA shape (it is bound by shared-memory latency, not by issue)
#ifndef MFCC_F32_WARPS
#define MFCC_F32_WARPS 8
#endif
#ifndef MFCC_F32_FT
#define MFCC_F32_FT 32
#endif
#ifndef MFCC_F32_NFR
#define MFCC_F32_NFR 4               // frames a warp carries through the chain together (two per half-warp)
#endif
constexpr int WARPS_F   = MFCC_F32_WARPS;
constexpr int THREADS_F = WARPS_F * 32;
constexpr int FT_F      = MFCC_F32_FT;
constexpr int RAWN_F    = ((HOP * (FT_F - 1) + NFFT + 15) + 7) & ~7;
#ifndef MFCC_F32_CTAS
#define MFCC_F32_CTAS 2
#endif
constexpr int NFR_F = MFCC_F32_NFR;
constexpr int CTAS_PER_SM_F = MFCC_F32_CTAS;

struct RunParams {
    const int16_t *x;
    long long clip_stride, n_samples, n_frames;
    void *out;            // int16 (integer mode) or float (fp32 mode)
    long long out_clip_stride;
    const DevTables *tab;
    const mfccf::DevTablesF *ftab;
    long long n_tiles;
    int tiles_per_clip;
    int nceptrums;
    int aligned;          // rows are 16-byte aligned: TMA path
    mfcc_b200_taps taps;  // all NULL unless requested (n_clips == 1)
};

// ------------------------------------------------------------------ PTX helpers
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "MBAR_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, 0x989680;\n"
        "@p bra MBAR_DONE;\n"
        "bra MBAR_WAIT;\n"
        "MBAR_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// Orders this thread's earlier generic-proxy accesses to shared memory (made visible to it by the
// preceding __syncthreads) before later async-proxy writes (the bulk copy that refills the buffer).
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
// TMA 1-D bulk copy global -> shared, completion on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void tma_load_1d(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// sign-extending 16-bit shared load (SASS LDS.S16; plain C++ compiles to LDS.U16 + PRMT)
__device__ __forceinline__ int lds_s16(uint32_t saddr) {
    int v;
    asm volatile("ld.shared.s16 %0, [%1];" : "=r"(v) : "r"(saddr));
    return v;
}

// ------------------------------------------------------------------ tile geometry
struct Tile {
    const int16_t *clip;
    long long ci, f0, a0;    // clip index, first frame, sample index of raw[0] (multiple of 8, may be -8)
    int nf, len, lo, hi;     // frames, samples needed in raw[], TMA covers raw[lo, hi)
};

// tile `ti` of clip `ci` (the persistent kernels step (ci, ti) by the grid size without dividing).
// A tile starts FT * HOP samples after the previous one, a multiple of 8: raw[0] is always the
// sample 8 before the tile's first frame, so every bulk copy is 16-byte aligned on both sides.
static_assert((ShapeBig::FT * HOP) % 8 == 0 && (ShapeSmall::FT * HOP) % 8 == 0 && (FT_F * HOP) % 8 == 0,
              "tiles must start on 16-byte boundaries of the clip");
static_assert((NFR_F == 2 || NFR_F == 4) && FT_F == NFR_F * WARPS_F && FT_F % 32 == 0, "a tile is one pass of every warp over whole 32-frame blocks");
static_assert((16 * (HOP / 2)) % 32 == 16, "frames 16 apart must lie 16 shared-memory banks apart");
constexpr int OFF0 = 8;                                      // raw[OFF0] = first sample of the tile's first frame
template <int FTT>
__device__ __forceinline__ Tile tile_geo(const RunParams &p, long long ci, int ti) {
    Tile g;
    g.ci = ci;
    g.f0 = (long long)ti * FTT;
    const long long left = p.n_frames - g.f0;
    g.nf = left < FTT ? (int)left : FTT;
    g.clip = p.x + ci * p.clip_stride;
    g.a0 = (long long)ti * (FTT * HOP) - OFF0;                // < 0 only for the first tile: the predecessor is 0
    g.len = HOP * (g.nf - 1) + NFFT + OFF0;                  // <= RAWN
    g.lo = ti == 0 ? OFF0 : 0;
    // the bulk copy moves whole 16-byte words that lie inside the clip
    const long long room = (p.n_samples & ~7LL) - g.a0;
    int hi = (g.len + 7) & ~7;
    if (room < hi) hi = room < g.lo ? g.lo : (int)room;
    if (!p.aligned) hi = g.lo;
    g.hi = hi;
    return g;
}

// position of a persistent CTA in the tile sequence: tile = ci * tiles_per_clip + ti
struct TileIter {
    long long tile, ci;
    int ti, dq, dr, tpc;
    __device__ __forceinline__ void init(const RunParams &p) {
        tpc = p.tiles_per_clip;
        tile = blockIdx.x;
        ci = (int)(blockIdx.x / (unsigned)tpc);
        ti = (int)(blockIdx.x % (unsigned)tpc);
        dq = (int)(gridDim.x / (unsigned)tpc);
        dr = (int)(gridDim.x % (unsigned)tpc);
    }
    __device__ __forceinline__ void next() {
        tile += gridDim.x;
        ci += dq;
        ti += dr;
        if (ti >= tpc) { ti -= tpc; ++ci; }
    }
};

// ------------------------------------------------------------------ one frame by one warp
// pre: pre-emphasised samples of the frame in shared memory (even halfword offset).
// TAPS: also store every stage's output (tp), the analogue of the bench collectors of
// mfcc/core/mfcc.py:176-183; compiled out of the production instantiation.
template <int NF, bool TAPS>
__device__ __forceinline__ void frame_warp(const int16_t *pre, unsigned *xbuf, const DevTables *tab, int lane,
                                           int16_t *out, int nceptrums, const mfcc_b200_taps &tp, long long k)
{
    // WindowHamming: top 16 bits of signed16 * unsigned9 = (y * curve) >> 9   window.py:74-75,84
    const int bt = (int)(__brev((unsigned)lane) >> 27);
    const uint32_t pre_s = smem_u32(pre + bt);
    int xr[16];
    unsigned cv[16];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const uint4 c4 = reinterpret_cast<const uint4 *>(tab->curveL)[q * 32 + lane];
        cv[4 * q] = c4.x; cv[4 * q + 1] = c4.y; cv[4 * q + 2] = c4.z; cv[4 * q + 3] = c4.w;
    }
#pragma unroll
    for (int l = 0; l < 16; ++l) {
        const int n = bt + 32 * brev4(l);
        xr[l] = (lds_s16(pre_s + 64u * brev4(l)) * (int)cv[l]) >> 9;
        if (TAPS && tp.d_window) tp.d_window[k * NFFT + n] = (int16_t)xr[l];
    }
    int re[8], im[8];
    fft512_warp(xr, xbuf, tab, lane, re, im);
    // PowerSpectrum: top 30 bits of the 32-bit sum   pow2.py:32, :64
    // where the lane's eight power values go in the padded layout: two 128-bit loads
    const uint4 pa0 = tab->pstore[lane], pa1 = tab->pstore[32 + lane];
    const unsigned pst[8] = {pa0.x, pa0.y, pa0.z, pa0.w, pa1.x, pa1.y, pa1.z, pa1.w};
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int b = lane + 32 * i;
        const unsigned pw = ((unsigned)(re[i] * re[i]) + (unsigned)(im[i] * im[i])) >> 2;
        xbuf[pst[i]] = pw;
        if (TAPS && tp.d_fft) { tp.d_fft[(k * NBINS + b) * 2] = (int16_t)re[i]; tp.d_fft[(k * NBINS + b) * 2 + 1] = (int16_t)im[i]; }
        if (TAPS && tp.d_power) tp.d_power[k * NBINS + b] = pw;
    }
    __syncwarp();
    const unsigned mel = mel_warp<NF>(xbuf, tab, lane);
    const unsigned lg = log2fix_tab(mel, tab->logfrac);
    const int c = dct_warp<NF>(lg, tab, lane);
    __syncwarp();   // xbuf is rewritten by the next frame
    if (TAPS && lane < NF) {
        if (tp.d_mel)  tp.d_mel[k * NF + lane] = (uint16_t)mel;
        if (tp.d_log2) tp.d_log2[k * NF + lane] = (uint16_t)lg;
        if (tp.d_dct)  tp.d_dct[k * NF + lane] = (int16_t)c;
    }
    if (lane < nceptrums) out[lane] = (int16_t)c;            // Discard, discard.py:27-55
}

// ------------------------------------------------------------------ fused kernel
template <class SH>
struct Smem {
    DevTables tab;
    alignas(16) int16_t raw[SH::RAWN];
    alignas(16) int16_t pre[SH::RAWN];
    alignas(16) unsigned xbuf[SH::WARPS][XW];
    alignas(8) uint64_t mbar;
};

// 228 KB of shared memory per SM, 1 KB reserved per resident CTA
static_assert(sizeof(Smem<ShapeBig>) + 1024 <= 233472 / ShapeBig::CTAS && sizeof(Smem<ShapeSmall>) + 1024 <= 233472 / ShapeSmall::CTAS,
              "the integer kernel must keep its CTAs resident");

template <class SH>
__device__ __forceinline__ void issue_tile(const RunParams &p, const TileIter &it, Smem<SH> *sm) {
    const Tile g = tile_geo<SH::FT>(p, it.ci, it.ti);
    const uint32_t bytes = (uint32_t)(g.hi - g.lo) * 2u;
    if (bytes) {
        fence_proxy_async();
        mbar_arrive_expect_tx(&sm->mbar, bytes);
        tma_load_1d(sm->raw + g.lo, g.clip + g.a0 + g.lo, bytes, &sm->mbar);
    } else {
        mbar_arrive(&sm->mbar);
    }
}

template <int NF, bool TAPS, class SH>
__global__ void __launch_bounds__(SH::THREADS, SH::CTAS) mfcc_int_fused(const RunParams p)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    Smem<SH> *sm = reinterpret_cast<Smem<SH> *>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    {   // tables: global -> shared, once per CTA
        const int4 *src = reinterpret_cast<const int4 *>(p.tab);
        int4 *dst = reinterpret_cast<int4 *>(&sm->tab);
        for (int i = tid; i < (int)(sizeof(DevTables) / 16); i += SH::THREADS) dst[i] = src[i];
    }
    if (tid == 0) mbar_init(&sm->mbar, 1);
    __syncthreads();

    TileIter it;
    it.init(p);
    if (tid == 0 && it.tile < p.n_tiles) issue_tile(p, it, sm);
    uint32_t phase = 0;

    while (it.tile < p.n_tiles) {
        const Tile g = tile_geo<SH::FT>(p, it.ci, it.ti);
        it.next();
        // samples the bulk copy does not cover: before the clip (zeros), the ragged
        // tail and everything past the end (zeros, software/main.c:135-144)
        for (int j = tid; j < g.lo; j += SH::THREADS) sm->raw[j] = 0;
        for (int j = g.hi + tid; j < g.len; j += SH::THREADS) {
            const long long s = g.a0 + j;
            sm->raw[j] = (s >= 0 && s < p.n_samples) ? g.clip[s] : (int16_t)0;
        }
        mbar_wait(&sm->mbar, phase);
        phase ^= 1;
        __syncthreads();

        // Preemph, once per sample: y = x + (prev >> 5) - prev, 16-bit   preemph.py:18-28
        // eight samples (one 128-bit word) per thread and step; whatever lies between
        // g.len and the end of the last word is never read back
        {
            const uint4 *r128 = reinterpret_cast<const uint4 *>(sm->raw);
            const unsigned *r32 = reinterpret_cast<const unsigned *>(sm->raw);
            uint4 *p128 = reinterpret_cast<uint4 *>(sm->pre);
            const int nq = (g.len + 7) >> 3;
            for (int i = tid; i < nq; i += SH::THREADS) {
                const uint4 w = r128[i];
                int sp = i ? hi16(r32[4 * i - 1]) : 0;
                const unsigned ws[4] = {w.x, w.y, w.z, w.w};
                unsigned o[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int s0 = lo16(ws[q]), s1 = hi16(ws[q]);
                    const int y0 = s0 + (sp >> 5) - sp;
                    const int y1 = s1 + (s0 >> 5) - s0;
                    o[q] = __byte_perm((unsigned)y0, (unsigned)y1, 0x5410);
                    sp = s1;
                }
                p128[i] = make_uint4(o[0], o[1], o[2], o[3]);
            }
        }
        __syncthreads();
        // raw[] is free again: prefetch the next tile while this one is computed
        if (tid == 0 && it.tile < p.n_tiles) issue_tile(p, it, sm);
        int16_t *otile = static_cast<int16_t *>(p.out) + g.ci * p.out_clip_stride + g.f0 * p.nceptrums;
        for (int fl = warp; fl < g.nf; fl += SH::WARPS)
            frame_warp<NF, TAPS>(sm->pre + OFF0 + HOP * fl, sm->xbuf[warp], &sm->tab, lane,
                                 otile + fl * p.nceptrums, p.nceptrums, p.taps, g.f0 + fl);
        __syncthreads();
    }
}

// ------------------------------------------------------------------ fp32 fused kernel
// Same tiling as mfcc_int_fused.  There is no shared pre-emphasis pass (each lane forms
// y = x - 0.96875 xp for its own samples), so the raw tile is read during the compute
// phase and is double buffered: the bulk copy of tile i+1 is issued before tile i is
// computed.
struct SmemF {
    mfccf::DevTablesF tab;
    alignas(16) int16_t raw[2][RAWN_F];
    alignas(16) float xbuf[WARPS_F][NFR_F][mfccf::XWF];
    alignas(8) uint64_t mbar[2];          // full[buf]: the tile's samples have arrived (one arrival + the copy's bytes)
    alignas(8) uint64_t empty[2];         // empty[buf]: every warp is done with the tile in raw[buf] (WARPS_F arrivals)
};

static_assert(sizeof(SmemF) + 1024 <= 233472 / CTAS_PER_SM_F, "the fp32 kernel must keep CTAS_PER_SM_F CTAs resident");

__device__ __forceinline__ void issue_tile_f(const RunParams &p, const TileIter &it, SmemF *sm, int buf) {
    const Tile g = tile_geo<FT_F>(p, it.ci, it.ti);
    const uint32_t bytes = (uint32_t)(g.hi - g.lo) * 2u;
    if (bytes) {
        fence_proxy_async();
        mbar_arrive_expect_tx(&sm->mbar[buf], bytes);
        tma_load_1d(sm->raw[buf] + g.lo, g.clip + g.a0 + g.lo, bytes, &sm->mbar[buf]);
    } else {
        mbar_arrive(&sm->mbar[buf]);
    }
}

template <int NF>
__global__ void __launch_bounds__(THREADS_F, CTAS_PER_SM_F) mfcc_f32_fused(const RunParams p)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    SmemF *sm = reinterpret_cast<SmemF *>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    {
        const int4 *src = reinterpret_cast<const int4 *>(p.ftab);
        int4 *dst = reinterpret_cast<int4 *>(&sm->tab);
        for (int i = tid; i < (int)(sizeof(mfccf::DevTablesF) / 16); i += THREADS_F) dst[i] = src[i];
    }
    // The padded power area has gaps no frame ever writes; the mel walk reads them with
    // zero weights, so they must hold finite numbers (0 * NaN would poison the sum).
    for (int i = tid; i < WARPS_F * NFR_F * mfccf::XWF; i += THREADS_F) (&sm->xbuf[0][0][0])[i] = 0.0f;
    if (tid == 0) {
        mbar_init(&sm->mbar[0], 1); mbar_init(&sm->mbar[1], 1);
        mbar_init(&sm->empty[0], WARPS_F); mbar_init(&sm->empty[1], WARPS_F);
    }
    __syncthreads();

    // No block-wide barrier in the tile loop.  Warp 0 is also the producer: before it starts tile i it
    // waits until every warp has left tile i-1 (empty[]), fills what the bulk copy of tile i+1 does not
    // cover (zeros before a clip, the ragged tail) and issues that copy into the buffer tile i-1 used.
    // The other warps only ever wait for data (full[]), so the warps of a CTA drift up to a tile apart
    // instead of meeting twice per tile.
    TileIter ti, tn;
    ti.init(p);
    tn = ti;
    auto produce = [&](const TileIter &t, int buf) {          // warp 0, all lanes
        const Tile g = tile_geo<FT_F>(p, t.ci, t.ti);
        int16_t *raw = sm->raw[buf];
        for (int j = lane; j < g.lo; j += 32) raw[j] = 0;
        for (int j = g.hi + lane; j < g.len; j += 32) {
            const long long s = g.a0 + j;
            raw[j] = (s >= 0 && s < p.n_samples) ? g.clip[s] : (int16_t)0;
        }
        __syncwarp();
        if (lane == 0) issue_tile_f(p, t, sm, buf);           // its arrive releases the fills above
    };
    if (warp == 0 && tn.tile < p.n_tiles) produce(tn, 0);
    tn.next();
    for (int it = 0; ti.tile < p.n_tiles; ++it) {
        const int buf = it & 1;
        const Tile g = tile_geo<FT_F>(p, ti.ci, ti.ti);
        ti.next();
        const int16_t *raw = sm->raw[buf];
        if (warp == 0 && tn.tile < p.n_tiles) {
            // tile it+1 goes where tile it-1 was: completion (it-1)/2 of empty[buf^1]
            if (it >= 1) mbar_wait(&sm->empty[buf ^ 1], (uint32_t)((it - 1) >> 1) & 1u);
            produce(tn, buf ^ 1);
        }
        tn.next();
        mbar_wait(&sm->mbar[buf], (uint32_t)(it >> 1) & 1u);
        float *otile = static_cast<float *>(p.out) + g.ci * p.out_clip_stride + g.f0 * p.nceptrums;
        // a warp takes NFR_F frames together, two per half-warp: slot i = warp + WARPS_F q (q = 0..NP-1) is
        // frame 32 (i / 16) + i % 16 for the lower half-warp and the frame 16 later for the upper one (their
        // samples lie 16 banks apart); a frame past the end of the tile is computed on whatever the buffer
        // holds and not stored
        constexpr int NP = NFR_F / 2;
        const int h = lane >> 4;
        if (warp < g.nf) {
            const int16_t *fr[NP];
            float *xh[NP], *xb[NFR_F];
            int fi[NFR_F];
#pragma unroll
            for (int q = 0; q < NP; ++q) {
                const int i = warp + WARPS_F * q;
                fi[2 * q] = 32 * (i >> 4) + (i & 15);
                fi[2 * q + 1] = fi[2 * q] + 16;
                fr[q] = raw + OFF0 + HOP * (fi[2 * q] + 16 * h);
                xh[q] = sm->xbuf[warp][2 * q] + h * mfccf::XWF;
                xb[2 * q] = sm->xbuf[warp][2 * q];
                xb[2 * q + 1] = sm->xbuf[warp][2 * q + 1];
            }
            float c[NFR_F];
            mfccf::power_pair_f32<NP>(fr, xh, &sm->tab, lane);
            mfccf::cepstrum_warp_f32<NF, NFR_F>(xb, &sm->tab, lane, c);
#pragma unroll
            for (int f = 0; f < NFR_F; ++f)
                if (lane < p.nceptrums && fi[f] < g.nf) otile[fi[f] * p.nceptrums + lane] = c[f];
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm->empty[buf]);
    }
}

// ------------------------------------------------------------------ stage kernels
// Each stage of the chain on its own, for the per-stage operator API.  They share the
// device functions of the fused kernel, so a stage test also tests the fused path.
__global__ void k_preemph(const int16_t *x, int16_t *y, long long n, int prev) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int p = i ? (int)x[i - 1] : prev;
        y[i] = (int16_t)((int)x[i] + (p >> 5) - p);
    }
}

__global__ void k_frame(const int16_t *y, int16_t *fr, long long n, long long n_frames) {
    const long long total = n_frames * NFFT;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long k = i / NFFT, s = k * HOP + (i - k * NFFT);
        fr[i] = s < n ? y[s] : (int16_t)0;
    }
}

__global__ void k_window(const int16_t *fr, int16_t *w, long long n_frames, const DevTables *tab) {
    const long long total = n_frames * NFFT;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x)
        w[i] = (int16_t)(((int)fr[i] * (int)tab->curveL[curve_index((int)(i & (NFFT - 1)))]) >> 9);
}

__global__ void k_power(const int16_t *f, uint32_t *pw, long long n_frames) {
    const long long total = n_frames * NBINS;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int re = f[2 * i], im = f[2 * i + 1];
        pw[i] = ((unsigned)(re * re) + (unsigned)(im * im)) >> 2;
    }
}

__global__ void k_log2(const uint16_t *m, uint16_t *o, long long total, const DevTables *tab) {
    // through the same table as the fused kernel (read from global memory here)
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x)
        o[i] = (uint16_t)log2fix_tab(m[i], tab->logfrac);
}

// 128-bit forms of the element-wise stages (eight int16 per thread and step); the launcher
// takes them when the tensors are 16-byte aligned, else the scalar kernels above.
__device__ __forceinline__ unsigned pk2(int lo, int hi) { return __byte_perm((unsigned)lo, (unsigned)hi, 0x5410); }

__global__ void k_preemph_v(const int16_t *x, int16_t *y, long long n, int prev) {
    const long long nv = n >> 3;
    const uint4 *x4 = reinterpret_cast<const uint4 *>(x);
    uint4 *y4 = reinterpret_cast<uint4 *>(y);
    for (long long j = blockIdx.x * (long long)blockDim.x + threadIdx.x; j < nv; j += (long long)gridDim.x * blockDim.x) {
        const uint4 w = x4[j];
        int p = j ? (int)x[8 * j - 1] : prev;
        const unsigned ws[4] = {w.x, w.y, w.z, w.w};
        unsigned o[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int s0 = lo16(ws[q]), s1 = hi16(ws[q]);
            o[q] = pk2(s0 + (p >> 5) - p, s1 + (s0 >> 5) - s0);
            p = s1;
        }
        y4[j] = make_uint4(o[0], o[1], o[2], o[3]);
    }
    if (blockIdx.x == 0)                                        // ragged tail
        for (long long i = 8 * nv + threadIdx.x; i < n; i += blockDim.x) {
            const int p = i ? (int)x[i - 1] : prev;
            y[i] = (int16_t)((int)x[i] + (p >> 5) - p);
        }
}

// hop and nfft are even: a frame starts on a 32-bit word of y
__global__ void k_frame_v(const int16_t *y, int16_t *fr, long long n, long long n_frames) {
    const long long total = n_frames * (NFFT / 8);
    const unsigned *y32 = reinterpret_cast<const unsigned *>(y);
    uint4 *f4 = reinterpret_cast<uint4 *>(fr);
    for (long long j = blockIdx.x * (long long)blockDim.x + threadIdx.x; j < total; j += (long long)gridDim.x * blockDim.x) {
        const long long k = j / (NFFT / 8), s = k * HOP + 8 * (j - k * (NFFT / 8));
        unsigned o[4];
        if (s + 8 <= n) {
#pragma unroll
            for (int q = 0; q < 4; ++q) o[q] = y32[(s >> 1) + q];
        } else {
#pragma unroll
            for (int q = 0; q < 4; ++q)
                o[q] = pk2(s + 2 * q < n ? (int)y[s + 2 * q] : 0, s + 2 * q + 1 < n ? (int)y[s + 2 * q + 1] : 0);
        }
        f4[j] = make_uint4(o[0], o[1], o[2], o[3]);
    }
}

__global__ void k_window_v(const int16_t *fr, int16_t *w, long long n_frames, const DevTables *tab) {
    __shared__ uint4 curve4[NFFT / 8];
    for (int i = threadIdx.x; i < NFFT; i += blockDim.x)
        reinterpret_cast<uint16_t *>(curve4)[i] = (uint16_t)tab->curveL[curve_index(i)];
    __syncthreads();
    const long long total = n_frames * (NFFT / 8);
    const uint4 *f4 = reinterpret_cast<const uint4 *>(fr);
    uint4 *w4 = reinterpret_cast<uint4 *>(w);
    for (long long j = blockIdx.x * (long long)blockDim.x + threadIdx.x; j < total; j += (long long)gridDim.x * blockDim.x) {
        const uint4 a = f4[j], c = curve4[j & (NFFT / 8 - 1)];
        const unsigned as[4] = {a.x, a.y, a.z, a.w}, cs[4] = {c.x, c.y, c.z, c.w};
        unsigned o[4];
#pragma unroll
        for (int q = 0; q < 4; ++q)
            o[q] = pk2((lo16(as[q]) * (int)(cs[q] & 0xffffu)) >> 9, (hi16(as[q]) * (int)(cs[q] >> 16)) >> 9);
        w4[j] = make_uint4(o[0], o[1], o[2], o[3]);
    }
}

__global__ void k_power_v(const int16_t *f, uint32_t *pw, long long n_frames) {
    const long long total = n_frames * (NBINS / 4);
    const uint4 *f4 = reinterpret_cast<const uint4 *>(f);
    uint4 *p4 = reinterpret_cast<uint4 *>(pw);
    for (long long j = blockIdx.x * (long long)blockDim.x + threadIdx.x; j < total; j += (long long)gridDim.x * blockDim.x) {
        const uint4 a = f4[j];
        const unsigned as[4] = {a.x, a.y, a.z, a.w};
        unsigned o[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int re = lo16(as[q]), im = hi16(as[q]);
            o[q] = ((unsigned)(re * re) + (unsigned)(im * im)) >> 2;
        }
        p4[j] = make_uint4(o[0], o[1], o[2], o[3]);
    }
}

__global__ void k_log2_v(const uint16_t *m, uint16_t *o, long long total, const DevTables *tab) {
    __shared__ uint16_t frac[2048];
    for (int i = threadIdx.x; i < 2048 / 8; i += blockDim.x)
        reinterpret_cast<uint4 *>(frac)[i] = reinterpret_cast<const uint4 *>(tab->logfrac)[i];
    __syncthreads();
    const long long nv = total >> 3;
    const uint4 *m4 = reinterpret_cast<const uint4 *>(m);
    uint4 *o4 = reinterpret_cast<uint4 *>(o);
    for (long long j = blockIdx.x * (long long)blockDim.x + threadIdx.x; j < nv; j += (long long)gridDim.x * blockDim.x) {
        const uint4 a = m4[j];
        const unsigned as[4] = {a.x, a.y, a.z, a.w};
        unsigned r[4];
#pragma unroll
        for (int q = 0; q < 4; ++q)
            r[q] = log2fix_tab(as[q] & 0xffffu, frac) | (log2fix_tab(as[q] >> 16, frac) << 16);
        o4[j] = make_uint4(r[0], r[1], r[2], r[3]);
    }
    if (blockIdx.x == 0)
        for (long long i = 8 * nv + threadIdx.x; i < total; i += blockDim.x) o[i] = (uint16_t)log2fix_tab(m[i], frac);
}

struct StageSmem {
    DevTables tab;
    alignas(16) unsigned xbuf[WARPS][XW];
};

// warp-per-frame stages: FFT, MEL, DCT
template <int STAGE, int NF>
__global__ void __launch_bounds__(THREADS) k_warp_stage(const void *in, void *out, long long n_frames, const DevTables *gtab)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    StageSmem *sm = reinterpret_cast<StageSmem *>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    {
        const int4 *src = reinterpret_cast<const int4 *>(gtab);
        int4 *dst = reinterpret_cast<int4 *>(&sm->tab);
        for (int i = tid; i < (int)(sizeof(DevTables) / 16); i += THREADS) dst[i] = src[i];
    }
    __syncthreads();
    unsigned *xbuf = sm->xbuf[warp];
    for (long long k = blockIdx.x * (long long)WARPS + warp; k < n_frames; k += (long long)gridDim.x * WARPS) {
        if (STAGE == MFCC_B200_STAGE_FFT) {
            const int16_t *w = reinterpret_cast<const int16_t *>(in) + k * NFFT;
            int16_t *o = reinterpret_cast<int16_t *>(out) + k * NBINS * 2;
            const int bt = (int)(__brev((unsigned)lane) >> 27);
            int xr[16], re[8], im[8];
#pragma unroll
            for (int l = 0; l < 16; ++l) xr[l] = w[bt + 32 * brev4(l)];
            fft512_warp(xr, xbuf, &sm->tab, lane, re, im);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int b = lane + 32 * i;
                o[2 * b] = (int16_t)re[i]; o[2 * b + 1] = (int16_t)im[i];
            }
        } else if (STAGE == MFCC_B200_STAGE_MEL) {
            const uint32_t *pw = reinterpret_cast<const uint32_t *>(in) + k * NBINS;
#pragma unroll
            for (int i = 0; i < 8; ++i) xbuf[sm->tab.paddr[lane + 32 * i]] = pw[lane + 32 * i];
            __syncwarp();
            const unsigned mel = mel_warp<NF>(xbuf, &sm->tab, lane);
            __syncwarp();
            if (lane < NF) reinterpret_cast<uint16_t *>(out)[k * NF + lane] = (uint16_t)mel;
        } else {
            // DCTStream's sink is signed 16 bits (dct_stream.py:11); the chain only ever
            // feeds it 15-bit log values, the stage operator may be given any int16
            const int16_t *lg = reinterpret_cast<const int16_t *>(in) + k * NF;
            const unsigned L = (unsigned)(int)lg[lane & (NF - 1)];
            const int c = dct_warp<NF>(L, &sm->tab, lane);
            if (lane < NF) reinterpret_cast<int16_t *>(out)[k * NF + lane] = (int16_t)c;
        }
    }
}

// ------------------------------------------------------------------ streaming step
// One hop per stream: state = [prev | nfft-hop raw samples], see include/mfcc_b200.h.
constexpr int STATE_N = NFFT - HOP + 2;      // 344 int16 per stream (343 used, even pitch)

struct StepSmem {
    DevTables tab;
    alignas(16) int16_t pre[WARPS][NFFT];
    alignas(16) unsigned xbuf[WARPS][XW];
    alignas(16) int16_t st[WARPS][STATE_N];      // the warps' state rows, then their new state rows
    alignas(16) int16_t hop[WARPS][HOP + 6];     // the warps' hops (row pitch 176: rows stay 16-byte aligned)
};

// A CTA takes WARPS consecutive streams at a time.  Their state rows (688 bytes each, contiguous) and
// hops (340 bytes each, contiguous) are staged through shared memory with 128-bit accesses by the
// whole CTA -- 512-byte requests per warp instead of 64-byte ones -- which matters most when `hop`
// (or `out`) is pinned HOST memory that the kernel reads over PCIe itself (MFCC.step with host
// tensors): the step is then one kernel whose PCIe reads overlap its arithmetic.
template <int NF>
__global__ void __launch_bounds__(THREADS) mfcc_int_step(int16_t *state, const int16_t *hop, int n_streams,
                                                         int16_t *out, int nceptrums, const DevTables *gtab)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    StepSmem *sm = reinterpret_cast<StepSmem *>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    {
        const int4 *src = reinterpret_cast<const int4 *>(gtab);
        int4 *dst = reinterpret_cast<int4 *>(&sm->tab);
        for (int i = tid; i < (int)(sizeof(DevTables) / 16); i += THREADS) dst[i] = src[i];
    }
    const mfcc_b200_taps none = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    constexpr int HIST = NFFT - HOP;                      // 342
    constexpr int ST_V = STATE_N / 8;                     // 43 128-bit words per state row
    const bool st16 = (reinterpret_cast<uintptr_t>(state) & 15) == 0;
    for (int sb = blockIdx.x * WARPS; sb < n_streams; sb += gridDim.x * WARPS) {
        const int rows = n_streams - sb < WARPS ? n_streams - sb : WARPS;
        __syncthreads();                                  // tables are in; the previous block's rows have left
        // ---- stage the block's state rows and hops
        {
            const int16_t *gs = state + (long long)sb * STATE_N;
            if (st16) {
                const uint4 *g4 = reinterpret_cast<const uint4 *>(gs);
                uint4 *s4 = reinterpret_cast<uint4 *>(&sm->st[0][0]);
                for (int i = tid; i < rows * ST_V; i += THREADS) s4[i] = g4[i];
            } else {
                for (int i = tid; i < rows * STATE_N; i += THREADS) (&sm->st[0][0])[i] = gs[i];
            }
            const int16_t *gh = hop + (long long)sb * HOP;
            const int nh = rows * HOP;                    // halfwords, contiguous in global memory
            if ((reinterpret_cast<uintptr_t>(gh) & 15) == 0) {
                const uint4 *g4 = reinterpret_cast<const uint4 *>(gh);
                for (int i = tid; i < nh / 8; i += THREADS) {
                    const uint4 v = g4[i];
                    const unsigned w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                    for (int q = 0; q < 4; ++q) {         // a 128-bit word may straddle two rows: 32-bit stores (HOP is even)
                        const int e = 8 * i + 2 * q, r = e / HOP, c = e - r * HOP;
                        *reinterpret_cast<unsigned *>(&sm->hop[r][c]) = w[q];
                    }
                }
                for (int e = (nh / 8) * 8 + tid; e < nh; e += THREADS) sm->hop[e / HOP][e % HOP] = gh[e];
            } else {
                for (int e = tid; e < nh; e += THREADS) sm->hop[e / HOP][e % HOP] = gh[e];
            }
        }
        __syncthreads();
        if (warp < rows) {
            const int16_t *st = sm->st[warp];             // st[0] = prev, st[1..342] = history
            const int16_t *hp = sm->hop[warp];
            int16_t *pre = sm->pre[warp];
            // raw(n) for n = -1 .. 511: state then hop
            int16_t nv[11];                               // this lane's share of the new state
#pragma unroll
            for (int r = 0; r < 16; ++r) {
                const int n = lane + 32 * r;
                const int cur = n < HIST ? st[1 + n] : hp[n - HIST];
                const int prv = n <= HIST ? st[n] : hp[n - 1 - HIST];
                pre[n] = (int16_t)(cur + (prv >> 5) - prv);
            }
            // new state = raw[HOP-1 .. 511]
#pragma unroll
            for (int r = 0; r < 11; ++r) {
                const int i = lane + 32 * r;              // 0..342
                const int n = i + HOP - 1;
                nv[r] = (i <= HIST) ? (n < HIST ? st[1 + n] : hp[n - HIST]) : (int16_t)0;
            }
            __syncwarp();
#pragma unroll
            for (int r = 0; r < 11; ++r) {
                const int i = lane + 32 * r;
                if (i < STATE_N) sm->st[warp][i] = nv[r];  // element 343 is the row's padding: written as 0
            }
            frame_warp<NF, false>(pre, sm->xbuf[warp], &sm->tab, lane, out + (long long)(sb + warp) * nceptrums, nceptrums,
                                  none, 0);
        }
        __syncthreads();
        // ---- the new state rows leave
        {
            int16_t *gs = state + (long long)sb * STATE_N;
            if (st16) {
                uint4 *g4 = reinterpret_cast<uint4 *>(gs);
                const uint4 *s4 = reinterpret_cast<const uint4 *>(&sm->st[0][0]);
                for (int i = tid; i < rows * ST_V; i += THREADS) g4[i] = s4[i];
            } else {
                for (int i = tid; i < rows * STATE_N; i += THREADS) gs[i] = (&sm->st[0][0])[i];
            }
        }
    }
}

// fp32 streaming step: same state layout as the integer mode (raw samples).  A warp takes two
// streams, one per half-warp through the spectrum (power_pair_f32).
constexpr int STEP_RAW_PITCH = NFFT + 32;       // halfwords; 272 words: the two rows lie 16 banks apart
struct StepSmemF {
    mfccf::DevTablesF tab;
    alignas(16) int16_t raw[WARPS][2][STEP_RAW_PITCH];   // raw[w][h][1] = sample before the frame, frame from [2]
    alignas(16) float xbuf[WARPS][2][mfccf::XWF];
};

template <int NF>
__global__ void __launch_bounds__(THREADS) mfcc_f32_step(int16_t *state, const int16_t *hop, int n_streams,
                                                         float *out, int nceptrums, const mfccf::DevTablesF *gtab)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    StepSmemF *sm = reinterpret_cast<StepSmemF *>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    {
        const int4 *src = reinterpret_cast<const int4 *>(gtab);
        int4 *dst = reinterpret_cast<int4 *>(&sm->tab);
        for (int i = tid; i < (int)(sizeof(mfccf::DevTablesF) / 16); i += THREADS) dst[i] = src[i];
    }
    for (int i = tid; i < WARPS * 2 * mfccf::XWF; i += THREADS) (&sm->xbuf[0][0][0])[i] = 0.0f;      // see mfcc_f32_fused
    __syncthreads();
    constexpr int HIST = NFFT - HOP;                      // 342
    const int h = lane >> 4;
    for (int s0 = 2 * (blockIdx.x * WARPS + warp); s0 < n_streams; s0 += 2 * gridDim.x * WARPS) {
#pragma unroll
        for (int f = 0; f < 2; ++f) {
            const int s = s0 + f < n_streams ? s0 + f : s0;   // an odd last stream is computed twice, stored once
            int16_t *st = state + (long long)s * STATE_N;     // st[0] = prev, st[1..342] = history
            const int16_t *hp = hop + (long long)s * HOP;
            int16_t *raw = sm->raw[warp][f];
            // raw(n) for n = -1 .. 511 at raw[2 + n]
            for (int i = lane; i <= NFFT; i += 32) {
                const int n = i - 1;
                raw[2 + n] = (n < HIST) ? st[1 + n] : hp[n - HIST];
            }
        }
        __syncwarp();
        // new state = raw(HOP-1 .. 511)
#pragma unroll
        for (int f = 0; f < 2; ++f)
            if (s0 + f < n_streams) {
                int16_t *st = state + (long long)(s0 + f) * STATE_N;
                for (int i = lane; i <= HIST; i += 32) st[i] = sm->raw[warp][f][2 + i + HOP - 1];
            }
        float *xb[2] = {sm->xbuf[warp][0], sm->xbuf[warp][1]};
        float c[2];
        const int16_t *fr[1] = {sm->raw[warp][0] + h * STEP_RAW_PITCH + 2};
        float *xh[1] = {sm->xbuf[warp][0] + h * mfccf::XWF};
        mfccf::power_pair_f32<1>(fr, xh, &sm->tab, lane);
        mfccf::cepstrum_warp_f32<NF, 2>(xb, &sm->tab, lane, c);
#pragma unroll
        for (int f = 0; f < 2; ++f)
            if (lane < nceptrums && s0 + f < n_streams) out[(long long)(s0 + f) * nceptrums + lane] = c[f];
        __syncwarp();
    }
}

// ------------------------------------------------------------------ host side
thread_local std::string g_err;
std::atomic<long long> g_launches{0};

int fail(int code, const std::string &msg) { g_err = msg; return code; }
#define CUDA_TRY(expr)                                                                         \
    do {                                                                                       \
        cudaError_t e_ = (expr);                                                               \
        if (e_ != cudaSuccess)                                                                 \
            return fail(MFCC_B200_ECUDA, std::string(#expr) + ": " + cudaGetErrorString(e_));  \
    } while (0)

// Staging pipeline of mfcc_b200_run_host: chunks of clips travel host -> device on one stream,
// are computed on a second and travel back on a third, chained by events, with up to
// HOST_SLOTS chunks in flight so that both copy engines and the SMs stay busy at once.
constexpr int HOST_SLOTS_MAX = 8;
struct HostPipe {
    cudaStream_t s_in = nullptr, s_run = nullptr, s_out = nullptr;
    int16_t *d_in[HOST_SLOTS_MAX] = {nullptr};
    void *d_out[HOST_SLOTS_MAX] = {nullptr};
    cudaEvent_t ev_in[HOST_SLOTS_MAX] = {nullptr};    // chunk has arrived in d_in[slot]
    cudaEvent_t ev_run[HOST_SLOTS_MAX] = {nullptr};   // kernel done: d_in[slot] free, d_out[slot] full
    cudaEvent_t ev_out[HOST_SLOTS_MAX] = {nullptr};   // d_out[slot] has left for the host
    size_t in_cap = 0, out_cap = 0;
    int slots_alloc = 0;
    long long chunk_bytes = 16LL << 20;               // samples per chunk (mfcc_b200_host_config)
    int slots = 4;
    int fail_at_chunk = -1;                           // test hook: mfcc_b200_debug_fail_chunk
};

struct Handle {
    mfcc_b200_cfg cfg;
    DevTables *d_tab = nullptr;
    mfccf::DevTablesF *d_ftab = nullptr;
    int device = 0;
    int sm_count = 0;
    HostPipe hp;                                      // staging for mfcc_b200_run_host
    // state and staging of mfcc_b200_step_host
    cudaStream_t s_step = nullptr;
    void *s_state = nullptr, *s_out = nullptr;
    int16_t *s_hop = nullptr;
    int s_streams = 0;
};

// Which twiddle each DCT lane needs: ownership of positions through the 2NF-point
// sub-FFT (mirrors tests/lane_model.py: dct_plan).
int build_dct_plan(int nf, const int32_t *tw /* [2nf][2] */, DevTables &t) {
    int S = 0;
    while ((1 << S) < 2 * nf) ++S;
    std::vector<int> A(nf), x0(nf);
    for (int j = 0; j < nf; ++j) A[j] = 2 * j;
    for (int s = 1; s < S; ++s) {
        for (int j = 0; j < nf; ++j) {
            const int bit = (j >> (s - 1)) & 1, partner = j ^ (1 << (s - 1));
            if (!bit && A[partner] != A[j] + (1 << s)) return -1;
            x0[j] = bit ? A[partner] + (1 << (s - 1)) : A[j];
        }
        for (int j = 0; j < nf; ++j) {
            const int k = x0[j] % (1 << s);
            const int idx = k * ((2 * nf) >> s);
            t.dct_tw[s * 32 + j] = make_int2(2 * tw[2 * idx], 2 * tw[2 * idx + 1]);
        }
        A = x0;
    }
    for (int j = 0; j < nf; ++j) {
        if (A[j] != j) return -1;                   // lane j must own bin j
        t.dct_twf[j] = make_int2(2 * tw[2 * j], 2 * tw[2 * j + 1]);
    }
    return 0;
}

int build_tables(const mfcc_b200_cfg &c, DevTables &t) {
    memset(&t, 0, sizeof(t));
    const int32_t *tw = c.tw_fft;
    for (int k = 0; k < 8; ++k)
        if (tw[2 * 32 * k] != H_TWA_RE[k] || tw[2 * 32 * k + 1] != H_TWA_IM[k])
            return fail(MFCC_B200_EUNSUPPORTED, "tw_fft does not match the twiddles compiled into stages 0-3");
    auto dbl = [&](int idx) { return make_int2(2 * tw[2 * idx], 2 * tw[2 * idx + 1]); };
    for (int l = 0; l < 16; ++l) {
        t.twB[l] = dbl(16 * l);                                     // stage 4
        t.twB[16 + l] = dbl(8 * l);                                 // stage 5
        for (int mm = 0; mm < 2; ++mm) t.twB[32 + mm * 16 + l] = dbl(64 * mm + 4 * l);   // stage 6
        for (int mm = 0; mm < 4; ++mm) t.twB[64 + mm * 16 + l] = dbl(32 * mm + 2 * l);   // stage 7
    }
    // the -j rotation used for the upper half of stages 5..7 must agree with the ROM
    for (int k = 0; k < 128; ++k)
        if (tw[2 * (k + 128)] != tw[2 * k + 1] || tw[2 * (k + 128) + 1] != -tw[2 * k])
            return fail(MFCC_B200_EUNSUPPORTED, "tw_fft is not a quarter-circle decode (fft.py:55-59)");
    for (int b = 0; b < 256; ++b) t.tw8[b] = dbl(b);
    for (int n = 0; n < NFFT; ++n) {
        if (c.window_curve[n] < 0 || c.window_curve[n] > 511)
            return fail(MFCC_B200_EUNSUPPORTED, "window_curve out of the 9-bit range");
        t.curveL[curve_index(n)] = (uint32_t)c.window_curve[n];
    }
    // 32 filters: the one-wide-multiply form of mel_warp() applies when every non-closing weight of
    // segment 0 is zero (see there); otherwise the two-sum form
    bool onewide = MFCC_MEL_ONEWIDE && c.nfilters == 32;
    for (int b = 0; b < NBINS && !c.mel_close[b]; ++b)
        if (c.mel_weight[b] != 0) onewide = false;
    t.mel_onewide = onewide ? 1 : 0;
    int g = 0, first = 0, seg0_len = 0, pad_next = 0;
    bool bank_used[32] = {false};
    const int maxseg = (c.nfilters == 32) ? 21 : 42;
    for (int b = 0; b < NBINS; ++b) {
        if (c.mel_weight[b] >> 30) return fail(MFCC_B200_EUNSUPPORTED, "mel_weight wider than 30 bits");
        if (c.mel_close[b]) {
            if (g >= 40) return fail(MFCC_B200_EUNSUPPORTED, "too many mel segments");
            t.seg_close[g] = (int16_t)b;
            const int len = b - first + 1;
            // lane g-1 walks segment g; lane 0 starts at bin 0 and takes segment 0 first,
            // with its wa weights in the wb slot (both end up in filter 0)
            const int lane_of = g == 0 ? 0 : g - 1;
            const int j0 = g == 1 ? seg0_len : 0;
            if (g == 0) seg0_len = len;
            // padded power layout: segment g starts at the first free address whose bank no
            // earlier segment start took (segment 1 follows segment 0 directly: one walk)
            int base = pad_next;
            if (g != 1) {
                const int nl = c.nfilters == 32 ? 32 : 16;
                while (bank_used[base % nl]) ++base;
                bank_used[base % nl] = true;
            }
            pad_next = base + len;
            for (int j = 0; j < len; ++j) t.paddr[first + j] = (uint16_t)(base + j);
            if (g == 0) t.seg_walk[0] = (int16_t)base;
            else if (g > 1) t.seg_walk[g - 1] = (int16_t)base;
            if (j0 + len > maxseg || g > c.nfilters) return fail(MFCC_B200_EUNSUPPORTED, "mel segment too long for the kernel's table");
            for (int j = 0; j < len; ++j) {
                const bool closing = (first + j == b);
                const uint32_t w = c.mel_weight[first + j];
                const uint32_t wa = closing ? (1u << 30) : w, wb = closing ? 0u : (1u << 30) - w;
                const int st = j0 + j;
                const int at = c.nfilters == 32 ? ((st >> 1) * 32 + lane_of) * 2 + (st & 1) : st * c.nfilters + lane_of;
                if (onewide) t.melab[at] = g == 0 ? make_uint2(0u, closing ? 1u : 0u) : make_uint2(wa, 1u);
                else t.melab[at] = g == 0 ? make_uint2(0u, wa) : make_uint2(wa, wb);
            }
            first = b + 1;
            ++g;
        }
    }
    if (g != c.nfilters + 1 || first != NBINS)
        return fail(MFCC_B200_EUNSUPPORTED, "mel tables must describe nfilters+1 segments ending at the last bin");
    for (int l = 0; l < 32; ++l) {
        uint32_t ps[8];
        for (int i = 0; i < 8; ++i) ps[i] = t.paddr[l + 32 * i];
        t.pstore[l] = make_uint4(ps[0], ps[1], ps[2], ps[3]);
        t.pstore[32 + l] = make_uint4(ps[4], ps[5], ps[6], ps[7]);
    }
    if (pad_next + maxseg + 1 > XW) return fail(MFCC_B200_EUNSUPPORTED, "padded power layout does not fit the exchange buffer");
    for (int z = 0; z < 2048; ++z) t.logfrac[z] = (uint16_t)log2fix_frac(2048u + z);
    t.nseg = g;
    t.mel_lo_bit = c.mel_lo_bit;
    if (build_dct_plan(c.nfilters, c.tw_dct, t) != 0)
        return fail(MFCC_B200_EUNSUPPORTED, "internal: DCT lane plan failed");
    return MFCC_B200_OK;
}

int build_tables_f(const mfcc_b200_cfg &c, mfccf::DevTablesF &t) {
    memset(&t, 0, sizeof(t));
    const double PI = 3.14159265358979323846;
    const int nf = c.nfilters, nb = mfccf::NBINS1;
    const int maxseg = (nf == 32) ? 21 : 42;
    for (int n = 0; n < NFFT; ++n) t.window[n] = c.f_window[n];
    for (int k2 = 0; k2 < 16; ++k2)
        for (int n1 = 0; n1 < 16; ++n1)
            t.tw16[k2 * 16 + n1] = make_float2((float)cos(2 * PI * (n1 * k2) / 256), (float)-sin(2 * PI * (n1 * k2) / 256));
    for (int k = 0; k <= 128; ++k)
        t.tw512[k] = make_float2((float)cos(2 * PI * k / 512), (float)-sin(2 * PI * k / 512));
    // DCT: the kernel uses D[k][nf-1-n] = (-1)^k D[k][n]; check it on the matrix we were given
    for (int k = 0; k < nf; ++k)
        for (int n = 0; n < nf / 2; ++n) {
            const float a = c.f_dct[k * nf + n], b = c.f_dct[k * nf + nf - 1 - n];
            if (fabsf(((k & 1) ? -b : b) - a) > 1e-6f)
                return fail(MFCC_B200_EUNSUPPORTED, "f_dct is not a DCT-II matrix (mirror symmetry missing)");
            t.dct[n * 32 + k] = a;
        }
    // Mel: every bin may feed two consecutive filters.  Segment g holds the bins whose
    // filters are {g-1, g}; its bins get the rise weight of g and the fall weight of g-1.
    std::vector<int> seg(nb);
    int cur = 0;
    for (int b = 0; b < nb; ++b) {
        int lo = -1, hi = -1;
        for (int m = 0; m < nf; ++m)
            if (c.f_mel[(size_t)m * nb + b] != 0.0f) { if (lo < 0) lo = m; hi = m; }
        int g = cur;
        if (lo >= 0) {
            if (hi - lo > 1) return fail(MFCC_B200_EUNSUPPORTED, "f_mel: a bin feeds more than two consecutive filters");
            if (hi > lo) g = hi;
            else {
                // one filter only: on its falling side (weight going down, or already in
                // the next segment) the bin belongs to segment lo+1, else to segment lo
                const bool falling = b > 0 && c.f_mel[(size_t)lo * nb + b] < c.f_mel[(size_t)lo * nb + b - 1];
                g = (falling || cur == lo + 1) ? lo + 1 : (cur > lo ? cur : lo);
            }
            if (g < cur || g > lo + 1) return fail(MFCC_B200_EUNSUPPORTED, "f_mel: filters are not ordered triangles");
        }
        seg[b] = cur = g;
    }
    if (cur > nf) return fail(MFCC_B200_EUNSUPPORTED, "f_mel: too many segments");
    // padded layout: segment g (lane g-1) starts at an address whose bank no other
    // segment start uses, so the lanes' walks never collide
    std::vector<int> first(nf + 2, -1), len(nf + 2, 0), base(nf + 2, 0);
    for (int b = 0; b < nb; ++b) { if (first[seg[b]] < 0) first[seg[b]] = b; ++len[seg[b]]; }
    bool used[32] = {false};
    int next = 0;
    const int nlanes = (nf == 32) ? 32 : 16;
    for (int g = 0; g <= nf; ++g) {
        // lane g-1 walks segment g; lane 0 walks segments 0 and 1 back to back
        if ((g <= 1 ? len[0] + len[1] : len[g]) > maxseg)
            return fail(MFCC_B200_EUNSUPPORTED, "f_mel: segment too long for the kernel's table");
        int a = next;
        if (g != 1) {
            while (used[a % nlanes]) ++a;
            used[a % nlanes] = true;
        }
        base[g] = a;
        next = a + len[g];
        if (g != 1) t.seg_base[g == 0 ? 0 : g - 1] = (uint16_t)a;
    }
    if (next + maxseg > mfccf::PW_WORDS) return fail(MFCC_B200_EUNSUPPORTED, "f_mel: padded power layout does not fit");
    for (int b = 0; b < nb; ++b) {
        const int g = seg[b], j = b - first[g];
        t.paddr[b] = (uint16_t)(base[g] + j);
        const float rise = (g < nf) ? c.f_mel[(size_t)g * nb + b] : 0.0f;
        const float fall = (g >= 1) ? c.f_mel[(size_t)(g - 1) * nb + b] : 0.0f;
        if (g == 0) t.melw[j * nf + 0] = make_float2(0.0f, rise);              // rise of filter 0, in the fall slot
        else t.melw[((g == 1 ? len[0] : 0) + j) * nf + (g - 1)] = make_float2(rise, fall);
    }
    for (int k1 = 0; k1 < 8; ++k1)
        for (int l = 0; l < 16; ++l) {
            const int k = 16 * k1 + l;
            t.pst[k1 * 16 + l] = (uint32_t)t.paddr[k] | ((uint32_t)t.paddr[256 - k] << 16);
        }
    t.mel_floor = c.f_mel_floor > 0.0f ? c.f_mel_floor : 0.0f;
    // the kernel's mel walk drops the lower half-warp after MEL_J0 steps when it has nothing left
    t.mel_tail_lane = 16;
    if (nf == 32)
        for (int j = mfccf::MEL_J0; j < maxseg; ++j)
            for (int l = 0; l < 16; ++l)
                if (t.melw[j * nf + l].x != 0.0f || t.melw[j * nf + l].y != 0.0f) t.mel_tail_lane = 0;
    return MFCC_B200_OK;
}

template <typename K>
int set_smem(K kernel, size_t bytes) {
    CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
    return MFCC_B200_OK;
}

// (Re)allocate the staging pipeline; on failure everything is released again and the pipe is
// left empty (the next call starts from scratch).
void host_pipe_free(HostPipe &hp, bool streams) {
    for (int i = 0; i < HOST_SLOTS_MAX; ++i) {
        if (hp.d_in[i]) cudaFree(hp.d_in[i]);
        if (hp.d_out[i]) cudaFree(hp.d_out[i]);
        hp.d_in[i] = nullptr; hp.d_out[i] = nullptr;
        if (streams) {
            for (cudaEvent_t *ev : {&hp.ev_in[i], &hp.ev_run[i], &hp.ev_out[i]})
                if (*ev) { cudaEventDestroy(*ev); *ev = nullptr; }
        }
    }
    hp.in_cap = hp.out_cap = 0;
    hp.slots_alloc = 0;
    if (streams) {
        for (cudaStream_t *st : {&hp.s_in, &hp.s_run, &hp.s_out})
            if (*st) { cudaStreamDestroy(*st); *st = nullptr; }
    }
}

int host_pipe_alloc(HostPipe &hp, size_t in_need, size_t out_need, int slots) {
    host_pipe_free(hp, false);
    cudaError_t e = cudaSuccess;
    for (cudaStream_t *st : {&hp.s_in, &hp.s_run, &hp.s_out})
        if (e == cudaSuccess && !*st) e = cudaStreamCreateWithFlags(st, cudaStreamNonBlocking);
    for (int i = 0; i < slots && e == cudaSuccess; ++i) {
        for (cudaEvent_t *ev : {&hp.ev_in[i], &hp.ev_run[i], &hp.ev_out[i]})
            if (e == cudaSuccess && !*ev) e = cudaEventCreateWithFlags(ev, cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaMalloc(&hp.d_in[i], in_need);
        if (e == cudaSuccess) e = cudaMalloc(&hp.d_out[i], out_need);
    }
    if (e != cudaSuccess) {
        host_pipe_free(hp, false);
        return fail(MFCC_B200_ECUDA, std::string("mfcc_b200_run_host: staging buffers: ") + cudaGetErrorString(e));
    }
    hp.in_cap = in_need; hp.out_cap = out_need; hp.slots_alloc = slots;
    return MFCC_B200_OK;
}

int check_handle(void *h) {
    if (!h) return fail(MFCC_B200_EINVAL, "null handle");
    return MFCC_B200_OK;
}

inline int grid_1d(long long total, int block, int sm_count) {
    long long g = (total + block - 1) / block;
    const long long cap = (long long)sm_count * 16;
    if (g > cap) g = cap;
    return g < 1 ? 1 : (int)g;
}

}  // namespace

// shared with mfcc_post.cu
namespace mfccb_host {
int fail_ext(int code, const char *msg) { return fail(code, msg); }
void count_launch() { g_launches.fetch_add(1); }
}  // namespace mfccb_host

// =================================================================== C ABI
extern "C" {

int mfcc_b200_version(void) { return 1; }
const char *mfcc_b200_last_error(void) { return g_err.c_str(); }
int64_t mfcc_b200_launch_count(void) { return g_launches.load(); }

int mfcc_b200_create(const mfcc_b200_cfg *cfg, void **handle)
{
    if (!cfg || !handle) return fail(MFCC_B200_EINVAL, "null argument");
    *handle = nullptr;
    if (cfg->struct_size != (int32_t)sizeof(mfcc_b200_cfg)) return fail(MFCC_B200_EINVAL, "mfcc_b200_cfg size mismatch");
    if (cfg->mode != MFCC_B200_MODE_INT && cfg->mode != MFCC_B200_MODE_FP32) return fail(MFCC_B200_EINVAL, "unknown mode");
    if (cfg->mode == MFCC_B200_MODE_FP32) {
        if (!cfg->f_window || !cfg->f_mel || !cfg->f_dct) return fail(MFCC_B200_EINVAL, "fp32 mode needs f_window, f_mel and f_dct");
        if (cfg->f_preemph != 0.96875f)
            return fail(MFCC_B200_EUNSUPPORTED, "the fp32 kernels cover f_preemph = 0.96875 (notebook/MFCC-INT.ipynb cell 2)");
        if (!(cfg->f_mel_floor >= 0.0f)) return fail(MFCC_B200_EINVAL, "f_mel_floor must be >= 0");
    }
    if (cfg->nfft != NFFT || cfg->hop != HOP) return fail(MFCC_B200_EUNSUPPORTED, "kernels cover nfft=512, hop=170");
    if (cfg->nfilters != 16 && cfg->nfilters != 32) return fail(MFCC_B200_EUNSUPPORTED, "kernels cover nfilters 16 and 32");
    if (cfg->nceptrums < 1 || cfg->nceptrums > cfg->nfilters) return fail(MFCC_B200_EINVAL, "nceptrums out of range");
    if (cfg->window_shift != 9 || cfg->pow_shift != 2 || cfg->mel_wbits != 30 || cfg->mel_out_bits != 16 ||
        cfg->log_precision != 11 || cfg->log_out_bits != 15)
        return fail(MFCC_B200_EUNSUPPORTED, "kernels cover the widths MFCC.elaborate uses (mfcc/core/mfcc.py:38-88)");
    if (cfg->mel_lo_bit < 30 || cfg->mel_lo_bit > 40) return fail(MFCC_B200_EUNSUPPORTED, "mel_lo_bit out of range");
    if (!cfg->window_curve || !cfg->tw_fft || !cfg->tw_dct || !cfg->mel_weight || !cfg->mel_close)
        return fail(MFCC_B200_EINVAL, "missing table");

    Handle *h = new Handle();
    h->cfg = *cfg;
    h->cfg.window_curve = nullptr; h->cfg.tw_fft = nullptr; h->cfg.tw_dct = nullptr;
    h->cfg.mel_weight = nullptr; h->cfg.mel_close = nullptr;
    DevTables t;
    int rc = build_tables(*cfg, t);
    if (rc) { delete h; return rc; }
    mfccf::DevTablesF *tf = nullptr;
    if (cfg->mode == MFCC_B200_MODE_FP32) {
        tf = new mfccf::DevTablesF();
        rc = build_tables_f(*cfg, *tf);
        if (rc) { delete tf; delete h; return rc; }
    }
    h->cfg.f_window = nullptr; h->cfg.f_mel = nullptr; h->cfg.f_dct = nullptr;
    cudaError_t e = cudaGetDevice(&h->device);
    if (e == cudaSuccess) e = cudaDeviceGetAttribute(&h->sm_count, cudaDevAttrMultiProcessorCount, h->device);
    if (e == cudaSuccess) e = cudaMalloc(&h->d_tab, sizeof(DevTables));
    if (e == cudaSuccess) e = cudaMemcpy(h->d_tab, &t, sizeof(DevTables), cudaMemcpyHostToDevice);
    if (e == cudaSuccess && tf) e = cudaMalloc(&h->d_ftab, sizeof(mfccf::DevTablesF));
    if (e == cudaSuccess && tf) e = cudaMemcpy(h->d_ftab, tf, sizeof(mfccf::DevTablesF), cudaMemcpyHostToDevice);
    delete tf;
    if (e != cudaSuccess) {
        if (h->d_tab) cudaFree(h->d_tab);
        if (h->d_ftab) cudaFree(h->d_ftab);
        delete h;
        return fail(MFCC_B200_ECUDA, std::string("mfcc_b200_create: ") + cudaGetErrorString(e));
    }
    rc = set_smem(mfcc_int_fused<32, false, ShapeBig>, sizeof(Smem<ShapeBig>));
    if (!rc) rc = set_smem(mfcc_int_fused<16, false, ShapeBig>, sizeof(Smem<ShapeBig>));
    if (!rc) rc = set_smem(mfcc_int_fused<32, false, ShapeSmall>, sizeof(Smem<ShapeSmall>));
    if (!rc) rc = set_smem(mfcc_int_fused<16, false, ShapeSmall>, sizeof(Smem<ShapeSmall>));
    if (!rc) rc = set_smem(mfcc_int_fused<32, true, ShapeSmall>, sizeof(Smem<ShapeSmall>));
    if (!rc) rc = set_smem(mfcc_int_fused<16, true, ShapeSmall>, sizeof(Smem<ShapeSmall>));
    if (!rc) rc = set_smem(mfcc_int_step<32>, sizeof(StepSmem));
    if (!rc) rc = set_smem(mfcc_int_step<16>, sizeof(StepSmem));
    if (!rc) rc = set_smem(k_warp_stage<MFCC_B200_STAGE_FFT, 32>, sizeof(StageSmem));
    if (!rc) rc = set_smem(k_warp_stage<MFCC_B200_STAGE_MEL, 32>, sizeof(StageSmem));
    if (!rc) rc = set_smem(k_warp_stage<MFCC_B200_STAGE_MEL, 16>, sizeof(StageSmem));
    if (!rc) rc = set_smem(k_warp_stage<MFCC_B200_STAGE_DCT, 32>, sizeof(StageSmem));
    if (!rc) rc = set_smem(k_warp_stage<MFCC_B200_STAGE_DCT, 16>, sizeof(StageSmem));
    if (!rc) rc = set_smem(mfcc_f32_fused<32>, sizeof(SmemF));
    if (!rc) rc = set_smem(mfcc_f32_fused<16>, sizeof(SmemF));
    if (!rc) rc = set_smem(mfcc_f32_step<32>, sizeof(StepSmemF));
    if (!rc) rc = set_smem(mfcc_f32_step<16>, sizeof(StepSmemF));
    if (rc) { cudaFree(h->d_tab); if (h->d_ftab) cudaFree(h->d_ftab); delete h; return rc; }
    *handle = h;
    return MFCC_B200_OK;
}

void mfcc_b200_destroy(void *handle)
{
    Handle *h = static_cast<Handle *>(handle);
    if (!h) return;
    host_pipe_free(h->hp, true);
    if (h->s_step) cudaStreamDestroy(h->s_step);
    if (h->s_state) cudaFree(h->s_state);
    if (h->s_hop) cudaFree(h->s_hop);
    if (h->s_out) cudaFree(h->s_out);
    if (h->d_tab) cudaFree(h->d_tab);
    if (h->d_ftab) cudaFree(h->d_ftab);
    delete h;
}

int mfcc_b200_run(void *handle, const int16_t *d_x, int64_t clip_stride, int64_t n_samples, int32_t n_clips,
                  int64_t n_frames, void *d_out, int64_t out_clip_stride, const mfcc_b200_taps *taps, void *stream)
{
    if (int rc = check_handle(handle)) return rc;
    Handle *h = static_cast<Handle *>(handle);
    if (n_clips < 0 || n_frames < 0 || n_samples < 0) return fail(MFCC_B200_EINVAL, "negative size");
    if (n_clips == 0 || n_frames == 0) return MFCC_B200_OK;
    if (!d_x || !d_out) return fail(MFCC_B200_EINVAL, "null buffer");
    if (n_clips > 1 && (clip_stride < n_samples || out_clip_stride < n_frames * h->cfg.nceptrums))
        return fail(MFCC_B200_EINVAL, "row pitch smaller than the row");
    if (taps && n_clips != 1) return fail(MFCC_B200_EINVAL, "taps need n_clips == 1");
    if (taps && h->cfg.mode == MFCC_B200_MODE_FP32) return fail(MFCC_B200_EUNSUPPORTED, "taps are an integer-mode feature");
    if (HOP * (n_frames - 1) + NFFT > (1LL << 40)) return fail(MFCC_B200_EINVAL, "clip too long");

    RunParams p;
    memset(&p, 0, sizeof(p));
    p.x = d_x; p.clip_stride = clip_stride; p.n_samples = n_samples; p.n_frames = n_frames;
    p.out = d_out; p.out_clip_stride = out_clip_stride;
    p.tab = h->d_tab; p.ftab = h->d_ftab; p.nceptrums = h->cfg.nceptrums;
    const bool f32 = h->cfg.mode == MFCC_B200_MODE_FP32;
    // integer mode: the large shape once the batch fills the machine four times over with its tiles
    // (and its partly empty last tile per clip does not cost more than the shape gains)
    const long long tpc_big = (n_frames + ShapeBig::FT - 1) / ShapeBig::FT;
    const long long tpc_small = (n_frames + ShapeSmall::FT - 1) / ShapeSmall::FT;
    const long long big_tiles = tpc_big * n_clips;
    const bool big = !f32 && !taps && big_tiles >= 4LL * h->sm_count * ShapeBig::CTAS &&
                     tpc_big * ShapeBig::FT * 100 <= tpc_small * ShapeSmall::FT * 105;
    const int ft = f32 ? FT_F : big ? ShapeBig::FT : ShapeSmall::FT;
    p.tiles_per_clip = (int)((n_frames + ft - 1) / ft);
    p.n_tiles = (long long)p.tiles_per_clip * n_clips;
    p.aligned = ((reinterpret_cast<uintptr_t>(d_x) & 15) == 0) && (n_clips == 1 || (clip_stride % 8) == 0);
    if (taps) p.taps = *taps;
    long long grid = (long long)h->sm_count * (f32 ? CTAS_PER_SM_F : big ? ShapeBig::CTAS : ShapeSmall::CTAS);
    if (grid > p.n_tiles) grid = p.n_tiles;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (f32) {
        if (h->cfg.nfilters == 32) mfcc_f32_fused<32><<<(unsigned)grid, THREADS_F, sizeof(SmemF), s>>>(p);
        else                       mfcc_f32_fused<16><<<(unsigned)grid, THREADS_F, sizeof(SmemF), s>>>(p);
    } else {
        const bool nf32 = h->cfg.nfilters == 32;
        using SB = ShapeBig;
        using SS = ShapeSmall;
        if (taps && nf32)  mfcc_int_fused<32, true, SS><<<(unsigned)grid, SS::THREADS, sizeof(Smem<SS>), s>>>(p);
        else if (taps)     mfcc_int_fused<16, true, SS><<<(unsigned)grid, SS::THREADS, sizeof(Smem<SS>), s>>>(p);
        else if (big && nf32) mfcc_int_fused<32, false, SB><<<(unsigned)grid, SB::THREADS, sizeof(Smem<SB>), s>>>(p);
        else if (big)      mfcc_int_fused<16, false, SB><<<(unsigned)grid, SB::THREADS, sizeof(Smem<SB>), s>>>(p);
        else if (nf32)     mfcc_int_fused<32, false, SS><<<(unsigned)grid, SS::THREADS, sizeof(Smem<SS>), s>>>(p);
        else               mfcc_int_fused<16, false, SS><<<(unsigned)grid, SS::THREADS, sizeof(Smem<SS>), s>>>(p);
    }
    g_launches.fetch_add(1);
    CUDA_TRY(cudaGetLastError());
    return MFCC_B200_OK;
}

int mfcc_b200_host_config(void *handle, int64_t chunk_bytes, int32_t slots)
{
    if (int rc = check_handle(handle)) return rc;
    Handle *h = static_cast<Handle *>(handle);
    if (chunk_bytes < 0 || slots < 0 || slots > HOST_SLOTS_MAX || slots == 1)
        return fail(MFCC_B200_EINVAL, "chunk_bytes >= 0 and slots in 2..8 (0 keeps the current value)");
    if (chunk_bytes) h->hp.chunk_bytes = chunk_bytes;
    if (slots) h->hp.slots = slots;
    return MFCC_B200_OK;
}

int mfcc_b200_debug_fail_chunk(void *handle, int32_t chunk)
{
    if (int rc = check_handle(handle)) return rc;
    static_cast<Handle *>(handle)->hp.fail_at_chunk = chunk;
    return MFCC_B200_OK;
}

int mfcc_b200_host_alloc(void **p, int64_t bytes, int32_t write_combined)
{
    if (!p || bytes <= 0) return fail(MFCC_B200_EINVAL, "bad argument");
    *p = nullptr;
    CUDA_TRY(cudaHostAlloc(p, (size_t)bytes, write_combined ? cudaHostAllocWriteCombined : cudaHostAllocDefault));
    return MFCC_B200_OK;
}

void mfcc_b200_host_free(void *p)
{
    if (p) cudaFreeHost(p);
}

int mfcc_b200_run_host(void *handle, const int16_t *h_x, int64_t clip_stride, int64_t n_samples, int32_t n_clips,
                       int64_t n_frames, void *h_out, int64_t out_clip_stride)
{
    if (int rc = check_handle(handle)) return rc;
    Handle *h = static_cast<Handle *>(handle);
    HostPipe &hp = h->hp;
    if (n_clips < 0 || n_frames < 0 || n_samples < 0) return fail(MFCC_B200_EINVAL, "negative size");
    if (n_clips == 0 || n_frames == 0) return MFCC_B200_OK;
    if (!h_x || !h_out) return fail(MFCC_B200_EINVAL, "null buffer");
    if (n_clips > 1 && (clip_stride < n_samples || out_clip_stride < n_frames * h->cfg.nceptrums))
        return fail(MFCC_B200_EINVAL, "row pitch smaller than the row");
    const int nc = h->cfg.nceptrums;
    const long long in_row = n_samples ? ((n_samples + 7) & ~7LL) : 8;   // device row pitch: 16-byte multiple (TMA)
    const long long out_row = n_frames * nc;
    long long per = hp.chunk_bytes / (in_row * 2);                       // clips per chunk
    if (per < 1) per = 1;
    if (per > n_clips) per = n_clips;
    const long long n_chunks = (n_clips + per - 1) / per;
    const int slots = (int)(n_chunks < hp.slots ? n_chunks : hp.slots);
    const size_t esz = h->cfg.mode == MFCC_B200_MODE_FP32 ? sizeof(float) : sizeof(int16_t);   // output element
    const size_t in_need = (size_t)per * in_row * 2, out_need = (size_t)per * out_row * esz;
    if (in_need > hp.in_cap || out_need > hp.out_cap || slots > hp.slots_alloc)
        if (int rc = host_pipe_alloc(hp, in_need > hp.in_cap ? in_need : hp.in_cap,
                                     out_need > hp.out_cap ? out_need : hp.out_cap, slots > hp.slots_alloc ? slots : hp.slots_alloc))
            return rc;
    // rows packed at the device pitch on the host too: one plain copy per chunk, else a pitched one
    const bool in_1d = n_clips == 1 || clip_stride == in_row;
    const bool out_1d = n_clips == 1 || out_clip_stride == out_row;
    unsigned char *out8 = static_cast<unsigned char *>(h_out);
    cudaError_t e = cudaSuccess;
    const char *what = "";
    int rc = MFCC_B200_OK;
    long long done = 0;
    for (long long c = 0; c < n_chunks && e == cudaSuccess && rc == MFCC_B200_OK; ++c) {
        const int slot = (int)(c % slots);
        const long long n = (n_clips - done) < per ? (n_clips - done) : per;
        // host -> device, once the kernel that last read d_in[slot] is done
        if (c >= slots) e = cudaStreamWaitEvent(hp.s_in, hp.ev_run[slot], 0), what = "cudaStreamWaitEvent";
        if (e == cudaSuccess && n_samples) {
            what = "host->device copy";
            if (in_1d)
                e = cudaMemcpyAsync(hp.d_in[slot], h_x + done * clip_stride, (size_t)((n - 1) * in_row + n_samples) * 2,
                                    cudaMemcpyHostToDevice, hp.s_in);
            else
                e = cudaMemcpy2DAsync(hp.d_in[slot], (size_t)in_row * 2, h_x + done * clip_stride, (size_t)clip_stride * 2,
                                      (size_t)n_samples * 2, (size_t)n, cudaMemcpyHostToDevice, hp.s_in);
        }
        if (e == cudaSuccess && hp.fail_at_chunk == (int)c) e = cudaErrorUnknown, what = "injected failure (mfcc_b200_debug_fail_chunk)";
        if (e == cudaSuccess) e = cudaEventRecord(hp.ev_in[slot], hp.s_in), what = "cudaEventRecord";
        // kernel, once the chunk is in and the copy that last read d_out[slot] is done
        if (e == cudaSuccess) e = cudaStreamWaitEvent(hp.s_run, hp.ev_in[slot], 0), what = "cudaStreamWaitEvent";
        if (e == cudaSuccess && c >= slots) e = cudaStreamWaitEvent(hp.s_run, hp.ev_out[slot], 0);
        if (e != cudaSuccess) break;
        rc = mfcc_b200_run(h, hp.d_in[slot], in_row, n_samples, (int32_t)n, n_frames, hp.d_out[slot], out_row, nullptr, hp.s_run);
        if (rc) break;
        e = cudaEventRecord(hp.ev_run[slot], hp.s_run), what = "cudaEventRecord";
        // device -> host
        if (e == cudaSuccess) e = cudaStreamWaitEvent(hp.s_out, hp.ev_run[slot], 0), what = "cudaStreamWaitEvent";
        if (e == cudaSuccess) {
            what = "device->host copy";
            unsigned char *dst = out8 + (size_t)done * out_clip_stride * esz;
            if (out_1d)
                e = cudaMemcpyAsync(dst, hp.d_out[slot], (size_t)n * out_row * esz, cudaMemcpyDeviceToHost, hp.s_out);
            else
                e = cudaMemcpy2DAsync(dst, (size_t)out_clip_stride * esz, hp.d_out[slot], (size_t)out_row * esz,
                                      (size_t)out_row * esz, (size_t)n, cudaMemcpyDeviceToHost, hp.s_out);
        }
        if (e == cudaSuccess) e = cudaEventRecord(hp.ev_out[slot], hp.s_out), what = "cudaEventRecord";
        done += n;
    }
    // Whatever happened, nothing of this call may still be in flight when it returns: the caller
    // owns h_x and h_out and may free them at once.
    const cudaError_t e1 = cudaStreamSynchronize(hp.s_in), e2 = cudaStreamSynchronize(hp.s_run),
                      e3 = cudaStreamSynchronize(hp.s_out);
    if (rc) return rc;                                                   // message set by mfcc_b200_run
    if (e != cudaSuccess) return fail(MFCC_B200_ECUDA, std::string("mfcc_b200_run_host: ") + what + ": " + cudaGetErrorString(e));
    for (cudaError_t es : {e1, e2, e3})
        if (es != cudaSuccess) return fail(MFCC_B200_ECUDA, std::string("mfcc_b200_run_host: cudaStreamSynchronize: ") + cudaGetErrorString(es));
    return MFCC_B200_OK;
}

int mfcc_b200_stage(void *handle, int32_t stage, const void *d_in, void *d_out, int64_t n, int64_t n_frames,
                    int32_t prev, void *stream)
{
    if (int rc = check_handle(handle)) return rc;
    Handle *h = static_cast<Handle *>(handle);
    if (stage < 0 || stage >= MFCC_B200_STAGE_COUNT) return fail(MFCC_B200_EINVAL, "unknown stage");
    if (h->cfg.mode != MFCC_B200_MODE_INT) return fail(MFCC_B200_EUNSUPPORTED, "stage operators are an integer-mode feature");
    if (n < 0 || n_frames < 0) return fail(MFCC_B200_EINVAL, "negative size");
    const long long rows = (stage == MFCC_B200_STAGE_FRAME) ? n_frames : n;
    if (rows == 0) return MFCC_B200_OK;
    if (!d_in || !d_out) return fail(MFCC_B200_EINVAL, "null buffer");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    const int nf = h->cfg.nfilters;
    const int B = 256;
    const uintptr_t ai = reinterpret_cast<uintptr_t>(d_in), ao = reinterpret_cast<uintptr_t>(d_out);
    const bool v16 = ((ai | ao) & 15) == 0;                     // both tensors take 128-bit accesses
    switch (stage) {
    case MFCC_B200_STAGE_PREEMPH:
        if (v16 && n >= 8) k_preemph_v<<<grid_1d(n / 8, B, h->sm_count), B, 0, s>>>((const int16_t *)d_in, (int16_t *)d_out, n, prev);
        else k_preemph<<<grid_1d(n, B, h->sm_count), B, 0, s>>>((const int16_t *)d_in, (int16_t *)d_out, n, prev);
        break;
    case MFCC_B200_STAGE_FRAME:
        if ((ai & 3) == 0 && (ao & 15) == 0)
            k_frame_v<<<grid_1d(n_frames * (NFFT / 8), B, h->sm_count), B, 0, s>>>((const int16_t *)d_in, (int16_t *)d_out, n, n_frames);
        else k_frame<<<grid_1d(n_frames * NFFT, B, h->sm_count), B, 0, s>>>((const int16_t *)d_in, (int16_t *)d_out, n, n_frames);
        break;
    case MFCC_B200_STAGE_WINDOW:
        if (v16) k_window_v<<<grid_1d(n * (NFFT / 8), B, h->sm_count), B, 0, s>>>((const int16_t *)d_in, (int16_t *)d_out, n, h->d_tab);
        else k_window<<<grid_1d(n * NFFT, B, h->sm_count), B, 0, s>>>((const int16_t *)d_in, (int16_t *)d_out, n, h->d_tab);
        break;
    case MFCC_B200_STAGE_POWER:
        if (v16) k_power_v<<<grid_1d(n * (NBINS / 4), B, h->sm_count), B, 0, s>>>((const int16_t *)d_in, (uint32_t *)d_out, n);
        else k_power<<<grid_1d(n * NBINS, B, h->sm_count), B, 0, s>>>((const int16_t *)d_in, (uint32_t *)d_out, n);
        break;
    case MFCC_B200_STAGE_LOG2:
        if (v16 && n * nf >= 8) k_log2_v<<<grid_1d(n * nf / 8, B, h->sm_count), B, 0, s>>>((const uint16_t *)d_in, (uint16_t *)d_out, n * nf, h->d_tab);
        else k_log2<<<grid_1d(n * nf, B, h->sm_count), B, 0, s>>>((const uint16_t *)d_in, (uint16_t *)d_out, n * nf, h->d_tab);
        break;
    default: {
        const int grid = grid_1d(n, WARPS, h->sm_count);
        if (stage == MFCC_B200_STAGE_FFT)
            k_warp_stage<MFCC_B200_STAGE_FFT, 32><<<grid, THREADS, sizeof(StageSmem), s>>>(d_in, d_out, n, h->d_tab);
        else if (stage == MFCC_B200_STAGE_MEL && nf == 32)
            k_warp_stage<MFCC_B200_STAGE_MEL, 32><<<grid, THREADS, sizeof(StageSmem), s>>>(d_in, d_out, n, h->d_tab);
        else if (stage == MFCC_B200_STAGE_MEL)
            k_warp_stage<MFCC_B200_STAGE_MEL, 16><<<grid, THREADS, sizeof(StageSmem), s>>>(d_in, d_out, n, h->d_tab);
        else if (nf == 32)
            k_warp_stage<MFCC_B200_STAGE_DCT, 32><<<grid, THREADS, sizeof(StageSmem), s>>>(d_in, d_out, n, h->d_tab);
        else
            k_warp_stage<MFCC_B200_STAGE_DCT, 16><<<grid, THREADS, sizeof(StageSmem), s>>>(d_in, d_out, n, h->d_tab);
    }
    }
    g_launches.fetch_add(1);
    CUDA_TRY(cudaGetLastError());
    return MFCC_B200_OK;
}

int mfcc_b200_state_bytes(void *handle, int32_t n_streams, int64_t *bytes)
{
    if (int rc = check_handle(handle)) return rc;
    if (!bytes || n_streams < 0) return fail(MFCC_B200_EINVAL, "bad argument");
    *bytes = (int64_t)n_streams * STATE_N * (int64_t)sizeof(int16_t);
    return MFCC_B200_OK;
}

int mfcc_b200_step(void *handle, void *d_state, const int16_t *d_hop, int32_t n_streams, void *d_out, void *stream)
{
    if (int rc = check_handle(handle)) return rc;
    Handle *h = static_cast<Handle *>(handle);
    if (n_streams < 0) return fail(MFCC_B200_EINVAL, "negative size");
    if (n_streams == 0) return MFCC_B200_OK;
    if (!d_state || !d_hop || !d_out) return fail(MFCC_B200_EINVAL, "null buffer");
    const bool f32s = h->cfg.mode == MFCC_B200_MODE_FP32;      // the fp32 kernel takes two streams per warp
    int grid = f32s ? (n_streams + 2 * WARPS - 1) / (2 * WARPS) : (n_streams + WARPS - 1) / WARPS;
    const int cap = h->sm_count * 8;
    if (grid > cap) grid = cap;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (h->cfg.mode == MFCC_B200_MODE_FP32) {
        if (h->cfg.nfilters == 32)
            mfcc_f32_step<32><<<grid, THREADS, sizeof(StepSmemF), s>>>((int16_t *)d_state, d_hop, n_streams, (float *)d_out,
                                                                        h->cfg.nceptrums, h->d_ftab);
        else
            mfcc_f32_step<16><<<grid, THREADS, sizeof(StepSmemF), s>>>((int16_t *)d_state, d_hop, n_streams, (float *)d_out,
                                                                        h->cfg.nceptrums, h->d_ftab);
    } else if (h->cfg.nfilters == 32)
        mfcc_int_step<32><<<grid, THREADS, sizeof(StepSmem), s>>>((int16_t *)d_state, d_hop, n_streams, (int16_t *)d_out,
                                                                   h->cfg.nceptrums, h->d_tab);
    else
        mfcc_int_step<16><<<grid, THREADS, sizeof(StepSmem), s>>>((int16_t *)d_state, d_hop, n_streams, (int16_t *)d_out,
                                                                   h->cfg.nceptrums, h->d_tab);
    g_launches.fetch_add(1);
    CUDA_TRY(cudaGetLastError());
    return MFCC_B200_OK;
}

int mfcc_b200_step_host(void *handle, int32_t n_streams, const int16_t *h_hop, void *h_out, int32_t reset)
{
    if (int rc = check_handle(handle)) return rc;
    Handle *h = static_cast<Handle *>(handle);
    if (n_streams < 1) return fail(MFCC_B200_EINVAL, "n_streams must be >= 1");
    const size_t esz = h->cfg.mode == MFCC_B200_MODE_FP32 ? sizeof(float) : sizeof(int16_t);
    const size_t state_bytes = (size_t)n_streams * STATE_N * sizeof(int16_t);
    if (!h->s_step) CUDA_TRY(cudaStreamCreateWithFlags(&h->s_step, cudaStreamNonBlocking));
    if (h->s_streams != n_streams) {
        if (h->s_state) cudaFree(h->s_state);
        if (h->s_hop) cudaFree(h->s_hop);
        if (h->s_out) cudaFree(h->s_out);
        h->s_state = h->s_out = nullptr; h->s_hop = nullptr; h->s_streams = 0;
        CUDA_TRY(cudaMalloc(&h->s_state, state_bytes));
        CUDA_TRY(cudaMalloc(&h->s_hop, (size_t)n_streams * HOP * sizeof(int16_t)));
        CUDA_TRY(cudaMalloc(&h->s_out, (size_t)n_streams * h->cfg.nceptrums * esz));
        h->s_streams = n_streams;
        reset = 1;
    }
    if (reset) CUDA_TRY(cudaMemsetAsync(h->s_state, 0, state_bytes, h->s_step));      // the reset of every stream
    if (!h_hop) {                                                                    // reset only
        CUDA_TRY(cudaStreamSynchronize(h->s_step));
        return MFCC_B200_OK;
    }
    if (!h_out) return fail(MFCC_B200_EINVAL, "null buffer");
    CUDA_TRY(cudaMemcpyAsync(h->s_hop, h_hop, (size_t)n_streams * HOP * sizeof(int16_t), cudaMemcpyHostToDevice, h->s_step));
    if (int rc = mfcc_b200_step(h, h->s_state, h->s_hop, n_streams, h->s_out, h->s_step)) return rc;
    CUDA_TRY(cudaMemcpyAsync(h_out, h->s_out, (size_t)n_streams * h->cfg.nceptrums * esz, cudaMemcpyDeviceToHost, h->s_step));
    CUDA_TRY(cudaStreamSynchronize(h->s_step));
    return MFCC_B200_OK;
}

}  // extern "C"
