/*
 * TEST INFRASTRUCTURE -- the TIMED CPU baseline of bench.py ("cpu_baseline", "--impl reference").
 *
 * Same arithmetic and same entry point as the checker oracle/mfcc_ref.c (which stays the
 * plain, line-by-line restatement of the RTL that the parity tests use), written the way a
 * CPU implementation that is meant to be fast would be:
 *   - built with  gcc -O3 -march=native -fopenmp   (oracle/Makefile, target libmfcc_ref_fast.so)
 *   - pre-emphasis once per sample of the clip, not once per frame position
 *   - bit reversal and per-stage twiddles from tables built once per call
 *   - butterflies in 32-bit arithmetic on split re/im arrays with unit-stride inner loops that
 *     gcc vectorises (a 16-bit value times a 15-bit twiddle is below 2^30, the sum of two such
 *     products plus the bias fits 32 bits)
 *   - FilterBank accumulators in 64 bits (only bits 31..46 are read, filterbank.py:122: arithmetic
 *     modulo 2^64 gives the same bits as the RTL's 65-bit registers)
 * tests/test_oracle.py requires it to equal the checker bit for bit.  Only tests/ and bench.py
 * may load it; the product library never links or calls it.
 *
 * Cites: the same reference lines as oracle/mfcc_ref.c (mfcc/core/preemph.py:18-28,
 * window.py:74-84, misc/fft.py:93-96 + :152-191 + :413-418, pow2.py:32 + :64,
 * filterbank.py:90-142, log.py:57-102 + :123-133, dct_stream.py:29-37, misc/discard.py:27-55).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef struct {
    int32_t nfft, hop, nfilters, nceptrums, window_shift, pow_shift, mel_wbits, mel_lo_bit, mel_out_bits,
            log_precision, log_out_bits, reserved;
    const int32_t  *window_curve;
    const int32_t  *tw_fft;
    const int32_t  *tw_dct;
    const uint64_t *mel_step;
} mfcc_ref_cfg;

typedef struct { void *window, *fft, *power, *mel, *log2, *dct; } mfcc_ref_taps;   /* not supported here */

static inline int32_t wrap16(int32_t v) { return (int32_t)(int16_t)(uint16_t)(uint32_t)v; }

/* tables shared by all threads of one call */
typedef struct {
    int nfft, nb, dsz, nbd;
    uint16_t *rev;           /* [nfft] bit reversal */
    int32_t *twr, *twi;      /* per stage s, entries k = 0..2^s-1 at offset 2^s - 1: the twiddle of butterfly k */
    uint16_t *drev;          /* [dsz] */
    int32_t *dtwr, *dtwi;
    uint32_t *melw;          /* [nfft/2] weight of bin b (filterbank.py:90-115) */
    uint8_t *melclose;       /* [nfft/2] bin closes its segment (peak or last) */
} fast_tabs;

static unsigned bitrev(unsigned n, int bits)
{
    unsigned r = 0;
    for (int i = 0; i < bits; i++) { r = (r << 1) | (n & 1); n >>= 1; }
    return r;
}

static void stage_twiddles(int size, int nb, const int32_t *tw, int32_t *twr, int32_t *twi)
{
    for (int s = 0; s < nb; s++) {
        const int half = 1 << s, tstep = (size / 2) >> s;
        for (int k = 0; k < half; k++) { twr[half - 1 + k] = tw[2 * k * tstep]; twi[half - 1 + k] = tw[2 * k * tstep + 1]; }
    }
}

/* radix-2 DIT stages s0..nb-1 in place on split arrays holding the bit-reversed input */
static void fft_stages(int32_t *restrict vr, int32_t *restrict vi, int size, int nb, int s0,
                       const int32_t *restrict twr, const int32_t *restrict twi)
{
    for (int s = s0; s < nb; s++) {
        const int half = 1 << s;
        const int32_t *restrict wr = twr + half - 1, *restrict wi = twi + half - 1;
        for (int g = 0; g < size; g += 2 * half) {
            int32_t *restrict ar = vr + g, *restrict ai = vi + g, *restrict br = vr + g + half, *restrict bi = vi + g + half;
#pragma omp simd
            for (int k = 0; k < half; k++) {
                const int32_t tr = (br[k] * wr[k] - bi[k] * wi[k] + 8191) >> 14;
                const int32_t ti = (bi[k] * wr[k] + br[k] * wi[k] + 8191) >> 14;
                const int32_t xr = ar[k], xi = ai[k];
                ar[k] = (int32_t)(int16_t)((xr + tr) >> 1);
                ai[k] = (int32_t)(int16_t)((xi + ti) >> 1);
                br[k] = (int32_t)(int16_t)((xr - tr) >> 1);
                bi[k] = (int32_t)(int16_t)((xi - ti) >> 1);
            }
        }
    }
}

static uint32_t log2fix(uint32_t m, int prec, int out_bits)
{
    uint64_t x = (m == 0) ? ((uint64_t)1 << prec) : ((uint64_t)m << prec);
    uint32_t o = 0, b = 1u << (prec - 1);
    while (x >> (prec + 1)) { x >>= 1; o += 1u << prec; }
    uint64_t z = x;
    for (int cnt = prec - 1; cnt != 0; cnt--) {
        uint64_t c = z * z;
        if ((c >> (2 * prec + 1)) & 1) { z = c >> (prec + 1); o += b; }
        else                           { z = c >> prec; }
        b >>= 1;
    }
    return o & ((1u << out_bits) - 1);
}

static void one_clip(const mfcc_ref_cfg *c, const fast_tabs *t, const int16_t *x, int64_t n_samples, int64_t n_frames,
                     int16_t *out, int16_t *pre, int32_t *scratch)
{
    const int nfft = c->nfft, nbins = nfft / 2, nf = c->nfilters, dsz = t->dsz;
    int32_t *vr = scratch, *vi = scratch + nfft, *dr = vi + nfft, *di = dr + dsz;
    const int64_t need = n_frames ? (int64_t)c->hop * (n_frames - 1) + nfft : 0;
    /* Preemph over the stream, zeros past the end of the clip (software/main.c:135-144) */
    int32_t prev = 0;
    for (int64_t i = 0; i < need; i++) {
        const int32_t cur = i < n_samples ? x[i] : 0;
        pre[i] = (int16_t)(cur + (prev >> 5) - prev);
        prev = cur;
    }
    /* the FilterBank registers persist from frame to frame; after the last bin of a frame
     * rega = 0 and regb = d_last << wbits (filterbank.py:119-134), which only ever reaches
     * the suppressed output of segment 0 */
    uint64_t rega = 0, regb = 0;
    const uint32_t omask = (1u << c->mel_out_bits) - 1;
    for (int64_t k = 0; k < n_frames; k++) {
        const int16_t *fr = pre + (int64_t)c->hop * k;
        for (int n = 0; n < nfft; n++) {
            const int32_t w = wrap16(((int32_t)fr[n] * c->window_curve[n]) >> c->window_shift);
            vr[t->rev[n]] = w;
        }
        memset(vi, 0, sizeof(int32_t) * nfft);
        fft_stages(vr, vi, nfft, t->nb, 0, t->twr, t->twi);
        uint32_t mel[64];
        int nout = 0, adr = 0;
        for (int b = 0; b < nbins; b++) {
            const uint64_t d = ((uint32_t)(vr[b] * vr[b]) + (uint32_t)(vi[b] * vi[b])) >> c->pow_shift;
            if (t->melclose[b]) {
                if (adr != 0 && nout < nf) mel[nout++] = (uint32_t)(regb >> c->mel_lo_bit) & omask;
                regb = rega + (d << c->mel_wbits);
                rega = 0;
                adr++;
            } else {
                const uint64_t cc = d * t->melw[b];
                rega += cc;
                regb += (d << c->mel_wbits) - cc;
            }
        }
        for (; nout < nf; nout++) mel[nout] = 0;
        memset(dr, 0, sizeof(int32_t) * 2 * dsz);
        for (int n = 0; n < nf; n++) {
            const int32_t lg = (int32_t)log2fix(mel[n], c->log_precision, c->log_out_bits);
            dr[t->drev[2 * n + 1]] = lg;
            dr[t->drev[dsz - 1 - 2 * n]] = lg;
        }
        fft_stages(dr, di, dsz, t->nbd, 0, t->dtwr, t->dtwi);
        for (int n = 0; n < c->nceptrums; n++) out[k * c->nceptrums + n] = (int16_t)dr[n];
    }
}

int mfcc_ref_run(const mfcc_ref_cfg *cfg, const int16_t *x, int64_t clip_stride, int64_t n_samples,
                 int32_t n_clips, int64_t n_frames, int16_t *out, int64_t out_clip_stride,
                 const mfcc_ref_taps *taps, int32_t n_threads)
{
    if (!cfg || !x || !out || n_clips < 0 || n_frames < 0 || taps) return -1;
    if (cfg->nfilters > 64 || 2 * cfg->mel_wbits > 64 || cfg->mel_lo_bit + cfg->mel_out_bits > 64) return -1;
    fast_tabs t;
    t.nfft = cfg->nfft; t.dsz = 4 * cfg->nfilters;
    t.nb = 0; while ((1 << t.nb) < t.nfft) t.nb++;
    t.nbd = 0; while ((1 << t.nbd) < t.dsz) t.nbd++;
    const int nbins = t.nfft / 2;
    t.rev = (uint16_t *)malloc(sizeof(uint16_t) * (t.nfft + t.dsz)); t.drev = t.rev + t.nfft;
    t.twr = (int32_t *)malloc(sizeof(int32_t) * 2 * (t.nfft + t.dsz)); t.twi = t.twr + t.nfft;
    t.dtwr = t.twi + t.nfft; t.dtwi = t.dtwr + t.dsz;
    t.melw = (uint32_t *)malloc(sizeof(uint32_t) * nbins); t.melclose = (uint8_t *)malloc(nbins);
    for (int n = 0; n < t.nfft; n++) t.rev[n] = (uint16_t)bitrev((unsigned)n, t.nb);
    for (int n = 0; n < t.dsz; n++) t.drev[n] = (uint16_t)bitrev((unsigned)n, t.nbd);
    stage_twiddles(t.nfft, t.nb, cfg->tw_fft, t.twr, t.twi);
    stage_twiddles(t.dsz, t.nbd, cfg->tw_dct, t.dtwr, t.dtwi);
    {   /* the weight generator of filterbank.py:90-115 does not depend on the data */
        const uint64_t wtop = ((uint64_t)1 << cfg->mel_wbits) - 1;
        const uint64_t accmask = (cfg->mel_wbits * 2 >= 64) ? ~(uint64_t)0 : (((uint64_t)1 << (2 * cfg->mel_wbits)) - 1);
        uint64_t acc = 0; int adr = 0;
        for (int b = 0; b < nbins; b++) {
            const uint64_t w = acc >> cfg->mel_wbits;
            const int close = (w == wtop) || (b == nbins - 1);
            t.melw[b] = (uint32_t)w; t.melclose[b] = (uint8_t)close;
            if (close) { adr++; acc = 0; } else acc = (acc + cfg->mel_step[adr]) & accmask;
        }
    }
    const int64_t need = n_frames ? (int64_t)cfg->hop * (n_frames - 1) + cfg->nfft : 0;
    const size_t words = (size_t)2 * t.nfft + 2 * t.dsz + 16;
    int nthr = n_threads > 0 ? n_threads : 1;
#ifdef _OPENMP
    #pragma omp parallel num_threads(nthr)
#endif
    {
        int32_t *scratch = (int32_t *)malloc(words * sizeof(int32_t));
        int16_t *pre = (int16_t *)malloc((size_t)(need + 8) * sizeof(int16_t));
#ifdef _OPENMP
        #pragma omp for schedule(dynamic, 1)
#endif
        for (int32_t ci = 0; ci < n_clips; ci++)
            one_clip(cfg, &t, x + (int64_t)ci * clip_stride, n_samples, n_frames,
                     out + (int64_t)ci * out_clip_stride, pre, scratch);
        free(scratch); free(pre);
    }
    (void)nthr;
    free(t.rev); free(t.twr); free(t.melw); free(t.melclose);
    return 0;
}

int mfcc_ref_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
