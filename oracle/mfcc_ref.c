/*
 * TEST INFRASTRUCTURE -- CPU oracle (plain C) for the MFCC hot path.
 *
 * A restatement of the integer arithmetic of the reference's nMigen RTL, one
 * function per stage, each citing the file:line it follows under
 * /root/reference.  It is the checker for the CUDA path and the timed CPU
 * baseline of bench.py ("kind": "port").  Only tests/, __graft_entry__.smoke()
 * and bench.py may load it; the product library never links or calls it.
 *
 * Tables (window curve, twiddles, mel weights) are passed in by the caller and
 * are built by oracle/ref_np.py the way the reference builds them at elaboration
 * time; this file holds only run-time arithmetic.
 *
 * Pinning: checked bit for bit, stage by stage, against vectors made by executing the
 * reference's own RTL source (tests/golden/rtl_*.npz, tests/test_rtl_golden.py);
 * see the header of oracle/ref_np.py and DESIGN.md ("Oracle and pinning").
 *
 * Build: make -C oracle      (gcc -O2 -fopenmp -shared -fPIC)
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef struct {
    int32_t nfft;            /* MFCC(nfft)                       mfcc/core/mfcc.py:20 */
    int32_t hop;             /* stepsize = nfft//3               mfcc/core/mfcc.py:43 */
    int32_t nfilters;        /* MFCC(nfilters)                   mfcc/core/mfcc.py:21 */
    int32_t nceptrums;       /* MFCC(nceptrums)                  mfcc/core/mfcc.py:21 */
    int32_t window_shift;    /* precision+1 = 9                  mfcc/core/window.py:20,84 */
    int32_t pow_shift;       /* 32 - width_output = 2            mfcc/core/pow2.py:64 */
    int32_t mel_wbits;       /* width_mul = 30                   mfcc/core/filterbank.py:49-52 */
    int32_t mel_lo_bit;      /* maxvalrange - (gain+width_output) = 31   filterbank.py:71,122 */
    int32_t mel_out_bits;    /* width_output = 16                mfcc/core/mfcc.py:70 */
    int32_t log_precision;   /* 11                               mfcc/core/log.py:114 */
    int32_t log_out_bits;    /* 15                               mfcc/core/mfcc.py:82 */
    int32_t reserved;
    const int32_t  *window_curve;  /* [nfft]           multipliers, window.py:59-115 */
    const int32_t  *tw_fft;        /* [nfft/2][2]      decoded ROM, fft.py:43-59 */
    const int32_t  *tw_dct;        /* [2*nfilters][2]  decoded ROM of the 4*nfilters FFT */
    const uint64_t *mel_step;      /* [nfilters+1]     filterbank.py:22-34 */
} mfcc_ref_cfg;

typedef struct {              /* optional per-stage outputs of clip 0, any may be NULL */
    int16_t  *window;         /* [K][nfft]        */
    int16_t  *fft;            /* [K][nfft/2][2]   (re, im) */
    uint32_t *power;          /* [K][nfft/2]      */
    uint16_t *mel;            /* [K][nfilters]    */
    uint16_t *log2;           /* [K][nfilters]    */
    int16_t  *dct;            /* [K][nfilters]    */
} mfcc_ref_taps;

static inline int32_t wrap16(int32_t v) { return (int32_t)(int16_t)(uint16_t)(uint32_t)v; }

/* Preemph: source = sink + (odata >> 5) - odata, 16-bit.  mfcc/core/preemph.py:18-28 */
static inline int32_t preemph1(int32_t x, int32_t prev) { return wrap16(x + (prev >> 5) - prev); }

/* Butterfly(width=16, bias_width=14, scale_bit=1).  mfcc/misc/fft.py:93-96, :140-192.
 * The three-multiplier form of the RTL equals re*re - im*im etc. exactly in 33 bits;
 * the unsigned [14:] slice only matters modulo 2^17 because bits 1..16 of the sum
 * are kept, so signed floor shifts give the same bits. */
static inline void butterfly(int32_t *x0r, int32_t *x0i, int32_t *x1r, int32_t *x1i,
                             int32_t twr, int32_t twi)
{
    const int64_t bias = (1 << 13) - 1;
    int64_t pr = (int64_t)*x1r * twr - (int64_t)*x1i * twi + bias;
    int64_t pi = (int64_t)*x1i * twr + (int64_t)*x1r * twi + bias;
    int64_t tr = pr >> 14, ti = pi >> 14;
    int64_t ar = *x0r, ai = *x0i;
    *x0r = wrap16((int32_t)((ar + tr) >> 1));
    *x0i = wrap16((int32_t)((ai + ti) >> 1));
    *x1r = wrap16((int32_t)((ar - tr) >> 1));
    *x1i = wrap16((int32_t)((ai - ti) >> 1));
}

static unsigned bitrev(unsigned n, int bits)
{
    unsigned r = 0;
    for (int i = 0; i < bits; i++) { r = (r << 1) | (n & 1); n >>= 1; }
    return r;
}

/* FFT(size): bit-reversed load (fft.py:413-418, :440-450), then the decimation-in-time
 * schedule the Scheduler realises (fft.py:254-258, :310-316, :328-342). In place on
 * vr/vi, which must already hold the bit-reversed input. tw = decoded ROM [size/2][2]. */
static void fft_stages(int32_t *vr, int32_t *vi, int size, int nb, const int32_t *tw)
{
    for (int s = 0; s < nb; s++) {
        int half = 1 << s;
        int tstep = (size / 2) >> s;
        for (int g = 0; g < size; g += 2 * half)
            for (int k = 0; k < half; k++) {
                int p = g + k, q = p + half;
                butterfly(&vr[p], &vi[p], &vr[q], &vi[q], tw[2 * k * tstep], tw[2 * k * tstep + 1]);
            }
    }
}

/* Log2Fix -> Log2FixCalc.  mfcc/core/log.py:123-126 (input), :57-62 (SHIFT-RIGHT),
 * :65 + :80-102 (precision-1 squarings), :133 (output slice). */
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

typedef unsigned __int128 u128;

static void one_clip(const mfcc_ref_cfg *c, const int16_t *x, int64_t n_samples, int64_t n_frames,
                     int16_t *out, const mfcc_ref_taps *taps, int32_t *scratch)
{
    const int nfft = c->nfft, nbins = nfft / 2, nf = c->nfilters, dsz = 4 * nf;
    int nb = 0, nbd = 0;
    while ((1 << nb) < nfft) nb++;
    while ((1 << nbd) < dsz) nbd++;
    int32_t *vr = scratch, *vi = scratch + nfft;
    int32_t *dr = vi + nfft, *di = dr + dsz;
    uint32_t *mel = (uint32_t *)(di + dsz);
    uint32_t *lg = mel + nf;
    const uint64_t wtop = ((uint64_t)1 << c->mel_wbits) - 1;
    const uint64_t accmask = (c->mel_wbits * 2 >= 64) ? ~(uint64_t)0 : (((uint64_t)1 << (2 * c->mel_wbits)) - 1);

    /* FilterBank registers persist from frame to frame (filterbank.py:90, :96, :119-120) */
    uint64_t i_acc = 0; int filter_adr = 0; u128 rega = 0, regb = 0;

    for (int64_t k = 0; k < n_frames; k++) {
        int64_t base = (int64_t)c->hop * k;
        /* Preemph + Frame + WindowHamming, loaded bit-reversed for the FFT.
         * frame k = y[hop*k .. hop*k+nfft-1] (frame.py:117-148 with windowlen = nfft);
         * samples past the end of the clip are zeros, as the hosts feed them
         * (mfcc/core/mfcc.py:150-153, software/main.c:135-144). */
        for (int n = 0; n < nfft; n++) {
            int64_t i = base + n;
            int32_t cur = (i < n_samples) ? x[i] : 0;
            int32_t prv = (i - 1 >= 0 && i - 1 < n_samples) ? x[i - 1] : 0;
            int32_t y = preemph1(cur, prv);
            /* window.py:74-75, :84: top 16 bits of signed16 * unsigned9 */
            int32_t w = wrap16((y * c->window_curve[n]) >> c->window_shift);
            if (taps && taps->window) taps->window[k * nfft + n] = (int16_t)w;
            unsigned r = bitrev((unsigned)n, nb);
            vr[r] = w; vi[r] = 0;                       /* fft_stream.py:32-33 */
        }
        fft_stages(vr, vi, nfft, nb, c->tw_fft);

        /* FftStream emits bins 0..nfft/2-1 (fft_stream.py:24, :28); PowerSpectrum
         * (pow2.py:32, :64); FilterBank (filterbank.py:90-142). */
        int nout = 0;
        for (int b = 0; b < nbins; b++) {
            if (taps && taps->fft) {
                taps->fft[(k * nbins + b) * 2] = (int16_t)vr[b];
                taps->fft[(k * nbins + b) * 2 + 1] = (int16_t)vi[b];
            }
            uint32_t s = (uint32_t)(vr[b] * vr[b]) + (uint32_t)(vi[b] * vi[b]);
            uint64_t d = s >> c->pow_shift;
            if (taps && taps->power) taps->power[k * nbins + b] = (uint32_t)d;

            int last = (b == nbins - 1);
            uint64_t w = i_acc >> c->mel_wbits;
            int highest = (w == wtop);
            int adr = filter_adr;
            u128 cc = (u128)d * w;
            if (highest || last) { filter_adr = last ? 0 : adr + 1; i_acc = 0; }
            else                 { i_acc = (i_acc + c->mel_step[adr]) & accmask; }
            uint32_t o_data = (uint32_t)(regb >> c->mel_lo_bit) & ((1u << c->mel_out_bits) - 1);
            if (highest || last) {
                if (adr != 0 && nout < nf) mel[nout++] = o_data;
                regb = rega + ((u128)d << c->mel_wbits);
                rega = 0;
            } else {
                rega += cc;
                regb += ((u128)d << c->mel_wbits) - cc;
            }
        }
        for (; nout < nf; nout++) mel[nout] = 0;   /* cannot happen for realisable configs */

        /* Log2Fix on each mel value, then DCTStream fill (dct_stream.py:29-34) */
        memset(dr, 0, sizeof(int32_t) * 2 * dsz);
        for (int n = 0; n < nf; n++) {
            lg[n] = log2fix(mel[n], c->log_precision, c->log_out_bits);
            if (taps && taps->mel)  taps->mel[k * nf + n] = (uint16_t)mel[n];
            if (taps && taps->log2) taps->log2[k * nf + n] = (uint16_t)lg[n];
            dr[bitrev((unsigned)(2 * n + 1), nbd)] = (int32_t)lg[n];
            dr[bitrev((unsigned)(dsz - 1 - 2 * n), nbd)] = (int32_t)lg[n];
        }
        fft_stages(dr, di, dsz, nbd, c->tw_dct);
        /* real part of bins 0..nfilters-1 (dct_stream.py:36-37, :24, :28); Discard
         * keeps the first nceptrums (discard.py:27-55) */
        for (int n = 0; n < nf; n++) {
            if (taps && taps->dct) taps->dct[k * nf + n] = (int16_t)dr[n];
            if (n < c->nceptrums) out[k * c->nceptrums + n] = (int16_t)dr[n];
        }
    }
}

/* Whole chain MFCC.elaborate (mfcc/core/mfcc.py:38-104) on n_clips independent
 * streams, each starting from reset.  x: n_clips rows of n_samples int16, row pitch
 * clip_stride (samples).  out: n_clips rows of n_frames*nceptrums int16, row pitch
 * out_clip_stride (elements).  taps apply to clip 0.  Returns 0, or -1 on bad
 * arguments. */
int mfcc_ref_run(const mfcc_ref_cfg *cfg, const int16_t *x, int64_t clip_stride, int64_t n_samples,
                 int32_t n_clips, int64_t n_frames, int16_t *out, int64_t out_clip_stride,
                 const mfcc_ref_taps *taps, int32_t n_threads)
{
    if (!cfg || !x || !out || n_clips < 0 || n_frames < 0) return -1;
    size_t words = (size_t)2 * cfg->nfft + 8 * cfg->nfilters + 2 * cfg->nfilters + 16;
    int nthr = n_threads > 0 ? n_threads : 1;
#ifdef _OPENMP
    #pragma omp parallel num_threads(nthr)
#endif
    {
        int32_t *scratch = (int32_t *)malloc(words * sizeof(int32_t));
#ifdef _OPENMP
        #pragma omp for schedule(dynamic, 1)
#endif
        for (int32_t ci = 0; ci < n_clips; ci++)
            one_clip(cfg, x + (int64_t)ci * clip_stride, n_samples, n_frames,
                     out + (int64_t)ci * out_clip_stride, ci == 0 ? taps : NULL, scratch);
        free(scratch);
    }
    (void)nthr;
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

/* Single-stage entry points used by the stage parity tests. */
void mfcc_ref_fft(int32_t *vr, int32_t *vi, int32_t size, const int32_t *tw, int64_t n)
{
    int nb = 0; while ((1 << nb) < size) nb++;
    int32_t *tr = (int32_t *)malloc(sizeof(int32_t) * 2 * size), *ti = tr + size;
    for (int64_t f = 0; f < n; f++) {
        int32_t *r = vr + f * size, *i = vi + f * size;
        for (int p = 0; p < size; p++) { unsigned q = bitrev((unsigned)p, nb); tr[q] = r[p]; ti[q] = i[p]; }
        fft_stages(tr, ti, size, nb, tw);
        memcpy(r, tr, sizeof(int32_t) * size); memcpy(i, ti, sizeof(int32_t) * size);
    }
    free(tr);
}

void mfcc_ref_log2(const uint32_t *m, uint32_t *o, int64_t n, int32_t prec, int32_t out_bits)
{
    for (int64_t i = 0; i < n; i++) o[i] = log2fix(m[i], prec, out_bits);
}
