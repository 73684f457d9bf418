/*
 * mfcc_b200.h -- C ABI of the B200 MFCC path (libmfcc_b200.so).
 *
 * The reference (lambdaconcept/mfcc) has no FFI layer: its hot path is an nMigen
 * hardware pipeline whose interface is a chain of stream endpoints
 * (/root/reference/mfcc/core/mfcc.py:19-30, :90-104) and whose host side talks to
 * the FPGA over USB (software/main.c:100-177).  The entry points below are what a
 * host binding for this path replaces those with; each one cites the reference
 * interface it stands in for.  INTEGRATION.md shows the binding stubs.
 *
 * Conventions (same as the reference's C tools, software/ft601.h:43-50,
 * software/cepstrum.h:20-24): plain pointers and sizes, caller owns every buffer,
 * 0 on success and a negative code on failure, never aborts.
 * Pointers named d_* are DEVICE pointers; h_* are HOST pointers.
 * `stream` is a cudaStream_t passed as void* (NULL = default stream); every device
 * entry point is asynchronous on it.  A handle is not thread-safe; handles are
 * independent.
 */
#ifndef MFCC_B200_H
#define MFCC_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MFCC_B200_OK            0
#define MFCC_B200_EINVAL      (-1)   /* bad argument */
#define MFCC_B200_EUNSUPPORTED (-2)  /* configuration the kernels do not cover */
#define MFCC_B200_ECUDA       (-3)   /* CUDA runtime error, see mfcc_b200_last_error() */

#define MFCC_B200_MODE_INT   0   /* bit-exact fixed point (the FPGA's arithmetic) */
#define MFCC_B200_MODE_FP32  1   /* float32, notebook/MFCC-INT.ipynb structure */

/* Stage identifiers, in pipeline order (mfcc/core/mfcc.py:90-104). */
enum mfcc_b200_stage_id {
    MFCC_B200_STAGE_PREEMPH = 0,  /* Preemph         mfcc/core/preemph.py:9-30      int16[n]        -> int16[n]          */
    MFCC_B200_STAGE_FRAME   = 1,  /* Frame           mfcc/core/frame.py:49-155      int16[n]        -> int16[K][nfft]    */
    MFCC_B200_STAGE_WINDOW  = 2,  /* WindowHamming   mfcc/core/window.py:10-126     int16[K][nfft]  -> int16[K][nfft]    */
    MFCC_B200_STAGE_FFT     = 3,  /* FftStream       mfcc/core/fft_stream.py:6-73   int16[K][nfft]  -> int16[K][nfft/2][2] (re,im) */
    MFCC_B200_STAGE_POWER   = 4,  /* PowerSpectrum   mfcc/core/pow2.py:43-70        int16[K][nfft/2][2] -> uint32[K][nfft/2] */
    MFCC_B200_STAGE_MEL     = 5,  /* FilterBank      mfcc/core/filterbank.py:37-144 uint32[K][nfft/2] -> uint16[K][nfilters] */
    MFCC_B200_STAGE_LOG2    = 6,  /* Log2Fix         mfcc/core/log.py:107-139       uint16[K][nfilters] -> uint16[K][nfilters] */
    MFCC_B200_STAGE_DCT     = 7,  /* DCTStream       mfcc/core/dct_stream.py:6-73   uint16[K][nfilters] -> int16[K][nfilters] */
    MFCC_B200_STAGE_COUNT   = 8   /* Discard (mfcc/misc/discard.py) is a slice; the host does it */
};

/*
 * Configuration = the arguments of MFCC(width, nfft, samplerate, nfilters, nceptrums)
 * (mfcc/core/mfcc.py:19-26), the constants MFCC.elaborate hands to its stages
 * (:38-88) and the ROM contents those stages compute when the design is elaborated
 * (window.py:22-43, fft.py:28-36, filterbank.py:15-34).  The tables are HOST arrays,
 * copied by mfcc_b200_create(); mfcc_b200/tables.py builds them.
 */
typedef struct mfcc_b200_cfg {
    int32_t struct_size;     /* sizeof(mfcc_b200_cfg), for ABI checking */
    int32_t mode;            /* MFCC_B200_MODE_* */
    int32_t nfft;            /* 512 (the only size the kernels cover) */
    int32_t hop;             /* nfft/3 = 170            mfcc/core/mfcc.py:43 */
    int32_t nfilters;        /* 16 or 32                mfcc/core/mfcc.py:21 */
    int32_t nceptrums;       /* 1..nfilters             mfcc/core/mfcc.py:21 */
    int32_t window_shift;    /* precision+1 = 9         mfcc/core/window.py:20, :84 */
    int32_t pow_shift;       /* 32-width_output = 2     mfcc/core/pow2.py:64 */
    int32_t mel_wbits;       /* 30                      mfcc/core/filterbank.py:49-52 */
    int32_t mel_lo_bit;      /* first kept bit of the mel accumulator: 31 (32 filters),
                                32 (16 filters)         mfcc/core/filterbank.py:71, :122 */
    int32_t mel_out_bits;    /* 16                      mfcc/core/mfcc.py:70 */
    int32_t log_precision;   /* 11                      mfcc/core/log.py:114 */
    int32_t log_out_bits;    /* 15                      mfcc/core/mfcc.py:82 */
    int32_t reserved[3];
    const int32_t  *window_curve;  /* [nfft]            multiplier per sample position */
    const int32_t  *tw_fft;        /* [nfft/2][2]       decoded twiddle ROM (re, im) of the nfft-point FFT */
    const int32_t  *tw_dct;        /* [2*nfilters][2]   decoded twiddle ROM of the 4*nfilters-point FFT */
    const uint32_t *mel_weight;    /* [nfft/2]          weight applied to each bin */
    const uint8_t  *mel_close;     /* [nfft/2]          1 where a bin closes its segment */
    /* fp32 mode only (may be NULL in int mode) */
    const float    *f_window;      /* [nfft]                   */
    const float    *f_mel;         /* [nfilters][nfft/2+1]     */
    const float    *f_dct;         /* [nfilters][nfilters]     */
    float           f_preemph;     /* 0.96875                  */
    float           f_mel_floor;   /* mel energies are clamped to this before log2 (0 = off) */
} mfcc_b200_cfg;

/* Per-stage outputs of one clip (device pointers, any may be NULL).
 * The analogue of the collector processes of the reference's simulation bench,
 * mfcc/core/mfcc.py:176-183. */
typedef struct mfcc_b200_taps {
    int16_t  *d_window;   /* [K][nfft]        */
    int16_t  *d_fft;      /* [K][nfft/2][2]   */
    uint32_t *d_power;    /* [K][nfft/2]      */
    uint16_t *d_mel;      /* [K][nfilters]    */
    uint16_t *d_log2;     /* [K][nfilters]    */
    int16_t  *d_dct;      /* [K][nfilters]    */
} mfcc_b200_taps;

/* Replaces elaborating MFCC(...) (mfcc/core/mfcc.py:32-117) and loading the
 * bitstream: uploads the ROM tables to the current CUDA device. */
int  mfcc_b200_create(const mfcc_b200_cfg *cfg, void **handle);
void mfcc_b200_destroy(void *handle);

/*
 * Whole chain, fused, cepstra only.  Replaces the loop of mfcc_convert()
 * (software/main.c:128-166: write nfft then hop samples, read nceptrums words per
 * round) and the data path MFCC.sink -> MFCC.source (mfcc/core/mfcc.py:90-104).
 *   d_x      n_clips rows of n_samples int16, row pitch clip_stride (samples);
 *            every row is an independent stream that starts from reset
 *   n_frames frames to produce per clip.  Frame k covers samples hop*k .. hop*k+nfft-1;
 *            samples at or beyond n_samples read as zero, which is what the reference
 *            hosts feed after the end of the audio (software/main.c:135-144)
 *   d_out    n_clips rows of n_frames*nceptrums int16 (int mode) or float (fp32 mode),
 *            row pitch out_clip_stride (elements)
 *   taps     optional, requires n_clips == 1
 */
int  mfcc_b200_run(void *handle, const int16_t *d_x, int64_t clip_stride, int64_t n_samples,
                   int32_t n_clips, int64_t n_frames, void *d_out, int64_t out_clip_stride,
                   const mfcc_b200_taps *taps, void *stream);

/*
 * Same call with HOST buffers: the library stages the clips through pinned memory
 * and overlaps host->device copy, kernel and device->host copy in chunks of clips.
 * This is the entry point a host tool in the place of software/main.c binds.
 * Synchronous: returns when h_out is complete.
 */
int  mfcc_b200_run_host(void *handle, const int16_t *h_x, int64_t clip_stride, int64_t n_samples,
                        int32_t n_clips, int64_t n_frames, void *h_out, int64_t out_clip_stride);

/*
 * Staging pipeline of mfcc_b200_run_host: chunks of `chunk_bytes` of samples (default 16 MiB),
 * `slots` of them in flight (2..8, default 4) on three streams chained by events: host->device
 * copy, kernel, device->host copy.  0 keeps the current value.  Rows that are packed on the host
 * at the device pitch (clip_stride == n_samples rounded up to 8, out_clip_stride ==
 * n_frames*nceptrums) move with one plain copy per chunk and direction, anything else with one
 * pitched copy.  Whatever happens inside the call, no copy or kernel of it is in flight when it
 * returns (an error is returned after all three streams have drained).
 */
int  mfcc_b200_host_config(void *handle, int64_t chunk_bytes, int32_t slots);

/*
 * Page-locked host memory for the buffers of mfcc_b200_run_host / _step_host (what a host tool
 * in the place of software/main.c:115-127, which mallocs its sample buffer, would use).
 * write_combined != 0: uncached on the CPU side, for buffers the CPU only writes.
 */
int  mfcc_b200_host_alloc(void **p, int64_t bytes, int32_t write_combined);
void mfcc_b200_host_free(void *p);

/* Test hook: make chunk number `chunk` (0-based) of the next mfcc_b200_run_host calls fail after
 * its host->device copy has been queued (-1 = off).  tests/test_gpu_parity.py uses it to check
 * that an error inside the pipeline still leaves the caller's buffers quiescent. */
int  mfcc_b200_debug_fail_chunk(void *handle, int32_t chunk);

/*
 * One stage on its own (device buffers), the analogue of instantiating a single
 * stage module (constructors listed in the enum above).  n is the number of input
 * samples for PREEMPH and FRAME and the number of frames K otherwise; for FRAME
 * n_frames rows are produced (zero-extended like mfcc_b200_run).  `prev` is the
 * sample before d_in[0] for PREEMPH (0 after reset, preemph.py:18).
 */
int  mfcc_b200_stage(void *handle, int32_t stage, const void *d_in, void *d_out, int64_t n,
                     int64_t n_frames, int32_t prev, void *stream);

/*
 * Frame with any parameter set (mfcc/core/frame.py:49-60; the constructor defaults and the
 * reference's own bench use windowlen 400, stepsize 160, nfft 512, frame.py:50, :233-249):
 * d_out[k][j] = d_in[stepsize*k + j] for j < windowlen, 0 for windowlen <= j < nfft and for
 * samples at or beyond n.  Needs no handle.
 */
int  mfcc_b200_frame(const int16_t *d_in, int16_t *d_out, int64_t n, int64_t n_frames, int32_t windowlen,
                     int32_t stepsize, int32_t nfft, void *stream);

/*
 * PowerSpectrum(width=16, width_output) with any output width (mfcc/core/pow2.py:43-50; the
 * class default and the reference's own bench use 24, pow2.py:44, :74; MFCC uses 30):
 * d_out[i] = ((re*re + im*im) mod 2^32) >> (32 - width_output) for n_bins (re, im) pairs.
 */
int  mfcc_b200_power(const int16_t *d_in, uint32_t *d_out, int64_t n_bins, int32_t width_output, void *stream);

/*
 * Log2Fix(width, width_output) with any widths (mfcc/core/log.py:107-139; the reference's
 * own bench uses 37 and 20, log.py:143; MFCC uses 16 and 15): elementwise fixed-point log2
 * with precision = width_output - ceil(log2(width)) fraction bits, log2(0) = 0.
 */
int  mfcc_b200_log2fix(const uint64_t *d_in, uint32_t *d_out, int64_t n, int32_t width, int32_t width_output,
                       void *stream);

/*
 * WindowHamming(16, nfft, precision) with any nfft >= 8 (mfcc/core/window.py:10-20, :74-84):
 * d_out[r][n] = top 16 bits of d_in[r][n] * d_curve[n], d_curve[nfft] the multipliers the
 * quarter table decodes to (mfcc_b200.tables.window_curve).
 */
int  mfcc_b200_window(const int16_t *d_in, int16_t *d_out, int64_t n_rows, int32_t nfft, int32_t precision,
                      const uint16_t *d_curve, void *stream);

/*
 * FftStream(16, nfft) and DCTStream(16, nfft) with any power-of-two size (mfcc/core/
 * fft_stream.py:6-38, nfft 32..4096: d_in [n_rows][nfft] real -> d_out [n_rows][nfft/2][2];
 * mfcc/core/dct_stream.py:6-37, nfft 8..1024 inputs: d_in [n_rows][nfft] -> d_out
 * [n_rows][nfft], through the 4*nfft-point FFT of :29-34).  Both run the FFT core of
 * mfcc/misc/fft.py with its rounding butterfly; d_twiddle is the decoded twiddle ROM of that
 * core, int16 [size/2][2] = W(k) (re, im), size = nfft or 4*nfft (tables.twiddle_table).
 * Sizes below 32 points are refused: the reference's scheduler reads a stage's operands
 * before the previous stage has written them there (mfcc/misc/fft.py:251-308, :344).
 */
int  mfcc_b200_fft_stream(const int16_t *d_in, int16_t *d_out, int64_t n_rows, int32_t nfft,
                          const int16_t *d_twiddle, void *stream);
int  mfcc_b200_dct_stream(const int16_t *d_in, int16_t *d_out, int64_t n_rows, int32_t nfft,
                          const int16_t *d_twiddle, void *stream);

/*
 * FilterBank with any parameter set (mfcc/core/filterbank.py:37-63; the class defaults are
 * width 24 / output 24 / 16 taps, the reference's own bench uses width 31 / output 32,
 * filterbank.py:148; MFCC uses 30 / 16 / gain 18): the register-level algorithm of
 * filterbank.py:90-142 on d_power [n_frames][nbins] (width-bit unsigned values), giving
 * d_out [n_frames][ntap].  d_steps[nsteps] is the table of calc_filters() (:22-34) for the
 * filter points, maxvalrange and lo_bit = maxvalrange - (gain + width_output) as at :71, :122.
 */
int  mfcc_b200_filterbank(const uint32_t *d_power, uint32_t *d_out, int64_t n_frames, int32_t nbins, int32_t ntap,
                          int32_t width_mul, int32_t maxvalrange, int32_t lo_bit, int32_t width_output,
                          const uint64_t *d_steps, int32_t nsteps, void *stream);

/*
 * Streaming: one hop per stream per call, the mic2mfcc use of the core
 * (mfcc/targets/mic2mfcc.py:19-30).  d_state holds, per stream, the last nfft-hop
 * raw samples and the sample before them (what Preemph.odata and the Frame ring
 * hold, preemph.py:18-21, frame.py:65-73).  Zero state = the core after reset with
 * nfft-hop zeros already pushed in.
 *   d_hop  [n_streams][hop] int16     d_out [n_streams][nceptrums]
 * d_hop and d_out may be any pointers the device can dereference: device memory, or
 * page-locked host memory under unified addressing (the kernel then moves those bytes over
 * PCIe itself; writing d_out that way saves a copy operation per step, bench.py
 * "streaming_hostwrite").  A group of streams is stepped by offsetting all three pointers.
 */
int  mfcc_b200_state_bytes(void *handle, int32_t n_streams, int64_t *bytes);
int  mfcc_b200_step(void *handle, void *d_state, const int16_t *d_hop, int32_t n_streams,
                    void *d_out, void *stream);

/*
 * The same step with HOST buffers and the state kept inside the handle: what a plain-C host
 * without CUDA headers binds in place of the microphone front end (host/mic2mfcc.c).
 * The state is (re)allocated and zeroed when n_streams changes or reset != 0; h_hop == NULL
 * resets only.  Synchronous: returns when h_out [n_streams][nceptrums] is complete.
 */
int  mfcc_b200_step_host(void *handle, int32_t n_streams, const int16_t *h_hop, void *h_out, int32_t reset);

/*
 * ---- the steps after the path in the reference's tool chain ----
 * All buffers are device pointers owned by the caller.
 *
 * Cepstral lifter, software/lift.py:12-26, :38-39:
 *   d_out[r][n] = int16( d_lift[n] * d_in[r][n] ), float64 product, truncation toward zero
 *   like ndarray.astype(np.int16); d_lift[n] = 1 + (L/2) sin(pi n / L), L = 22 in the script.
 */
int  mfcc_b200_lifter(const int16_t *d_in, int16_t *d_out, int64_t n_rows, int32_t ncoeff,
                      const double *d_lift, void *stream);

/*
 * Per-coefficient standardisation over time, sklearn.preprocessing.scale(spec, axis=1)
 * followed by astype(np.int16) (software/genlibrosa.py:25-28), for n_clips matrices
 * [n_frames][ncoeff] of int16 (in_is_float = 0) or float32 (in_is_float = 1).
 */
int  mfcc_b200_standardize(const void *d_in, int32_t in_is_float, int16_t *d_out, int64_t n_frames,
                           int32_t ncoeff, int32_t n_clips, void *stream);

/* spec.astype(np.int16) of float32 features (software/genlibrosa.py:27): truncation toward zero. */
int  mfcc_b200_to_int16(const float *d_in, int16_t *d_out, int64_t n, void *stream);

/*
 * Serial framing of cepstrum columns, MagicInserter (mfcc/misc/magic.py:27-39) and the
 * byte order of the UART sender (mfcc/targets/mic2mfcc.py:58-62); what
 * cepstrum_get_column() parses (software/cepstrum.c:15-68): per column the 16-bit magic
 * (0xa55a) then ncoeff words, high byte first.  d_bytes holds n_cols * (ncoeff+1) * 2 bytes.
 */
int  mfcc_b200_uart_frame(const int16_t *d_cols, uint8_t *d_bytes, int64_t n_cols, int32_t ncoeff,
                          uint32_t magic, void *stream);

/*
 * Voice-activity gate of the host (software/cepstrum.c:13, :93-183), one step per new
 * column per stream: d_ring [n_streams][window] int16 keeps the last `window` (93 in
 * software/recv.c:17-18) values of coefficient 0, d_head [n_streams] the ring position;
 * d_power gets cepstrum_eval_power()'s sum (C int, wraps), d_active = power >= threshold
 * (POWER_THRESHOLD = 100000000).  Zero-initialised state = a ring of silence.
 */
int  mfcc_b200_vad_step(int16_t *d_ring, int32_t *d_head, const int16_t *d_cols, int32_t n_streams,
                        int32_t ncoeff, int32_t window, int32_t threshold, int32_t *d_power,
                        uint8_t *d_active, void *stream);

/*
 * ---- the librosa parameter set of the comparison script ----
 * software/genlibrosa.py:21-24 computes librosa.feature.mfcc(samples, sr=16000,
 * hop_length=170, n_mfcc=32) for every clip; librosa's defaults fill in the rest:
 * n_fft 2048, Hann window, centred frames over reflect padding, power spectrogram,
 * 128 Slaney mel filters, power_to_db (ref 1, amin 1e-10, top_db 80), orthonormal
 * DCT-II.  Tables are HOST arrays copied by create (mfcc_b200/tables.py lr_* build them).
 * Output is float32, time-major [n_frames][n_mfcc] per clip (librosa returns the
 * transpose).
 */
typedef struct mfcc_b200_librosa_cfg {
    int32_t struct_size;     /* sizeof(mfcc_b200_librosa_cfg) */
    int32_t n_fft;           /* 2048 (the size the kernels cover) */
    int32_t hop;             /* 170                     software/genlibrosa.py:23 */
    int32_t n_mels;          /* 1..128, librosa default 128 */
    int32_t n_mfcc;          /* 1..min(64, n_mels), 32 in software/genlibrosa.py:24 */
    int32_t center;          /* 1: frame t is centred on sample t*hop over reflect padding */
    float   amin;            /* 1e-10 */
    float   top_db;          /* 80; negative = no floor (librosa top_db=None) */
    const float *window;     /* [n_fft], already divided by the int16 full scale 32768 */
    const float *mel;        /* [n_mels][n_fft/2+1] dense filter matrix */
    const float *dct;        /* [n_mfcc][n_mels] */
} mfcc_b200_librosa_cfg;

typedef struct mfcc_b200_librosa_taps {   /* device pointers, any may be NULL; n_clips == 1 */
    float *d_power;          /* [K][n_fft/2+1] */
    float *d_logmel;         /* [K][n_mels], before the top_db floor */
} mfcc_b200_librosa_taps;

int     mfcc_b200_librosa_create(const mfcc_b200_librosa_cfg *cfg, void **handle);
void    mfcc_b200_librosa_destroy(void *handle);
/* librosa's frame count for a clip: 1 + n_samples / hop when centred. */
int64_t mfcc_b200_librosa_frames(void *handle, int64_t n_samples);
/* Bytes of device scratch mfcc_b200_librosa_run needs (per-clip maxima + the log-mel rows of one
 * chunk of clips: the dB floor depends on the maximum over the whole clip, so the path has two
 * passes with the rows in between). */
int     mfcc_b200_librosa_workspace(void *handle, int32_t n_clips, int64_t n_frames, int64_t *bytes);
/* Size of a chunk's log-mel rows in bytes; 0 keeps the value.  Default: the whole batch is one
 * chunk (fastest).  With ~48 MiB the rows stay in the 126 MB L2 between the passes and every
 * chunk reuses the same lines: less than half the DRAM traffic and a workspace of that size, for
 * ~8 % fewer frames/s.  Changes what mfcc_b200_librosa_workspace returns. */
int     mfcc_b200_librosa_config(void *handle, int64_t chunk_bytes);
/* Default shape (128 mel bands, 32 coefficients) only: which kernel forms the DCT product. 1 (the
 * default): warp-level tensor-core MMAs, TF32 with head/tail splits, float32-class accuracy;
 * 0: the FMA-pipe kernel. */
int     mfcc_b200_librosa_dct_path(void *handle, int32_t tensor_cores);
int     mfcc_b200_librosa_run(void *handle, const int16_t *d_x, int64_t clip_stride, int64_t n_samples,
                              int32_t n_clips, int64_t n_frames, void *d_work, float *d_out,
                              int64_t out_clip_stride, const mfcc_b200_librosa_taps *taps, void *stream);

/*
 * No counterpart in the reference: added for callers that
 * cannot allocate device memory (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  mfcc_b200_librosa_run without the workspace argument: the per-clip maxima
 * and the log-mel rows between the two passes live in a file-scope static __device__ array of
 * MFCC_B200_LIBROSA_STATIC_BYTES (16 MiB; every GPU that loads the library holds it until the
 * process exits).  The call allocates nothing (no cudaMalloc / cudaFree, no allocating library
 * routine); the array's device address was taken once, at mfcc_b200_librosa_create, on the device
 * that was current then, which must be current again here.  Limit:
 *     n_clips * n_frames * n_mels * 4 + round_up(4 * n_clips, 256) <= MFCC_B200_LIBROSA_STATIC_BYTES
 * (about 32 700 frames of 128 bands per call: five one-minute clips); anything larger is
 * rejected with MFCC_B200_EINVAL and a message that says so.  Same kernels and the same results
 * as mfcc_b200_librosa_run; no taps.  ALL calls on a device share the one array (it keeps the last
 * call's contents; what must start clean is cleared by the call itself): calls that overlap on
 * the GPU -- different streams, different handles, different host threads -- race.  Serialise
 * them on one stream.
 */
#define MFCC_B200_LIBROSA_STATIC_BYTES (16 * 1024 * 1024)
int     mfcc_b200_librosa_run_static(void *handle, const int16_t *d_x, int64_t clip_stride, int64_t n_samples,
                                     int32_t n_clips, int64_t n_frames, float *d_out, int64_t out_clip_stride,
                                     void *stream);

/* Number of kernels this library has launched since it was loaded (bench.py's
 * gpu_launches). */
int64_t     mfcc_b200_launch_count(void);
const char *mfcc_b200_last_error(void);
int         mfcc_b200_version(void);

#ifdef __cplusplus
}
#endif
#endif /* MFCC_B200_H */
