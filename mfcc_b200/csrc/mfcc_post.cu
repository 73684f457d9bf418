// Post-processing and live-stream helpers of the B200 MFCC path (include/mfcc_b200.h,
// "next rows" f2 and f3 of SURVEY.md 8(f)): the cepstral lifter and the per-coefficient
// standardisation the reference's scripts apply to the cepstra, the UART column framing
// of the microphone build and the host's voice-activity gate.  All of them are
// element-wise or tiny per-stream reductions: HBM-bound by construction.
#include <cstdint>
#include <cuda_runtime.h>

#include "../../include/mfcc_b200.h"

namespace mfccb_host {
int fail_ext(int code, const char *msg);
void count_launch();
}
using mfccb_host::fail_ext;

namespace {

__device__ __forceinline__ int16_t trunc16(double v) {
    // ndarray.astype(np.int16) on float64: truncate toward zero, keep the low 16 bits
    return (int16_t)(long long)v;
}

// software/lift.py:12-26, :38-39: out = int16(lift[n] * c), float64 product
__global__ void k_lifter(const int16_t *in, int16_t *out, long long total, int ncoeff, const double *lift) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x)
        out[i] = trunc16(lift[i % ncoeff] * (double)in[i]);
}

// sklearn.preprocessing.scale over time, per clip and coefficient (software/genlibrosa.py:25-28):
// one CTA per (clip, coefficient) column; mean, then centred second moment, in float64.
template <typename T>
__global__ void k_standardize(const T *in, int16_t *out, long long n_frames, int ncoeff, long long clip_stride) {
    const int clip = blockIdx.y, c = blockIdx.x;
    const T *src = in + (long long)clip * clip_stride + c;
    int16_t *dst = out + (long long)clip * clip_stride + c;
    __shared__ double red[32];
    __shared__ double s_mean, s_std;
    auto block_sum = [&](double v) {
        for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
        __syncthreads();
        double t = 0.0;
        if (threadIdx.x < 32) {
            t = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.0;
            for (int o = 16; o; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        }
        __syncthreads();
        return t;                       // valid in warp 0
    };
    double acc = 0.0;
    for (long long k = threadIdx.x; k < n_frames; k += blockDim.x) acc += (double)src[k * ncoeff];
    double tot = block_sum(acc);
    if (threadIdx.x == 0) s_mean = tot / (double)n_frames;
    __syncthreads();
    const double mean = s_mean;
    acc = 0.0;
    for (long long k = threadIdx.x; k < n_frames; k += blockDim.x) {
        const double d = (double)src[k * ncoeff] - mean;
        acc += d * d;
    }
    tot = block_sum(acc);
    if (threadIdx.x == 0) {
        const double sd = sqrt(tot / (double)n_frames);
        s_std = sd < 10.0 * 2.220446049250313e-16 ? 1.0 : sd;     // _handle_zeros_in_scale
    }
    __syncthreads();
    const double sd = s_std;
    for (long long k = threadIdx.x; k < n_frames; k += blockDim.x)
        dst[k * ncoeff] = trunc16(((double)src[k * ncoeff] - mean) / sd);
}

// mfcc/misc/magic.py:27-39 + mfcc/targets/mic2mfcc.py:58-62: magic word, then the
// coefficients, every 16-bit word high byte first
__global__ void k_uart_frame(const int16_t *cols, uint8_t *bytes, long long n_cols, int ncoeff, unsigned magic) {
    const long long words = n_cols * (ncoeff + 1);
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < words; i += (long long)gridDim.x * blockDim.x) {
        const long long col = i / (ncoeff + 1);
        const int j = (int)(i - col * (ncoeff + 1));
        const unsigned w = j == 0 ? magic : (unsigned)(uint16_t)cols[col * ncoeff + j - 1];
        bytes[2 * i] = (uint8_t)(w >> 8);
        bytes[2 * i + 1] = (uint8_t)w;
    }
}

// software/cepstrum.c:93-183.  Per stream a ring of the last `window` values of
// coefficient 0 (the host keeps whole columns; only c0 is ever summed).  head = next
// write position = oldest column.  After the new column is stored, power = sum of c0^2
// over the columns the host's loop visits: ring elements i in [size/3, 2*size/3) step
// ncoeff of a ring of size = window * ncoeff, i.e. columns ceil((size/3)/ncoeff) ..
// relative to the oldest.  C `int` arithmetic: wraps modulo 2^32.
__global__ void k_vad_step(int16_t *ring, int *head, const int16_t *cols, int n_streams, int ncoeff, int window,
                           int first_col, int last_col, int threshold, int *power, uint8_t *active) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_streams) return;
    int16_t *r = ring + (long long)s * window;
    int h = head[s];
    r[h] = cols[(long long)s * ncoeff];
    h = h + 1 == window ? 0 : h + 1;
    head[s] = h;
    unsigned p = 0;
    for (int c = first_col; c < last_col; ++c) {
        int idx = h + c;
        if (idx >= window) idx -= window;
        const int v = r[idx];
        p += (unsigned)(v * v);
    }
    power[s] = (int)p;
    active[s] = (int)p >= threshold;
}

// ndarray.astype(np.int16) of a float32 array (software/genlibrosa.py:27)
__global__ void k_to_int16(const float *in, int16_t *out, long long total) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x)
        out[i] = trunc16((double)in[i]);
}

// Frame(width, windowlen, stepsize, nfft), mfcc/core/frame.py:49-155, any parameter set:
// frame k = y[stepsize k .. stepsize k + windowlen) followed by nfft - windowlen zeros
// (the `padding` mux, frame.py:77, :120); samples past the end of the stream read as zero
__global__ void k_frame_general(const int16_t *y, int16_t *fr, long long n, long long n_frames, int windowlen,
                                int stepsize, int nfft) {
    const long long total = n_frames * nfft;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long k = i / nfft;
        const int j = (int)(i - k * nfft);
        const long long s = k * stepsize + j;
        fr[i] = (j < windowlen && s < n) ? y[s] : (int16_t)0;
    }
}

// 128-bit form for even windowlen / stepsize, nfft a multiple of 8, y on a 32-bit and fr on a
// 16-byte boundary: eight outputs per thread and step, sample pairs as 32-bit words
__global__ void k_frame_general_v(const int16_t *y, int16_t *fr, long long n, long long n_frames, int windowlen,
                                  int stepsize, int nfft) {
    const int vpf = nfft >> 3;
    const long long total = n_frames * vpf;
    const unsigned *y32 = reinterpret_cast<const unsigned *>(y);
    uint4 *f4 = reinterpret_cast<uint4 *>(fr);
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const long long k = i / vpf;
        const int j = (int)(i - k * vpf) << 3;
        const long long s = k * stepsize + j;
        unsigned o[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int jj = j + 2 * q;
            const long long ss = s + 2 * q;
            if (jj + 1 < windowlen && ss + 1 < n) o[q] = y32[ss >> 1];
            else {
                const unsigned lo = (jj < windowlen && ss < n) ? (unsigned short)y[ss] : 0u;
                const unsigned hi = (jj + 1 < windowlen && ss + 1 < n) ? (unsigned short)y[ss + 1] : 0u;
                o[q] = lo | (hi << 16);
            }
        }
        f4[i] = make_uint4(o[0], o[1], o[2], o[3]);
    }
}

// PowerSpectrum(width=16, width_output), mfcc/core/pow2.py:32, :64: the 32-bit sum of the two
// squares (it wraps for re = im = -32768), top width_output bits
__global__ void k_power_general(const int16_t *f, uint32_t *pw, long long total, int shift) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int re = f[2 * i], im = f[2 * i + 1];
        pw[i] = ((unsigned)(re * re) + (unsigned)(im * im)) >> shift;
    }
}

// four bins per thread and step (128-bit loads and stores); the launcher adds the scalar tail
__global__ void k_power_general_v(const int16_t *f, uint32_t *pw, long long total4, int shift) {
    const uint4 *f4 = reinterpret_cast<const uint4 *>(f);
    uint4 *p4 = reinterpret_cast<uint4 *>(pw);
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total4; i += (long long)gridDim.x * blockDim.x) {
        const uint4 a = f4[i];
        const unsigned as[4] = {a.x, a.y, a.z, a.w};
        unsigned o[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int re = (int)(short)(as[q] & 0xffffu), im = (int)as[q] >> 16;
            o[q] = ((unsigned)(re * re) + (unsigned)(im * im)) >> shift;
        }
        p4[i] = make_uint4(o[0], o[1], o[2], o[3]);
    }
}

// Log2Fix(width, width_output) -> Log2FixCalc, mfcc/core/log.py:19-104, :107-139, any widths:
// the FSM's shift-right normalisation, then precision-1 squarings of a (precision+1)-bit value
__global__ void k_log2_general(const unsigned long long *in, unsigned *out, long long total, int precision,
                               int width_output) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const unsigned long long v = in[i];
        unsigned long long x = (v == 0 ? 1ull : v) << precision;          // log.py:123-126
        unsigned long long o = 0, b = 1ull << (precision - 1);
        while (x >> (precision + 1)) { x >>= 1; o += 1ull << precision; }  // SHIFT-RIGHT
        unsigned long long z = x;
        for (int cnt = precision - 1; cnt > 0; --cnt) {                   // CALC-1 / CALC-2
            z &= (2ull << precision) - 1;                                 // the multiplier takes precision+1 bits
            const unsigned long long c = z * z;
            if ((c >> (2 * precision + 1)) & 1) { z = c >> (precision + 1); o += b; }
            else z = c >> precision;
            b >>= 1;
        }
        out[i] = (unsigned)(width_output >= 32 ? o : o & ((1ull << width_output) - 1));   // log.py:133
    }
}

// FilterBank with any parameter set (mfcc/core/filterbank.py:37-144), register by register:
// the 2*wm-bit weight generator (:90-115) and the two maxvalrange-bit accumulators
// (:119-134), output taken from o_regb before the update and suppressed for segment 0
// (:139-140).  One thread per frame, 128-bit registers (maxvalrange reaches 68 bits for the
// reference bench's FilterBank(31, 32)).  Frames are independent: `last` clears the
// generator and the first peak of the next frame overwrites o_regb.
__global__ void k_filterbank_general(const uint32_t *power, uint32_t *out, long long n_frames, int nbins, int ntap,
                                     int wm, int maxvalrange, int lo_bit, int width_output,
                                     const unsigned long long *steps, int nsteps) {
    typedef unsigned __int128 u128;
    const u128 rmask = maxvalrange >= 128 ? ~(u128)0 : (((u128)1 << maxvalrange) - 1);
    const unsigned long long amask = (2 * wm >= 64) ? ~0ull : ((1ull << (2 * wm)) - 1);
    const unsigned long long top = (1ull << wm) - 1;
    const u128 omask = (((u128)1 << width_output) - 1);
    for (long long k = blockIdx.x * (long long)blockDim.x + threadIdx.x; k < n_frames; k += (long long)gridDim.x * blockDim.x) {
        const uint32_t *p = power + k * nbins;
        uint32_t *o = out + k * ntap;
        unsigned long long acc = 0;
        int adr = 0, nout = 0;
        u128 rega = 0, regb = 0;
        for (int b = 0; b < nbins; ++b) {
            const unsigned long long d = p[b];
            const bool last = b == nbins - 1;
            const unsigned long long w = acc >> wm;
            const bool highest = w == top;
            const u128 c = (u128)d * w;
            const int adr0 = adr;
            if (highest || last) { adr = last ? 0 : adr + 1; acc = 0; }
            else acc = (acc + (adr < nsteps ? steps[adr] : 0ull)) & amask;
            const uint32_t o_data = (uint32_t)((regb >> lo_bit) & omask);
            if (highest || last) {
                if (adr0 != 0 && nout < ntap) o[nout++] = o_data;
                regb = (rega + ((u128)d << wm)) & rmask;
                rega = 0;
            } else {
                rega = (rega + c) & rmask;
                regb = (regb + ((u128)d << wm) - c) & rmask;
            }
        }
        for (; nout < ntap; ++nout) o[nout] = 0;
    }
}

// WindowHamming(16, nfft, precision), mfcc/core/window.py:74-75, :84: the top 16 bits of the
// signed16 x unsigned(precision+1) product, curve[] from the caller (tables.window_curve).
__global__ void k_window_general(const int16_t *in, int16_t *out, long long total, int nfft, int shift,
                                 const uint16_t *curve) {
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x)
        out[i] = (int16_t)(((int)in[i] * (int)curve[i & (nfft - 1)]) >> shift);
}

// The FFT core of mfcc/misc/fft.py for any power-of-two size (32..4096; below 32 points the
// reference's scheduler has a read-before-write hazard, not reproduced), width 16: bit-reversed
// load (:413-418), log2(size) in-place stages of the rounding butterfly
//   t = (x1 * w + 8191) >> 14,  y0 = wrap16((x0 + t) >> 1),  y1 = wrap16((x0 - t) >> 1)
// (:93-96, :140-192), twiddle address k * (size/2 >> stage) (:310-316).  One CTA per row, the
// row in shared memory as packed (re, im) halves.  DCT = false: FftStream (fft_stream.py:30-38),
// real input [size], bins 0..size/2-1 out as (re, im).  DCT = true: DCTStream(16, size/4)
// (dct_stream.py:23-37), input L[size/4] placed at v[2n+1] and v[size-1-2n], real part of
// bins 0..size/4-1 out.
template <bool DCT>
__global__ void k_fft_general(const int16_t *in, int16_t *out, long long n_rows, int nb, const short2 *tw) {
    extern __shared__ unsigned v[];
    const int size = 1 << nb, half = size >> 1;
    const int n_in = DCT ? size >> 2 : size, n_out = DCT ? size >> 2 : half;
    for (long long row = blockIdx.x; row < n_rows; row += gridDim.x) {
        const int16_t *src = in + row * n_in;
        for (int n = threadIdx.x; n < size; n += blockDim.x) {
            int val;
            if (DCT) val = (n & 1) ? src[n < half ? (n >> 1) : ((size - 1 - n) >> 1)] : 0;
            else val = src[n];
            v[__brev((unsigned)n) >> (32 - nb)] = (unsigned)val & 0xffffu;          // imag = 0
        }
        __syncthreads();
        for (int s = 0; s < nb; ++s) {
            for (int b = threadIdx.x; b < half; b += blockDim.x) {
                const int k = b & ((1 << s) - 1);
                const int p0 = ((b >> s) << (s + 1)) + k, p1 = p0 + (1 << s);
                const short2 w = tw[k << (nb - 1 - s)];
                const unsigned a = v[p0], c = v[p1];
                const int x0r = (int16_t)(a & 0xffffu), x0i = (int)a >> 16;
                const int x1r = (int16_t)(c & 0xffffu), x1i = (int)c >> 16;
                const int tr = (x1r * w.x - x1i * w.y + 8191) >> 14;
                const int ti = (x1r * w.y + x1i * w.x + 8191) >> 14;
                v[p0] = ((unsigned)((x0r + tr) >> 1) & 0xffffu) | ((unsigned)((x0i + ti) >> 1) << 16);
                v[p1] = ((unsigned)((x0r - tr) >> 1) & 0xffffu) | ((unsigned)((x0i - ti) >> 1) << 16);
            }
            __syncthreads();
        }
        if (DCT) {
            for (int n = threadIdx.x; n < n_out; n += blockDim.x) out[row * n_out + n] = (int16_t)(v[n] & 0xffffu);
        } else {
            unsigned *dst = reinterpret_cast<unsigned *>(out) + row * n_out;         // (re, im) pairs
            for (int n = threadIdx.x; n < n_out; n += blockDim.x) dst[n] = v[n];
        }
        __syncthreads();
    }
}

inline int grid_for(long long total, int block) {
    long long g = (total + block - 1) / block;
    if (g > 148 * 16) g = 148 * 16;
    return g < 1 ? 1 : (int)g;
}

#define POST_LAUNCH_CHECK()                                                          \
    do {                                                                             \
        mfccb_host::count_launch();                                                  \
        cudaError_t e_ = cudaGetLastError();                                         \
        if (e_ != cudaSuccess) return fail_ext(MFCC_B200_ECUDA, cudaGetErrorString(e_)); \
    } while (0)

}  // namespace

extern "C" {

int mfcc_b200_lifter(const int16_t *d_in, int16_t *d_out, int64_t n_rows, int32_t ncoeff, const double *d_lift,
                     void *stream)
{
    if (n_rows < 0 || ncoeff < 1) return fail_ext(MFCC_B200_EINVAL, "bad size");
    if (n_rows == 0) return MFCC_B200_OK;
    if (!d_in || !d_out || !d_lift) return fail_ext(MFCC_B200_EINVAL, "null buffer");
    k_lifter<<<grid_for(n_rows * ncoeff, 256), 256, 0, (cudaStream_t)stream>>>(d_in, d_out, n_rows * ncoeff, ncoeff, d_lift);
    POST_LAUNCH_CHECK();
    return MFCC_B200_OK;
}

int mfcc_b200_frame(const int16_t *d_in, int16_t *d_out, int64_t n, int64_t n_frames, int32_t windowlen,
                    int32_t stepsize, int32_t nfft, void *stream)
{
    if (n < 0 || n_frames < 0) return fail_ext(MFCC_B200_EINVAL, "negative size");
    if (windowlen < 1 || stepsize < 1 || nfft < windowlen) return fail_ext(MFCC_B200_EINVAL, "need 1 <= windowlen <= nfft, stepsize >= 1");
    if (n_frames == 0) return MFCC_B200_OK;
    if (!d_out || (n > 0 && !d_in)) return fail_ext(MFCC_B200_EINVAL, "null buffer");
    const bool vec = !(windowlen & 1) && !(stepsize & 1) && !(nfft & 7) &&
                     !(reinterpret_cast<uintptr_t>(d_in) & 3) && !(reinterpret_cast<uintptr_t>(d_out) & 15);
    if (vec)
        k_frame_general_v<<<grid_for(n_frames * (nfft >> 3), 256), 256, 0, (cudaStream_t)stream>>>(d_in, d_out, n, n_frames,
                                                                                                 windowlen, stepsize, nfft);
    else
        k_frame_general<<<grid_for(n_frames * nfft, 256), 256, 0, (cudaStream_t)stream>>>(d_in, d_out, n, n_frames, windowlen,
                                                                                        stepsize, nfft);
    POST_LAUNCH_CHECK();
    return MFCC_B200_OK;
}

int mfcc_b200_window(const int16_t *d_in, int16_t *d_out, int64_t n_rows, int32_t nfft, int32_t precision,
                     const uint16_t *d_curve, void *stream)
{
    if (n_rows < 0 || nfft < 8 || (nfft & (nfft - 1)) || precision < 1 || precision > 14)
        return fail_ext(MFCC_B200_EINVAL, "need n_rows >= 0, nfft a power of two >= 8 and 1 <= precision <= 14");
    if (n_rows == 0) return MFCC_B200_OK;
    if (!d_in || !d_out || !d_curve) return fail_ext(MFCC_B200_EINVAL, "null buffer");
    k_window_general<<<grid_for(n_rows * nfft, 256), 256, 0, (cudaStream_t)stream>>>(d_in, d_out, n_rows * nfft, nfft,
                                                                                        precision + 1, d_curve);
    POST_LAUNCH_CHECK();
    return MFCC_B200_OK;
}

static int fft_general(bool dct, const int16_t *d_in, int16_t *d_out, int64_t n_rows, int32_t size,
                       const int16_t *d_twiddle, void *stream)
{
    if (n_rows < 0 || size < 32 || size > 4096 || (size & (size - 1)))
        return fail_ext(MFCC_B200_EINVAL, dct ? "need n_rows >= 0 and nfft a power of two in 8..1024"
                                              : "need n_rows >= 0 and nfft a power of two in 32..4096");
    if (n_rows == 0) return MFCC_B200_OK;
    if (!d_in || !d_out || !d_twiddle) return fail_ext(MFCC_B200_EINVAL, "null buffer");
    if (((uintptr_t)d_twiddle | (uintptr_t)d_out) & 3) return fail_ext(MFCC_B200_EINVAL, "d_twiddle and d_out must be 4-byte aligned");
    int nb = 0;
    while ((1 << nb) < size) ++nb;
    const int threads = size / 2 < 256 ? (size / 2 < 32 ? 32 : size / 2) : 256;
    const int grid = (int)(n_rows < 148 * 16 ? n_rows : 148 * 16);
    const size_t smem = (size_t)size * sizeof(unsigned);
    if (dct) k_fft_general<true><<<grid, threads, smem, (cudaStream_t)stream>>>(d_in, d_out, n_rows, nb, (const short2 *)d_twiddle);
    else k_fft_general<false><<<grid, threads, smem, (cudaStream_t)stream>>>(d_in, d_out, n_rows, nb, (const short2 *)d_twiddle);
    POST_LAUNCH_CHECK();
    return MFCC_B200_OK;
}

int mfcc_b200_fft_stream(const int16_t *d_in, int16_t *d_out, int64_t n_rows, int32_t nfft, const int16_t *d_twiddle, void *stream)
{
    return fft_general(false, d_in, d_out, n_rows, nfft, d_twiddle, stream);
}

int mfcc_b200_dct_stream(const int16_t *d_in, int16_t *d_out, int64_t n_rows, int32_t nfft, const int16_t *d_twiddle, void *stream)
{
    if (nfft < 8 || nfft > 1024) return fail_ext(MFCC_B200_EINVAL, "need n_rows >= 0 and nfft a power of two in 8..1024");
    return fft_general(true, d_in, d_out, n_rows, 4 * nfft, d_twiddle, stream);
}

int mfcc_b200_power(const int16_t *d_in, uint32_t *d_out, int64_t n_bins, int32_t width_output, void *stream)
{
    if (n_bins < 0 || width_output < 1 || width_output > 32) return fail_ext(MFCC_B200_EINVAL, "need n_bins >= 0 and 1 <= width_output <= 32");
    if (n_bins == 0) return MFCC_B200_OK;
    if (!d_in || !d_out) return fail_ext(MFCC_B200_EINVAL, "null buffer");
    const bool vec = !((reinterpret_cast<uintptr_t>(d_in) | reinterpret_cast<uintptr_t>(d_out)) & 15) && n_bins >= 4;
    const int64_t nv = vec ? (n_bins & ~(int64_t)3) : 0;               // bins done four at a time
    if (nv) k_power_general_v<<<grid_for(nv / 4, 256), 256, 0, (cudaStream_t)stream>>>(d_in, d_out, nv / 4, 32 - width_output);
    if (n_bins > nv)
        k_power_general<<<grid_for(n_bins - nv, 256), 256, 0, (cudaStream_t)stream>>>(d_in + 2 * nv, d_out + nv, n_bins - nv,
                                                                                      32 - width_output);
    POST_LAUNCH_CHECK();
    return MFCC_B200_OK;
}

int mfcc_b200_log2fix(const uint64_t *d_in, uint32_t *d_out, int64_t n, int32_t width, int32_t width_output, void *stream)
{
    if (n < 0 || width < 2 || width_output < 1 || width_output > 32) return fail_ext(MFCC_B200_EINVAL, "bad widths");
    int lg = 0;
    while ((1 << lg) < width) ++lg;                                   // ceil(log2(width)), log.py:114
    const int precision = width_output - lg;
    if (precision < 1 || precision > 30 || width + precision > 63)
        return fail_ext(MFCC_B200_EUNSUPPORTED, "need 1 <= width_output - ceil(log2(width)) <= 30 and width + precision <= 63");
    if (n == 0) return MFCC_B200_OK;
    if (!d_in || !d_out) return fail_ext(MFCC_B200_EINVAL, "null buffer");
    k_log2_general<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>((const unsigned long long *)d_in, d_out, n, precision,
                                                                       width_output);
    POST_LAUNCH_CHECK();
    return MFCC_B200_OK;
}

int mfcc_b200_filterbank(const uint32_t *d_power, uint32_t *d_out, int64_t n_frames, int32_t nbins, int32_t ntap,
                         int32_t width_mul, int32_t maxvalrange, int32_t lo_bit, int32_t width_output,
                         const uint64_t *d_steps, int32_t nsteps, void *stream)
{
    if (n_frames < 0 || nbins < 1 || ntap < 1 || nsteps < 1) return fail_ext(MFCC_B200_EINVAL, "bad size");
    if (width_mul < 1 || width_mul > 31 || width_output < 1 || width_output > 32 || maxvalrange < 1 || maxvalrange > 127 ||
        lo_bit < 0 || lo_bit + width_output > 128)
        return fail_ext(MFCC_B200_EUNSUPPORTED, "need width_mul <= 31, width_output <= 32, maxvalrange <= 127");
    if (n_frames == 0) return MFCC_B200_OK;
    if (!d_power || !d_out || !d_steps) return fail_ext(MFCC_B200_EINVAL, "null buffer");
    k_filterbank_general<<<grid_for(n_frames, 128), 128, 0, (cudaStream_t)stream>>>(
        d_power, d_out, n_frames, nbins, ntap, width_mul, maxvalrange, lo_bit, width_output,
        (const unsigned long long *)d_steps, nsteps);
    POST_LAUNCH_CHECK();
    return MFCC_B200_OK;
}

int mfcc_b200_to_int16(const float *d_in, int16_t *d_out, int64_t n, void *stream)
{
    if (n < 0) return fail_ext(MFCC_B200_EINVAL, "bad size");
    if (n == 0) return MFCC_B200_OK;
    if (!d_in || !d_out) return fail_ext(MFCC_B200_EINVAL, "null buffer");
    k_to_int16<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(d_in, d_out, n);
    POST_LAUNCH_CHECK();
    return MFCC_B200_OK;
}

int mfcc_b200_standardize(const void *d_in, int32_t in_is_float, int16_t *d_out, int64_t n_frames, int32_t ncoeff,
                          int32_t n_clips, void *stream)
{
    if (n_frames < 0 || ncoeff < 1 || n_clips < 0) return fail_ext(MFCC_B200_EINVAL, "bad size");
    if (n_frames == 0 || n_clips == 0) return MFCC_B200_OK;
    if (!d_in || !d_out) return fail_ext(MFCC_B200_EINVAL, "null buffer");
    if (n_clips > 65535) return fail_ext(MFCC_B200_EINVAL, "more than 65535 clips per call");
    const dim3 grid((unsigned)ncoeff, (unsigned)n_clips);
    const long long pitch = n_frames * ncoeff;
    if (in_is_float) k_standardize<float><<<grid, 256, 0, (cudaStream_t)stream>>>((const float *)d_in, d_out, n_frames, ncoeff, pitch);
    else             k_standardize<int16_t><<<grid, 256, 0, (cudaStream_t)stream>>>((const int16_t *)d_in, d_out, n_frames, ncoeff, pitch);
    POST_LAUNCH_CHECK();
    return MFCC_B200_OK;
}

int mfcc_b200_uart_frame(const int16_t *d_cols, uint8_t *d_bytes, int64_t n_cols, int32_t ncoeff, uint32_t magic,
                         void *stream)
{
    if (n_cols < 0 || ncoeff < 1) return fail_ext(MFCC_B200_EINVAL, "bad size");
    if (n_cols == 0) return MFCC_B200_OK;
    if (!d_cols || !d_bytes) return fail_ext(MFCC_B200_EINVAL, "null buffer");
    k_uart_frame<<<grid_for(n_cols * (ncoeff + 1), 256), 256, 0, (cudaStream_t)stream>>>(d_cols, d_bytes, n_cols, ncoeff, magic & 0xffffu);
    POST_LAUNCH_CHECK();
    return MFCC_B200_OK;
}

int mfcc_b200_vad_step(int16_t *d_ring, int32_t *d_head, const int16_t *d_cols, int32_t n_streams, int32_t ncoeff,
                       int32_t window, int32_t threshold, int32_t *d_power, uint8_t *d_active, void *stream)
{
    if (n_streams < 0 || ncoeff < 1 || window < 3) return fail_ext(MFCC_B200_EINVAL, "bad size");
    if (n_streams == 0) return MFCC_B200_OK;
    if (!d_ring || !d_head || !d_cols || !d_power || !d_active) return fail_ext(MFCC_B200_EINVAL, "null buffer");
    // elements i in [size/3, 2*size/3) step ncoeff of the column-major ring (cepstrum.c:166-178)
    const long long size = (long long)window * ncoeff;
    const long long lo = size / 3, hi = 2 * size / 3;
    if (lo % ncoeff) return fail_ext(MFCC_B200_EUNSUPPORTED, "window*ncoeff/3 must be a multiple of ncoeff (the host then sums c0)");
    const int first_col = (int)(lo / ncoeff);
    const int last_col = (int)((hi - lo + ncoeff - 1) / ncoeff) + first_col;
    k_vad_step<<<(n_streams + 127) / 128, 128, 0, (cudaStream_t)stream>>>(d_ring, d_head, d_cols, n_streams, ncoeff, window,
                                                                        first_col, last_col, threshold, d_power, d_active);
    POST_LAUNCH_CHECK();
    return MFCC_B200_OK;
}

}  // extern "C"
