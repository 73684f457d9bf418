/*
 * mic2mfcc on the GPU: what the microphone build of the reference puts on its serial line,
 * computed from raw audio on the host.
 *
 *     mic2mfcc [nceptrums] < audio.s16le > columns.bin
 *
 * The FPGA top (mfcc/targets/mic2mfcc.py:19-30, :53-74) feeds an I2S microphone into
 * MFCC(nfft=512, nfilters=32, nceptrums=16) and sends every cepstrum column over a UART as
 * the 16-bit magic 0xa55a followed by the coefficients, every word high byte first
 * (mfcc/misc/magic.py:27-39); software/cepstrum.c:15-68 parses exactly that.  This tool reads
 * 16 kHz mono 16-bit little-endian PCM from stdin, one 170-sample hop at a time, and writes
 * the same byte stream to stdout.  From reset the core's first frame is samples 0..511 and
 * every later one starts 170 further on (mfcc/core/frame.py).  A step of the library frames
 * the LAST 512 samples it was given, and 512 = 3 * 170 + 2, so the tool first pushes a
 * priming hop that ends in samples 0 and 1 (zeros before them, the reset value of the
 * pre-emphasis register included); the third full hop after that completes samples 0..511
 * and its column is the first one sent, like the hardware's.
 *
 * Plain C11 on top of include/mfcc_b200.h (mfcc_b200_step_host keeps the stream state on
 * the GPU); no CUDA headers needed.
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "mfcc_b200.h"
#include "mfcc_tables_32.h"

enum { NFFT = 512, STEPSIZE = 170, MAGIC = 0xa55a };

int main(int argc, char **argv)
{
    const int nc = argc > 1 ? atoi(argv[1]) : 16;
    void *h = NULL;
    int16_t hop[STEPSIZE], col[32];
    unsigned char wire[2 * 33];
    long hops = 0, sent = 0;
    if (nc < 1 || nc > 32) { fprintf(stderr, "Usage: %s [nceptrums 1..32] < audio.s16le > columns.bin\n", argv[0]); return 1; }
    mfcc_b200_cfg cfg = {
        .struct_size = sizeof cfg, .mode = MFCC_B200_MODE_INT, .nfft = NFFT, .hop = STEPSIZE,
        .nfilters = 32, .nceptrums = nc, .window_shift = 9, .pow_shift = 2, .mel_wbits = 30,
        .mel_lo_bit = MFCC_TABLES_MEL_LO_BIT, .mel_out_bits = 16, .log_precision = 11, .log_out_bits = 15,
        .window_curve = mfcc_window_curve, .tw_fft = mfcc_tw_fft, .tw_dct = mfcc_tw_dct,
        .mel_weight = mfcc_mel_weight, .mel_close = mfcc_mel_close, .f_preemph = 0.96875f };
    if (mfcc_b200_create(&cfg, &h)) { fprintf(stderr, "mfcc_b200_create: %s\n", mfcc_b200_last_error()); return 2; }
    memset(hop, 0, sizeof hop);
    if (fread(hop + STEPSIZE - 2, sizeof hop[0], 2, stdin) != 2) return 0;
    if (mfcc_b200_step_host(h, 1, hop, col, 1)) {
        fprintf(stderr, "mfcc_b200_step_host: %s\n", mfcc_b200_last_error());
        return 2;
    }
    while (fread(hop, sizeof hop[0], STEPSIZE, stdin) == STEPSIZE) {
        if (mfcc_b200_step_host(h, 1, hop, col, 0)) {
            fprintf(stderr, "mfcc_b200_step_host: %s\n", mfcc_b200_last_error());
            return 2;
        }
        if (++hops < 3) continue;                       /* window not full yet */
        wire[0] = MAGIC >> 8; wire[1] = MAGIC & 0xff;
        for (int i = 0; i < nc; i++) {
            wire[2 + 2 * i] = (unsigned char)((uint16_t)col[i] >> 8);
            wire[3 + 2 * i] = (unsigned char)col[i];
        }
        fwrite(wire, 1, 2 * (size_t)(nc + 1), stdout);
        ++sent;
    }
    mfcc_b200_destroy(h);
    fprintf(stderr, "%ld hops in, %ld columns out\n", hops, sent);
    return 0;
}
