/*
 * wav2mfcc on the GPU: the host tool of the reference (software/main.c) with the FPGA
 * round trips replaced by libmfcc_b200.so.
 *
 *     wav2mfcc <wavdir>
 *
 * Same contract as the reference tool: every file whose name ends in ".wav" under <wavdir>
 * (directories walked recursively, software/main.c:206-247) that is 16 kHz, mono, 16-bit PCM
 * (checks and messages of software/main.c:66-92) gets a headerless native-endian int16
 * file "<name>.mfcc" next to it with (N - 512) / 170 + 2 rows of 32 coefficients
 * (software/main.c:95, :128-166: the round in which the audio ends is zero-filled), one
 * row for files shorter than a frame.  Unlike the reference, a rejected file does not end
 * the walk of its directory.
 *
 * The ROM tables come from mfcc_tables_32.h (python -m mfcc_b200.dump_tables); everything
 * else is the C ABI of include/mfcc_b200.h.  Plain C11, no CUDA headers needed.
 */
#include <dirent.h>
#include <errno.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>

#include "mfcc_b200.h"
#include "mfcc_tables_32.h"

enum { NFFT = 512, STEPSIZE = 170, NCEPSTRUMS = 32, SAMPLERATE = 16000 };

static uint32_t rd32(const unsigned char *p) { return p[0] | p[1] << 8 | p[2] << 16 | (uint32_t)p[3] << 24; }
static uint16_t rd16(const unsigned char *p) { return (uint16_t)(p[0] | p[1] << 8); }

/* Reads a RIFF/WAVE file; returns the samples (malloc'd) or NULL after printing why not. */
static int16_t *wav_load(const char *path, size_t *nsamples)
{
    FILE *f = fopen(path, "rb");
    unsigned char hdr[12], ck[8], fmt[16];
    int have_fmt = 0;
    int16_t *data = NULL;
    if (!f) { printf("Failed to open: %s\n", path); return NULL; }
    if (fread(hdr, 1, 12, f) != 12 || memcmp(hdr, "RIFF", 4) || memcmp(hdr + 8, "WAVE", 4)) {
        printf("Unexpected format: not a RIFF/WAVE file (%s)\n", path);
        goto done;
    }
    while (fread(ck, 1, 8, f) == 8) {
        const uint32_t size = rd32(ck + 4);
        if (!memcmp(ck, "fmt ", 4)) {
            if (size < 16 || fread(fmt, 1, 16, f) != 16) break;
            have_fmt = 1;
            if (rd16(fmt) != 1) { printf("Unexpected format: %04x\n", rd16(fmt)); goto done; }
            if (rd16(fmt + 14) != 16) { printf("Unexpected samplesize: %d\n", rd16(fmt + 14) / 8); goto done; }
            if (rd32(fmt + 4) != SAMPLERATE) { printf("Unexpected samplerate: %u\n", rd32(fmt + 4)); goto done; }
            if (rd16(fmt + 2) != 1) { printf("Unexpected channel count: %d\n", rd16(fmt + 2)); goto done; }
            fseek(f, (long)(size - 16 + (size & 1)), SEEK_CUR);
        } else if (!memcmp(ck, "data", 4)) {
            if (!have_fmt) break;
            data = malloc(size ? size : 2);
            if (!data) { printf("Out of memory (%s)\n", path); goto done; }
            *nsamples = fread(data, 2, size / 2, f);      /* little-endian hosts, like the reference tool */
            goto done;
        } else {
            fseek(f, (long)(size + (size & 1)), SEEK_CUR);
        }
    }
    printf("Unexpected format: no fmt/data chunk (%s)\n", path);
done:
    fclose(f);
    return data;
}

static int convert(void *h, const char *in, const char *out)
{
    size_t n = 0;
    int16_t *x = wav_load(in, &n);
    if (!x) return -1;
    const int64_t rows = n < NFFT ? 1 : (int64_t)((n - NFFT) / STEPSIZE) + 2;
    int16_t *cep = malloc((size_t)rows * NCEPSTRUMS * sizeof *cep);
    int16_t zero[8] = {0};
    int rc = -1;
    if (cep) {
        /* an empty file still makes one zero-filled round (software/main.c:128-144) */
        rc = mfcc_b200_run_host(h, n ? x : zero, (int64_t)(n ? n : 8), (int64_t)n, 1, rows, cep, rows * NCEPSTRUMS);
        if (rc) printf("mfcc_b200_run_host: %s\n", mfcc_b200_last_error());
    }
    if (!rc) {
        FILE *f = fopen(out, "wb");
        if (!f) { printf("Failed to open %s: %s\n", out, strerror(errno)); rc = -1; }
        else { fwrite(cep, sizeof *cep, (size_t)rows * NCEPSTRUMS, f); fclose(f); }
    }
    free(cep);
    free(x);
    return rc;
}

static void walk(void *h, const char *dir, int *done, int *rejected)
{
    DIR *d = opendir(dir);
    struct dirent *e;
    if (!d) return;
    while ((e = readdir(d)) != NULL) {
        char path[4096], out[4096];
        struct stat st;
        const size_t len = strlen(e->d_name);
        if (!strcmp(e->d_name, ".") || !strcmp(e->d_name, "..")) continue;
        if (snprintf(path, sizeof path, "%s/%s", dir, e->d_name) >= (int)sizeof path || stat(path, &st)) continue;
        if (S_ISDIR(st.st_mode)) { walk(h, path, done, rejected); continue; }
        if (len < 4 || strcmp(e->d_name + len - 4, ".wav")) continue;
        snprintf(out, sizeof out, "%.*smfcc", (int)strlen(path) - 3, path);
        printf("%s %s\n", path, out);
        if (convert(h, path, out)) ++*rejected; else ++*done;
    }
    closedir(d);
}

int main(int argc, char **argv)
{
    void *h = NULL;
    int done = 0, rejected = 0;
    if (argc != 2) { printf("Usage: %s <wavdir>\n", argv[0]); return 1; }
    mfcc_b200_cfg cfg = {
        .struct_size = sizeof cfg, .mode = MFCC_B200_MODE_INT, .nfft = NFFT, .hop = STEPSIZE,
        .nfilters = 32, .nceptrums = NCEPSTRUMS, .window_shift = 9, .pow_shift = 2, .mel_wbits = 30,
        .mel_lo_bit = MFCC_TABLES_MEL_LO_BIT, .mel_out_bits = 16, .log_precision = 11, .log_out_bits = 15,
        .window_curve = mfcc_window_curve, .tw_fft = mfcc_tw_fft, .tw_dct = mfcc_tw_dct,
        .mel_weight = mfcc_mel_weight, .mel_close = mfcc_mel_close, .f_preemph = 0.96875f };
    if (mfcc_b200_create(&cfg, &h)) { printf("mfcc_b200_create: %s\n", mfcc_b200_last_error()); return 2; }
    walk(h, argv[1], &done, &rejected);
    mfcc_b200_destroy(h);
    printf("%d files converted, %d rejected\n", done, rejected);
    return 0;
}
