/* TEST INFRASTRUCTURE: a software FT601 for the reference's host tool.
 *
 * The reference's wav2mfcc tool (software/main.c) talks to the FPGA through four calls
 * declared in software/ft601.h:55-58.  This file implements those four calls by
 * forwarding the bytes to callbacks registered from Python (oracle/ref_wav2mfcc.py),
 * where a model of the FPGA top (mfcc/targets/wav2mfcc.py:27-44: bit 31 = reset, sample in
 * bits 15..0, one 32-bit word per cepstral coefficient back) answers them.  With it the
 * UNMODIFIED main.c + libwav compile into oracle/_ref/libwav2mfcc_ref.so and run here.
 */
#include <stddef.h>
#include "ft601.h"

typedef int (*bridge_io_fn)(void *data, int size);
static bridge_io_fn g_write, g_read;

void ft601_bridge_set(bridge_io_fn on_write, bridge_io_fn on_read)
{
    g_write = on_write;
    g_read = on_read;
}

int ft601_open(struct ft601_context *ctx)
{
    (void)ctx;
    return (g_write && g_read) ? FT601_SUCCESS : FT601_ERROR_NOT_FOUND;
}

void ft601_close(struct ft601_context *ctx) { (void)ctx; }

int ft601_write(struct ft601_context *ctx, void *data, int size)
{
    (void)ctx;
    return g_write ? g_write(data, size) : FT601_ERROR_OTHER;
}

int ft601_read(struct ft601_context *ctx, void *data, int size)
{
    (void)ctx;
    return g_read ? g_read(data, size) : FT601_ERROR_OTHER;
}
