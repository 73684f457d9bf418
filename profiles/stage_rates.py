"""Throughput of the per-stage kernels behind the operator API (integer mode), one long stream:
bytes read + written per launch over the launch time, against the measured HBM peak.

    python profiles/stage_rates.py [million_samples]
"""
import json
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mfcc_b200

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
n = int(float(sys.argv[1]) * 1e6) if len(sys.argv) > 1 else 32_000_000
dev = torch.device("cuda:0")
try:
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    peak = None
g = torch.Generator(device=dev).manual_seed(3)
x = (torch.randn(n, device=dev, generator=g) * 2000).clamp(-32768, 32767).to(torch.int16)
op = mfcc_b200.MFCC(nfilters=32, nceptrums=32, device=dev)


def nbytes(t):
    return t.numel() * t.element_size()


def timed(fn, arg, reps=5):
    out = fn(arg)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        out = fn(arg)
    e1.record()
    torch.cuda.synchronize()
    return out, e0.elapsed_time(e1) / reps


print("%-12s %10s %10s %10s %8s" % ("stage", "ms", "GB moved", "GB/s", "of peak"))
cur = x
for name, fn in (("preemph", op.preemph), ("frame", op.frame), ("window", op.window), ("fft_stream", op.fft_stream),
                 ("powspec", op.powspec), ("filterbank", op.filterbank), ("log2", op.log2), ("dct_stream", op.dct_stream)):
    out, ms = timed(fn, cur)
    gb = (nbytes(cur) + nbytes(out)) / 1e9
    print("%-12s %10.3f %10.3f %10.1f %8s" % (name, ms, gb, gb / ms * 1e3, "%.1f%%" % (100 * gb / ms * 1e3 / peak) if peak else "-"))
    cur = out
K = cur.shape[0]
_, ms = timed(op, x)
print("fused        %10.3f  (%d frames, %.4g frames/s; the whole chain in one launch)" % (ms, K, K / ms * 1e3))
# context for the write-heavy stages: what this GPU does for a pure write, a 1:1 copy and a 1:3 expansion
# (Frame reads every sample once from HBM and writes it three times)
buf = torch.empty(1_536_000_000 // 2, dtype=torch.int16, device=dev)
src = torch.empty(512_000_000 // 2, dtype=torch.int16, device=dev)
_, ms = timed(lambda t: t.zero_(), buf)
print("memset       %10.3f %10.3f %10.1f   (pure write)" % (ms, nbytes(buf) / 1e9, nbytes(buf) / 1e9 / ms * 1e3))
_, ms = timed(lambda t: buf[:t.numel()].copy_(t), src)
print("copy 1:1     %10.3f %10.3f %10.1f   (read + write)" % (ms, 2 * nbytes(src) / 1e9, 2 * nbytes(src) / 1e9 / ms * 1e3))
_, ms = timed(lambda t: buf.view(3, -1).copy_(t.expand(3, -1)), src)
print("expand 1:3   %10.3f %10.3f %10.1f   (one read, three writes: Frame's traffic shape)" % (
    ms, (nbytes(src) + nbytes(buf)) / 1e9, (nbytes(src) + nbytes(buf)) / 1e9 / ms * 1e3))
