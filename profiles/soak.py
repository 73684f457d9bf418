"""Full-size parity soak: N clips of the C2 mixture (60 s each, all four classes), the fused integer
kernel (large CTA shape) against the C checker bit for bit, every frame; then the same clips through
run_host.  Not a test (it takes a while on the CPU side): python profiles/soak.py [clips]"""
import os
import sys
import time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import mfcc_b200
from bench import synth_clips
from oracle import ref_c as RC

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
dev = torch.device("cuda:0")
op = mfcc_b200.MFCC(nfilters=32, nceptrums=32, device=dev)
cfg = RC.RefConfig(nfilters=32, nceptrums=32)
thr = RC.max_threads()
bad = frames = 0
t0 = time.time()
for first in range(0, n, 256):
    m = min(256, n - first)
    x = synth_clips(m, 960000, dev, first_clip=first)
    got = op(x).cpu().numpy()
    xh = x.cpu()
    host = op.run_host(xh.pin_memory()).numpy()
    want = RC.run(cfg, xh.numpy(), threads=thr)
    bad += int((got != want).sum()) + int((host != want).sum())
    frames += want.shape[0] * want.shape[1]
print("soak: %d clips, %d frames, %d coefficients compared twice (device path and run_host), mismatches: %d, %.0f s"
      % (n, frames, frames * 32, bad, time.time() - t0))
sys.exit(1 if bad else 0)
