"""Launch the fused integer and fp32 kernels a few times on a small batch (for ncu)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mfcc_b200
from bench import synth_clips
n = int(sys.argv[1]) if len(sys.argv) > 1 else 300
dev = torch.device("cuda:0")
x = synth_clips(n, 960000, dev)
for mode in ("int", "fp32"):
    op = mfcc_b200.MFCC(nfilters=32, nceptrums=32, mode=mode, mel_floor=1e-3 if mode == "fp32" else 0.0, device=dev)
    for _ in range(4):
        y = op(x)
    torch.cuda.synchronize()
