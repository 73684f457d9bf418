"""Time the fused integer and fp32 kernels on C2-shaped clips (for comparing library builds).

    python profiles/vbench.py [clips] [steps]
"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mfcc_b200
from bench import synth_clips

n = int(sys.argv[1]) if len(sys.argv) > 1 else 3000
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
dev = torch.device("cuda:0")
x = synth_clips(n, 960000, dev)
frames = n * 5645
out = {}
for mode in ("int", "fp32"):
    op = mfcc_b200.MFCC(nfilters=32, nceptrums=32, mode=mode, mel_floor=1e-3 if mode == "fp32" else 0.0, device=dev)
    for _ in range(3):
        y = op(x)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        y = op(x)
    e1.record()
    torch.cuda.synchronize()
    out[mode] = frames * steps / (e0.elapsed_time(e1) * 1e-3)
print("int %.4g frames/s | fp32 %.4g frames/s" % (out["int"], out["fp32"]))
