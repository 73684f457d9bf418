"""Launch the librosa-set kernels a few times on a small batch (for ncu / timing)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mfcc_b200
from bench import synth_clips
n = int(sys.argv[1]) if len(sys.argv) > 1 else 300
dev = torch.device("cuda:0")
x = synth_clips(n, 960000, dev)
op = mfcc_b200.LibrosaMFCC(device=dev)
K = op.num_frames(x.shape[1])
out = torch.empty((n, K, 32), dtype=torch.float32, device=dev)
for _ in range(3):
    op(x, out=out)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(3):
    op(x, out=out)
e1.record()
torch.cuda.synchronize()
print("librosa set: %d frames, %.3f ms/launch, %.4g frames/s" % (n * K, e0.elapsed_time(e1) / 3, n * K / (e0.elapsed_time(e1) / 3e3)))
