"""Launch the librosa-set kernels a few times on a small batch (for ncu / timing).

    python profiles/run_librosa.py [clips] [chunk_mb ...]

With chunk sizes given, the rate is measured for each (the log-mel rows kept between the two
passes: small enough and they stay in L2, profiles/r02_*_lr_chunks.txt)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import mfcc_b200
from bench import synth_clips
n = int(sys.argv[1]) if len(sys.argv) > 1 else 300
chunks = [float(a) for a in sys.argv[2:]] or [None]
dev = torch.device("cuda:0")
x = synth_clips(n, 960000, dev)
op = mfcc_b200.LibrosaMFCC(device=dev)
K = op.num_frames(x.shape[1])
out = torch.empty((n, K, 32), dtype=torch.float32, device=dev)
ref = None
for mb in chunks:
    if mb is not None:
        op.config(mb)
    for _ in range(3):
        op(x, out=out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        op(x, out=out)
    e1.record()
    torch.cuda.synchronize()
    same = ""
    if ref is None:
        ref = out.clone()
    else:
        same = " identical to the first: %s" % bool(torch.equal(ref, out))
    print("librosa set%s: %d frames, %.3f ms/launch, %.4g frames/s%s" % (
        "" if mb is None else " (chunk %g MiB)" % mb, n * K, e0.elapsed_time(e1) / 3, n * K / (e0.elapsed_time(e1) / 3e3), same))
