import sys, os
sys.path.insert(0, "/root/repo")
import torch, mfcc_b200
dev = torch.device("cuda:0")
hop = torch.randint(-3000, 3000, (4096, 170), dtype=torch.int16, device=dev)
for mode in ("int", "fp32"):
    op = mfcc_b200.MFCC(nfilters=32, nceptrums=32, mode=mode, mel_floor=1e-3 if mode == "fp32" else 0.0, device=dev)
    op.reset(4096)
    for _ in range(5):
        op.step(hop)
    torch.cuda.synchronize()
