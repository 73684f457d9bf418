import sys, os
sys.path.insert(0, os.getcwd())
import torch, mfcc_b200
dev = torch.device("cuda:0")
op = mfcc_b200.MFCC(nfilters=32, nceptrums=32, device=dev)
S = 4096
hop = torch.randint(-3000, 3000, (S, 170), dtype=torch.int16, device=dev)
out = torch.empty((S, 32), dtype=torch.int16, device=dev)
op.reset(S)
for _ in range(20): op.step(hop, out=out)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(200): op.step(hop, out=out)
e1.record(); torch.cuda.synchronize()
print("step kernel back to back: %.2f us per launch" % (e0.elapsed_time(e1) * 1e3 / 200))
