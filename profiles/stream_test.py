import sys, json
sys.path.insert(0, "/root/repo")
import torch, mfcc_b200
from bench import stream_latency
dev = torch.device("cuda:0")
torch.cuda.set_device(dev)
ops = {m: mfcc_b200.MFCC(nfilters=32, nceptrums=32, mode=m, mel_floor=1e-3 if m == "fp32" else 0.0, device=dev) for m in ("int", "fp32")}
for order in (("fp32", "int"), ("int", "fp32")):
    for m in order:
        r = stream_latency(ops[m], dev, 4096, 1000)
        print(m, {k: round(v, 1) for k, v in r.items() if k.endswith("p50") or k.endswith("p99")})
