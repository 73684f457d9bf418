"""Measured error of the two float paths against their float64 checkers:
the bundled WAV and C2's class-1 (Gaussian) and class-2 (sinusoids) clips.  GPU box only;
oracle/ is used here as the checker."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import mfcc_b200
from mfcc_b200 import post
from bench import synth_clips
from oracle import ref_float as F, ref_librosa as L, ref_host as H

dev = torch.device("cuda:0")
wav = np.load(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "f2bjrop1.npz"))["samples"]
c2 = synth_clips(4, 960000, dev).cpu().numpy()
clips = {"wav (178240 samples)": wav, "C2 class 1, 10 s": c2[1][:160000], "C2 class 2, 10 s": c2[2][:160000]}
opf = mfcc_b200.MFCC(nfilters=32, nceptrums=32, mode="fp32", device=dev)
opl = mfcc_b200.LibrosaMFCC(device=dev)
print("%-24s %-10s %10s %10s %10s %s" % ("clip", "path", "max|err|", "rms err", "max|ref|", "notes"))
for name, x in clips.items():
    xd = torch.from_numpy(np.ascontiguousarray(x)).to(dev)
    ref, t = F.mfcc(x, taps=True)
    got = opf(xd).cpu().numpy().astype(np.float64)
    ok = (t["mel"] >= 2.0 ** -10).all(axis=1)
    e = np.abs(got[ok] - ref[ok])
    print("%-24s %-10s %10.2e %10.2e %10.1f frames with every mel >= 2^-10: %d of %d; bound 2e-3 + 1e-4|ref|"
          % (name, "fp32", e.max(), np.sqrt((e ** 2).mean()), np.abs(ref[ok]).max(), ok.sum(), len(ok)))
    ref = L.mfcc(x)
    gf = opl(xd)
    got = gf.cpu().numpy().astype(np.float64)
    e = np.abs(got - ref)
    gi = post.to_int16(gf).cpu().numpy()
    si = post.scale(gf).cpu().numpy()
    ri, rs = np.trunc(ref).astype(np.int16), np.trunc(H.scale_time(ref)).astype(np.int16)
    print("%-24s %-10s %10.2e %10.2e %10.1f after astype(int16): %.4f%% equal, max diff %d; after scale+int16: %.4f%% equal, max diff %d"
          % (name, "librosa", e.max(), np.sqrt((e ** 2).mean()), np.abs(ref).max(),
             100 * (gi == ri).mean(), np.abs(gi.astype(int) - ri).max(), 100 * (si == rs).mean(), np.abs(si.astype(int) - rs).max()))
