"""Drive every kernel of libmfcc_b200.so once on small inputs, with the cheap internal
cross-checks (aligned vs unaligned rows, device vs host path, stage chain vs fused kernel).
Written as the driver for compute-sanitizer (memcheck / racecheck / synccheck); that tool is
closed on this GPU pool, so the script runs plain, as tests/test_gpu_parity.py::test_all_kernels_driver.
"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import mfcc_b200
from mfcc_b200 import post

dev = torch.device("cuda:0")
g = torch.Generator(device="cpu").manual_seed(5)
n = 512 + 170 * 53 + 31                                   # three tiles of 24 frames, ragged tail
x = torch.randint(-32768, 32768, (5, n), dtype=torch.int16, generator=g)
xd = x.to(dev)
xu = torch.zeros((3, n + 3), dtype=torch.int16, device=dev)[:, 1:n + 1]     # unaligned rows: plain-load path
xu.copy_(xd[:3])
for nf, nc in ((32, 32), (16, 16)):
    op = mfcc_b200.MFCC(nfilters=nf, nceptrums=nc, device=dev)
    y = op(xd)
    yu = op(xu)
    assert torch.equal(y[:3], yu)
    y1, taps = op(xd[0], taps=True)
    assert torch.equal(y1, y[0])
    yh = op.run_host(x)
    assert torch.equal(yh.to(dev), y)
    op.reset(7)
    for k in range(3):
        op.step(xd[:, 170 * k:170 * (k + 1)].repeat(2, 1)[:7].contiguous())
    # stage operators
    pre = op.preemph(xd[0])
    fr = op.frame(pre)
    w = op.window(fr)
    f = op.fft_stream(w)
    p = op.powspec(f)
    m = op.filterbank(p)
    l = op.log2(m)
    d = op.dct_stream(l)
    assert torch.equal(op.discard(d), y1)
    opf = mfcc_b200.MFCC(nfilters=nf, nceptrums=nc, mode="fp32", mel_floor=1e-3, device=dev)
    z = opf(xd)
    zu = opf(xu)
    assert torch.equal(z[:3], zu)
    opf.reset(7)
    opf.step(xd[:, :170].repeat(2, 1)[:7].contiguous())
    assert torch.isfinite(z).all()
y = mfcc_b200.MFCC(nfilters=32, nceptrums=32, device=dev)(xd)
post.lifter(y)
post.scale(y)
post.uart_frame(y[0, :, :16].contiguous())
gate = post.VoiceGate(5, ncepstrums=32)
for k in range(4):
    gate.step(y[:, k].contiguous())
for kw in (dict(), dict(n_mels=40, n_mfcc=13, hop_length=160), dict(center=False)):
    lr = mfcc_b200.LibrosaMFCC(device=dev, **kw)
    for m in (n, 700, 3):
        if kw.get("center", True) or m >= 2048:
            f = lr(xd[:, :m].contiguous())
            assert torch.isfinite(f).all()
    f1, t = lr(xd[0], taps=True)
    post.to_int16(f1)
    post.scale(f1)
# general kernels of the stage classes at sizes other than MFCC's
r = torch.randint(-32768, 32768, (9, 4096), dtype=torch.int16, generator=g).to(dev)
for nfft in (32, 256, 4096):
    mfcc_b200.FftStream(16, nfft, device=dev)(r[:, :nfft].contiguous())
    mfcc_b200.WindowHamming(16, nfft, 7, device=dev)(r[:, :nfft].contiguous())
for nin in (8, 64, 1024):
    mfcc_b200.DCTStream(16, nin, device=dev)(r[:, :nin].contiguous())
mfcc_b200.Frame(16, 400, 160, 512, device=dev)(r[0])
mfcc_b200.PowerSpectrum(device=dev)(r[:, :512].reshape(9, 256, 2))
mfcc_b200.Log2Fix(37, 20, device=dev)(r[:, :64].to(torch.int64).abs())
mfcc_b200.FilterBank(device=dev)(r[:, :256].to(torch.int32).abs())
torch.cuda.synchronize()
print("all_kernels.py: all kernels ran")
