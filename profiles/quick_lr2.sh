#!/bin/bash
# librosa set: parity, then rate and per-kernel time with the DCT on the tensor cores and on the FMA pipes
python -m pytest tests/test_librosa_mode.py -m gpu -x -q 2>&1 | tail -2
python - <<PY
import sys, os
sys.path.insert(0, os.getcwd())
import torch, mfcc_b200
from bench import synth_clips
dev = torch.device("cuda:0")
x = synth_clips(1200, 960000, dev)
op = mfcc_b200.LibrosaMFCC(device=dev)
K = op.num_frames(x.shape[1])
outs = {}
for tc in (True, False, True):
    op.dct_path(tc)
    out = torch.empty((1200, K, 32), dtype=torch.float32, device=dev)
    for _ in range(2): op(x, out=out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3): op(x, out=out)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    outs[tc] = out
    print("dct on %s: %.3f ms, %.4g frames/s" % ("tensor cores" if tc else "FMA pipes", ms, 1200 * K / ms * 1e3))
d = (outs[True] - outs[False]).abs()
print("max |tensor - fma| %.3g (values up to %.1f)" % (d.max().item(), outs[False].abs().max().item()))
PY
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,sm__inst_executed_pipe_tensor.sum --clock-control none -k regex:lr_dct -c 4 --csv --log-file gpurun_out/lr_dct.csv python profiles/run_librosa.py 300 > /dev/null 2>&1
python - <<PY
import csv
for r in csv.reader(open("gpurun_out/lr_dct.csv")):
    if len(r) > 10 and r[0].isdigit() and r[0] in ("0","3"): print(r[4][:40], r[-3], r[-2], r[-1])
PY
