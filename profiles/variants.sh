#!/bin/bash
# Compare library builds on the GPU box: for every build/variants/*.so, install it as the library,
# run the parity tests that cover the fused kernels, time them, and count instructions with ncu.
# usage (via gpurun): bash profiles/variants.sh [pytest-args]
cp mfcc_b200/libmfcc_b200.so /tmp/libmfcc_b200.orig.so
for v in build/variants/*.so; do
  echo "=== $v"
  cp $v mfcc_b200/libmfcc_b200.so
  python -m pytest tests/test_gpu_parity.py tests/test_rtl_golden.py -m gpu -x -q 2>&1 | tail -1
  python profiles/vbench.py 3000 10
  ncu --metrics smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum --clock-control none -k regex:fused -c 8 --csv --log-file gpurun_out/v_ncu.csv python profiles/run_both.py 300 > /dev/null 2>&1
  python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/v_ncu.csv")) if len(r)>10 and r[0].isdigit()]
frames=300*5645
by={}
for r in rows:
    by.setdefault((r[0],r[4][:50]),{})[r[-3]]=float(r[-1].replace(",",""))
for (i,k),m in [list(by.items())[3], list(by.items())[-1]]:
    print("  ", k, "inst/frame %.0f issue %.1f%% smem wavefronts/frame %.0f" % (m.get("smsp__inst_executed.sum",0)/frames, m.get("smsp__issue_active.avg.pct_of_peak_sustained_active",0), m.get("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",0)/frames))
PY
done
cp /tmp/libmfcc_b200.orig.so mfcc_b200/libmfcc_b200.so
