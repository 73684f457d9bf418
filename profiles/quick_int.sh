#!/bin/bash
# A/B of library builds (build/variants/*.so): parity, timing, ncu counts per frame of both fused kernels
cp mfcc_b200/libmfcc_b200.so /tmp/libmfcc_b200.orig.so
for v in build/variants/*.so; do
  echo "=== $v"
  cp $v mfcc_b200/libmfcc_b200.so
  python -m pytest tests/test_gpu_parity.py tests/test_rtl_golden.py tests/test_hypothesis.py tests/test_float_oracle.py -m gpu -x -q 2>&1 | tail -1
  python profiles/vbench.py 3000 10
  python profiles/vbench.py 3000 10
  ncu --metrics smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,sm__inst_executed_pipe_alu.sum,sm__inst_executed_pipe_fmaheavy.sum,sm__inst_executed_pipe_fma.sum,sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active,launch__registers_per_thread --clock-control none -k regex:fused -c 8 --csv --log-file gpurun_out/v_ncu.csv python profiles/run_both.py 300 > /dev/null 2>&1
  python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/v_ncu.csv")) if len(r)>10 and r[0].isdigit()]
frames=300*5645
by={}
for r in rows:
    by.setdefault((r[0],r[4][:50]),{})[r[-3]]=float(r[-1].replace(",",""))
for (i,k),m in [list(by.items())[3], list(by.items())[-1]]:
    g=lambda n: m.get(n,0)
    print("  ", k[22:], "regs %d inst/frame %.0f (alu %.0f fmaheavy %.0f fma %.0f) issue %.1f%% alu %.1f%% fmaheavy %.1f%% wavefronts/frame %.0f" % (g("launch__registers_per_thread"), g("smsp__inst_executed.sum")/frames, g("sm__inst_executed_pipe_alu.sum")/frames, g("sm__inst_executed_pipe_fmaheavy.sum")/frames, g("sm__inst_executed_pipe_fma.sum")/frames, g("smsp__issue_active.avg.pct_of_peak_sustained_active"), g("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active"), g("sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active"), g("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum")/frames))
PY
done
cp /tmp/libmfcc_b200.orig.so mfcc_b200/libmfcc_b200.so
