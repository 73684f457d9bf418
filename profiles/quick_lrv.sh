#!/bin/bash
# A/B of library builds on the librosa-set path: parity, rate, ncu counts per frame of lr_logmel
cp mfcc_b200/libmfcc_b200.so /tmp/libmfcc_b200.orig.so
for v in build/variants/*.so; do
  echo "=== $v"
  cp $v mfcc_b200/libmfcc_b200.so
  python -m pytest tests/test_librosa_mode.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -1
  python profiles/run_librosa.py 1200
  python profiles/run_librosa.py 1200
  ncu --metrics smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,gpu__time_duration.sum,launch__registers_per_thread --clock-control none -k regex:lr_logmel -c 3 --csv --log-file gpurun_out/lrv_ncu.csv python profiles/run_librosa.py 300 > /dev/null 2>&1
  python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/lrv_ncu.csv")) if len(r)>10 and r[0].isdigit()]
frames=300*5648
m={}
for r in rows:
    if r[0]=="2": m[r[-3]]=float(r[-1].replace(",",""))
print("   lr_logmel regs %d inst/frame %.0f issue %.1f%% wavefronts/frame %.0f time %s" % (m.get("launch__registers_per_thread",0), m.get("smsp__inst_executed.sum",0)/frames, m.get("smsp__issue_active.avg.pct_of_peak_sustained_active",0), m.get("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",0)/frames, m.get("gpu__time_duration.sum")))
PY
done
cp /tmp/libmfcc_b200.orig.so mfcc_b200/libmfcc_b200.so
