#!/bin/bash
# fp32 iteration: parity tests of the float path, timing of both fused kernels, ncu counts per frame.
python -m pytest tests/test_float_oracle.py tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -4
python profiles/vbench.py 3000 10
ncu --metrics smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum,smsp__inst_executed_op_shared_ld.sum,smsp__inst_executed_op_shared_st.sum,launch__registers_per_thread --clock-control none -k regex:fused -c 8 --csv --log-file gpurun_out/q_ncu.csv python profiles/run_both.py 300 > /dev/null 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/q_ncu.csv")) if len(r)>10 and r[0].isdigit()]
frames=300*5645
by={}
for r in rows:
    by.setdefault((r[0],r[4][:50]),{})[r[-3]]=float(r[-1].replace(",",""))
for (i,k),m in [list(by.items())[3], list(by.items())[-1]]:
    print("  ", k, "regs %d inst/frame %.0f issue %.1f%% smem wavefronts/frame %.0f (conflicts %.1f) lds %.0f sts %.0f" % (m.get("launch__registers_per_thread",0), m.get("smsp__inst_executed.sum",0)/frames, m.get("smsp__issue_active.avg.pct_of_peak_sustained_active",0), m.get("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",0)/frames, m.get("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",0)/frames, m.get("smsp__inst_executed_op_shared_ld.sum",0)/frames, m.get("smsp__inst_executed_op_shared_st.sum",0)/frames))
PY
