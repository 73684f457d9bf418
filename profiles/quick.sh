#!/bin/bash
# Quick GPU iteration: parity tests, a short bench line, and the dynamic instruction count of the fused kernels.
# usage (on the GPU box, via gpurun): bash profiles/quick.sh <tag>
tag=${1:-q}
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python bench.py --clips 3000 --steps 5 --e2e-clips 64 --no-cpu-baseline --stream-steps 500 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err || tail -5 gpurun_out/${tag}_bench.err
python - <<PY
import json
d=json.loads(open("gpurun_out/${tag}_bench.json").read().strip().splitlines()[-1])
print("int  fps %.4g  ms %.2f | fp32 fps %.4g | e2e %.4g | stream p50 %.1f us" % (d["value"], d["ms_per_step"], d.get("fp32",{}).get("value",0), d["e2e"]["value"], d.get("streaming",{}).get("device_us_p50",0)))
PY
ncu --metrics smsp__inst_executed.sum,gpu__time_duration.sum,sm__inst_executed_pipe_alu.sum,sm__inst_executed_pipe_fma.sum,sm__inst_executed_pipe_lsu.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:fused -c 8 --csv --log-file gpurun_out/${tag}_ncu.csv python profiles/run_both.py 300 > /dev/null 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/${tag}_ncu.csv")) if len(r)>10 and r[0].isdigit()]
frames=300*5645
by={}
for r in rows:
    by.setdefault((r[0],r[4][:60]),{})[r[-3]]=float(r[-1].replace(",",""))
for (i,k),m in [list(by.items())[3], list(by.items())[-1]]:
    print(k, "inst/frame %.0f" % (m.get("smsp__inst_executed.sum",0)/frames), {a.split(".")[0][-24:]:round(v,1) for a,v in m.items() if "pct" in a}, "alu/frame %.0f fma/frame %.0f lsu/frame %.0f" % (m.get("sm__inst_executed_pipe_alu.sum",0)/frames, m.get("sm__inst_executed_pipe_fma.sum",0)/frames, m.get("sm__inst_executed_pipe_lsu.sum",0)/frames))
PY
