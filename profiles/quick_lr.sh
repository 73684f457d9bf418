#!/bin/bash
# librosa set: parity, rate against the chunk size of the log-mel rows, DRAM bytes per frame of all its kernels
python -m pytest tests/test_librosa_mode.py -m gpu -x -q 2>&1 | tail -2
python profiles/run_librosa.py 1200 1000000 48 96 128 192 | tee gpurun_out/lr_chunks.txt
for mb in 1000000 48; do
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none --cache-control none -k regex:lr_ --csv --log-file gpurun_out/lr_dram_$mb.csv python profiles/run_librosa.py 600 $mb > /dev/null 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open("gpurun_out/lr_dram_$mb.csv")) if len(r)>10 and r[0].isdigit()]
tot={}
for r in rows:
    u=r[-2]; v=float(r[-1].replace(",",""))
    if "byte" in u: v*= {"byte":1,"Kbyte":1e3,"Mbyte":1e6,"Gbyte":1e9}[u]
    k=(r[4].split("(")[0][-24:], r[-3])
    tot[k]=tot.get(k,0)+v
frames=600*5648*6   # 6 calls (3 warm-up + 3 timed)
byk={}
for (k,m),v in tot.items():
    if "dram" in m: byk[k]=byk.get(k,0)+v
print("chunk $mb MiB: DRAM bytes/frame by kernel", {k: round(v/frames,1) for k,v in byk.items()}, "total %.1f" % (sum(byk.values())/frames))
PY
done
