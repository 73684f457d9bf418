#!/bin/bash
# Multi-GPU pass: bench.py at N = 2, 4, 8 (with the run_host chunk/slot sweep and the copy-only ceiling),
# plus the write-combined variant at 8.  usage (via gpurun --gpus 8): bash profiles/r02_scale.sh <tag>
tag=${1:-r02_a}
o=gpurun_out
nvidia-smi topo -m > $o/${tag}_topo.txt 2>&1
lscpu | head -30 > $o/${tag}_lscpu.txt 2>&1
numactl -H >> $o/${tag}_lscpu.txt 2>&1
free -g >> $o/${tag}_lscpu.txt 2>&1
for n in 8 4 2; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500+n)) \
      bench.py --gpus $n --steps 10 --e2e-sweep > $o/${tag}_bench_n$n.json 2> $o/${tag}_bench_n$n.err || tail -5 $o/${tag}_bench_n$n.err
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29611 \
    bench.py --gpus 8 --steps 10 --e2e-wc --no-c4 > $o/${tag}_bench_n8_wc.json 2> $o/${tag}_bench_n8_wc.err
python - <<PY
import json
for n in ("8","4","2","8_wc"):
    try:
        d=json.loads(open("$o/${tag}_bench_n%s.json"%n).read().strip().splitlines()[-1])
    except Exception as e:
        print(n,"failed",e); continue
    cc=d["e2e"]["copy_ceiling"]
    print("N=%s value %.4g e2e %.4g ceiling %.4g ratio %.3f h2d %.1f GB/s (alone %.1f) d2h alone %.1f c4 %s" % (n, d["value"], d["e2e"]["value"], cc["value"], cc["e2e_over_ceiling"], cc["h2d_gbs_all_gpus"], cc["h2d_alone_gbs_all_gpus"], cc["d2h_alone_gbs_all_gpus"], (d.get("c4") or {}).get("ms_per_pass")))
    print("   sweep", [(s["chunk_mb"], s["slots"], "%.3g"%s["value"]) for s in d["e2e"].get("sweep",[])])
PY
