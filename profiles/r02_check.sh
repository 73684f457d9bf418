#!/bin/bash
# First GPU pass of round 2: parity tests, smoke, the bench line with the e2e sweep, the reference arm.
# usage (via gpurun): bash profiles/r02_check.sh <tag>
tag=${1:-r02_a}
o=gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 | tee $o/${tag}_pytest.txt
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -6 | tee $o/${tag}_smoke.txt
python bench.py --e2e-sweep > $o/${tag}_bench_full.json 2> $o/${tag}_bench_full.err || tail -5 $o/${tag}_bench_full.err
python bench.py --impl reference --steps 5 > $o/${tag}_bench_reference.json 2> $o/${tag}_bench_reference.err
python - <<PY
import json
d=json.loads(open("$o/${tag}_bench_full.json").read().strip().splitlines()[-1])
print("int %.4g fps | e2e %.4g (ceiling %.4g, %.3f) | fp32 %.4g | lr %.4g | c4 %s | cpu %s" % (d["value"], d["e2e"]["value"], d["e2e"]["copy_ceiling"]["value"], d["e2e"]["copy_ceiling"]["e2e_over_ceiling"], d.get("fp32",{}).get("value",0), d.get("librosa_set",{}).get("value",0), d.get("c4"), d.get("cpu_baseline")))
print(d["e2e"].get("sweep"))
print(d["e2e"]["copy_ceiling"])
PY
