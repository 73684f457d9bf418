#!/usr/bin/env python
"""Executed warp-instructions per CUDA source line from an .ncu-rep (needs -lineinfo
and --import-source on).  usage: ncu_lines.py rep frames_per_launch [top]"""
import csv, subprocess, sys
rep, frames = sys.argv[1], float(sys.argv[2])
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
cur, ie, agg, src, ops = None, None, {}, {}, {}
last = None
for r in csv.reader(out.splitlines()):
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1].split("/")[-1]; continue
    if r[0] == "Line No":
        ie = r.index("Instructions Executed"); continue
    if r[0] == "Function Name" or ie is None or len(r) <= ie:
        continue
    if r[0].strip() and r[2] == "-":
        try:
            n = float(r[ie])
        except ValueError:
            continue
        last = (cur, int(r[0])); agg[last] = agg.get(last, 0) + n; src[last] = r[1].strip()
    elif r[2].startswith("0x") and last:
        op = r[3].split()[0] if not r[3].strip().startswith("@") else r[3].split()[1]
        op = op.split(".")[0]
        try:
            ops.setdefault(last, {}); ops[last][op] = ops[last].get(op, 0) + float(r[ie])
        except ValueError:
            pass
print("total warp-instr/frame %.1f" % (sum(agg.values()) / frames))
for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:top]:
    o = " ".join("%s:%.0f" % (a, b / frames) for a, b in sorted(ops.get(k, {}).items(), key=lambda t: -t[1])[:6])
    print("%-16s %4d %7.1f  %-70s | %s" % (k[0], k[1], v / frames, src[k][:70], o))
