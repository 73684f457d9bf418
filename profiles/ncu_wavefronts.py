#!/usr/bin/env python
"""Shared-memory wavefronts per CUDA source line from an .ncu-rep (needs -lineinfo and
--import-source on): where a shared-memory-bound kernel spends its one-wavefront-per-clock
budget.  usage: ncu_wavefronts.py rep frames_per_launch [top]"""
import csv, subprocess, sys
rep, frames = sys.argv[1], float(sys.argv[2])
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
cur, cols, agg, ideal, src, last = None, None, {}, {}, {}, None
for r in csv.reader(out.splitlines()):
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1].split("/")[-1]; continue
    if r[0] == "Line No":
        cols = {n: i for i, n in enumerate(r)}; continue
    if cols is None or len(r) <= cols["L1 Wavefronts Shared"]:
        continue
    if r[0].strip() and r[2] == "-":
        last = (cur, int(r[0])); src[last] = r[1].strip(); continue
    if r[2].startswith("0x") and last:
        try:
            w = float(r[cols["L1 Wavefronts Shared"]] or 0); wi = float(r[cols["L1 Wavefronts Shared Ideal"]] or 0)
        except ValueError:
            continue
        agg[last] = agg.get(last, 0) + w; ideal[last] = ideal.get(last, 0) + wi
tot = sum(agg.values())
print("shared wavefronts/frame %.1f (ideal %.1f)" % (tot / frames, sum(ideal.values()) / frames))
for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:top]:
    if v <= 0:
        break
    print("%-18s %5d %8.1f (ideal %6.1f)  %s" % (k[0], k[1], v / frames, ideal[k] / frames, src[k][:90]))
