#!/usr/bin/env python
"""Turn the raw outputs of profiles/final.sh (gpurun_out/<tag>_*) into the tracked
summaries under profiles/:  python profiles/collect.py <tag>"""
import csv
import json
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
G = os.path.join(ROOT, "gpurun_out")
P = os.path.join(ROOT, "profiles")
FRAMES = 300 * 5645


def raw(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = [r for r in csv.reader(out.splitlines()) if r]
    h = next(i for i, r in enumerate(rows) if r[0] == "ID")
    names, units, vals = rows[h], rows[h + 1], rows[h + 2]
    return dict(zip(names, vals)), dict(zip(names, units))


def to_bytes(v, u):
    return float(v.replace(",", "")) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u]


def main():
    global FRAMES
    tag = sys.argv[1]
    for suffix in ("bench_full.json", "bench_reference.json", "launches.csv", "pipe_bench.jsonl", "pytest.txt",
                   "librosa_rate.txt"):
        src = os.path.join(G, "%s_%s" % (tag, suffix))
        if os.path.exists(src):
            shutil.copy(src, os.path.join(P, "%s_%s" % (tag, suffix)))
    issue = {"source": "profiles/final.sh %s" % tag, "frames_per_profiled_launch": FRAMES}
    for k, name in (("int", "mfcc_int_fused"), ("f32", "mfcc_f32_fused"), ("lr", "lr_logmel")):
        rep = os.path.join(G, "%s_%s.ncu-rep" % (tag, k))
        if not os.path.exists(rep):
            continue
        FRAMES = 300 * (5648 if k == "lr" else 5645)
        with open(os.path.join(P, "%s_%s_ncu_full.txt" % (tag, k)), "w") as f:
            f.write(subprocess.run([sys.executable, os.path.join(P, "summarize_ncu.py"), rep, str(FRAMES)],
                                   capture_output=True, text=True).stdout)
        with open(os.path.join(P, "%s_%s_lines.txt" % (tag, k)), "w") as f:
            f.write(subprocess.run([sys.executable, os.path.join(P, "ncu_lines.py"), rep, str(FRAMES), "60"],
                                   capture_output=True, text=True).stdout)
        d, u = raw(rep)
        issue[name] = {
            "warp_inst_per_frame": float(d["smsp__inst_executed.sum"].replace(",", "")) / FRAMES,
            "dram_bytes_per_frame": (to_bytes(d["dram__bytes_read.sum"], u["dram__bytes_read.sum"])
                                     + to_bytes(d["dram__bytes_write.sum"], u["dram__bytes_write.sum"])) / FRAMES,
            "issue_active_pct": float(d["smsp__issue_active.avg.pct_of_peak_sustained_active"]),
            "lsu_wavefronts_per_frame": float(d.get("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "0").replace(",", "")) / FRAMES,
        }
    pb = os.path.join(G, "%s_pipe_bench.jsonl" % tag)
    if os.path.exists(pb):
        rows = [json.loads(l) for l in open(pb) if l.strip()]
        ops = {r["op"]: r for r in rows if "op" in r}
        issue["pipe_bench"] = {"device": rows[0], "imad_lane_inst_per_s": ops["IMAD (mad.lo.s32)"]["lane_inst_per_s"],
                               "mixed_lane_inst_per_s": ops["IMAD+SHF 1:1"]["lane_inst_per_s"],
                               "butterfly_mix_lane_inst_per_s": ops["IMAD:SHF:IADD 2:1:1"]["lane_inst_per_s"]}
    json.dump(issue, open(os.path.join(P, "r01_issue.json"), "w"), indent=1)
    print(json.dumps(issue, indent=1))


if __name__ == "__main__":
    main()
