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
                   "librosa_rate.txt", "accuracy.txt", "smoke.txt", "stage_rates.txt", "mma_bench.jsonl"):
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
        with open(os.path.join(P, "%s_%s_wavefronts.txt" % (tag, k)), "w") as f:
            f.write(subprocess.run([sys.executable, os.path.join(P, "ncu_wavefronts.py"), rep, str(FRAMES), "30"],
                                   capture_output=True, text=True).stdout)
        d, u = raw(rep)
        issue[name] = {
            "warp_inst_per_frame": float(d["smsp__inst_executed.sum"].replace(",", "")) / FRAMES,
            "dram_bytes_per_frame": (to_bytes(d["dram__bytes_read.sum"], u["dram__bytes_read.sum"])
                                     + to_bytes(d["dram__bytes_write.sum"], u["dram__bytes_write.sum"])) / FRAMES,
            "issue_active_pct": float(d["smsp__issue_active.avg.pct_of_peak_sustained_active"]),
            "lsu_wavefronts_per_frame": float(d.get("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "0").replace(",", "")) / FRAMES,
        }
    ll = os.path.join(G, "%s_launches.csv" % tag)
    if os.path.exists(ll):
        # per-kernel totals of the launch list (bench.py --steps 2 --warmup 3 under ncu: cold-cache,
        # serialised, so shares of the device time, not absolutes)
        agg, full = {}, []
        for r in csv.reader(open(ll)):
            if len(r) > 14 and r[0].isdigit() and r[12] == "gpu__time_duration.sum":
                name = r[4].replace("void ", "").replace("<unnamed>::", "").split("(")[0]
                a = agg.setdefault(name, [0, 0.0])
                a[0] += 1
                if "mfcc_int_fused" in name and r[8].replace(" ", "").startswith(("(592,", "(296,")):
                    full.append(float(r[14].replace(",", "")) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0}.get(r[13], 1e-6))
                a[1] += float(r[14].replace(",", "")) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0}.get(r[13], 1e-6)
        tot = sum(v[1] for v in agg.values())
        with open(os.path.join(P, "%s_launch_shares.txt" % tag), "w") as f:
            f.write("# %s: kernels of this library launched by `bench.py --steps 2 --warmup 3 --e2e-clips 256 "
                    "--stream-steps 20` under ncu (first 400 launches)\n" % tag)
            f.write("%-40s %8s %12s %8s\n" % ("kernel", "launches", "total ms", "share"))
            for name, (n, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
                f.write("%-40s %8d %12.3f %7.1f%%\n" % (name, n, ms, 100 * ms / tot))
            big = sorted(v for v in full if v > 20.0)
            f.write("# the timed step of bench.py is ONE launch of mfcc_int_fused<32, 0, ...> over 10 000 clips (100 %% of the "
                    "step);\n# its full-grid launches here (10 000 clips: warm-up + timed steps; 16 667 clips: the C4 leg) took "
                    "%s ms\n" % ", ".join("%.2f" % v for v in big))
    pb = os.path.join(G, "%s_pipe_bench.jsonl" % tag)
    if os.path.exists(pb):
        rows = [json.loads(l) for l in open(pb) if l.strip()]
        ops = {r["op"]: r for r in rows if "op" in r}
        issue["pipe_bench"] = {"device": rows[0], "imad_lane_inst_per_s": ops["IMAD (mad.lo.s32)"]["lane_inst_per_s"],
                               "mixed_lane_inst_per_s": ops["IMAD+SHF 1:1"]["lane_inst_per_s"],
                               "butterfly_mix_lane_inst_per_s": ops["IMAD:SHF:IADD 2:1:1"]["lane_inst_per_s"]}
    json.dump(issue, open(os.path.join(P, "%s_issue.json" % tag.split("_")[0]), "w"), indent=1)
    print(json.dumps(issue, indent=1))


if __name__ == "__main__":
    main()
