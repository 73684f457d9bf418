#!/usr/bin/env python
"""Summarise an .ncu-rep (ncu -i ... --page raw --csv) into the few lines we track.
usage: python profiles/summarize_ncu.py gpurun_out/prof.ncu-rep [frames_per_launch]"""
import csv, subprocess, sys
rep = sys.argv[1]
frames = float(sys.argv[2]) if len(sys.argv) > 2 else None
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = [r for r in csv.reader(raw.splitlines()) if r]
hdr = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
names, units = rows[hdr], rows[hdr + 1]
KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.sum", "sm__inst_executed_pipe_fma.sum", "sm__inst_executed_pipe_lsu.sum",
        "sm__inst_executed_pipe_fmaheavy.sum", "sm__inst_executed_pipe_fmalite.sum",
        "sm__warps_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__cycles_active.avg"]
for v in rows[hdr + 2:]:
    d = dict(zip(names, v))
    print("kernel:", d.get("Kernel Name"))
    for k in KEYS:
        if k in d:
            print("  %-68s %s %s" % (k, d[k], units[names.index(k)]))
    for k in names:
        if k.startswith("smsp__average_warps_issue_stalled") and k.endswith("_per_issue_active.ratio") \
                or k.startswith("smsp__average_warp_latency_issue_stalled") and k.endswith(".ratio"):
            try:
                if float(d[k]) >= 0.05:
                    print("  %-68s %s" % (k, d[k]))
            except ValueError:
                pass
    if frames:
        t = float(d["gpu__time_duration.sum"])
        tu = units[names.index("gpu__time_duration.sum")]
        t_s = t * {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0}.get(tu, 1e-9)
        rd = float(d["dram__bytes_read.sum"]) * {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1}[units[names.index("dram__bytes_read.sum")]]
        wr = float(d["dram__bytes_write.sum"]) * {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1}[units[names.index("dram__bytes_write.sum")]]
        print("  frames/launch %.0f  ->  %.3e frames/s under ncu; dram traffic %.1f B/frame (algorithmic 404);"
              " warp-instructions/frame %.0f" % (frames, frames / t_s, (rd + wr) / frames,
                                                  float(d["smsp__inst_executed.sum"]) / frames))
