#!/usr/bin/env python
"""Executed warp-instructions per source line: joins the SASS page of an ncu report
(ncu -i X.ncu-rep --page source --csv) with nvdisasm's line info of the same build.

    python profiles/dyn_lines.py <source.csv> <kernel-substring> <frames-in-launch> [lib.so]
"""
import collections
import csv
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def sass_with_lines(pat, lib):
    with tempfile.TemporaryDirectory() as td:
        subprocess.check_call(["cuobjdump", "-xelf", "all", lib], cwd=td, stdout=subprocess.DEVNULL)
        cub = [f for f in os.listdir(td) if f.endswith(".cubin")][0]
        dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(td, cub)], capture_output=True,
                             text=True).stdout
    sec, cur, out = None, ("?", 0), []
    for ln in dis.splitlines():
        m = re.match(r"\s*\.section\s+\.text\.(\S+?),", ln)
        if m:
            sec = m.group(1)
            continue
        if sec is None or pat not in sec:
            continue
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", ln)
        if m:
            out.append((cur, m.group(1)))
    return out


def main():
    src, pat, frames = sys.argv[1], sys.argv[2], float(sys.argv[3])
    lib = sys.argv[4] if len(sys.argv) > 4 else os.path.join(ROOT, "mfcc_b200", "libmfcc_b200.so")
    sass = sass_with_lines(pat, lib)
    rows = list(csv.reader(open(src)))
    hdr = rows[1]
    ci = hdr.index("Instructions Executed")
    body = rows[2:]
    assert len(body) == len(sass), (len(body), len(sass))
    per_line = collections.Counter()
    per_op = collections.defaultdict(collections.Counter)
    tot = 0
    for (loc, op), r in zip(sass, body):
        n = int(r[ci])
        assert op.split(".")[0] in r[1], (op, r[1])
        per_line[loc] += n
        per_op[loc][op.split(".")[0]] += n
        tot += n
    print("total %.1f warp-instructions per frame" % (tot / frames))
    thr = float(os.environ.get("MIN", "4"))
    for loc, n in sorted(per_line.items()):
        if n / frames >= thr:
            ops = " ".join("%s:%.0f" % (k, v / frames) for k, v in per_op[loc].most_common(6))
            print("%-18s %4d %7.1f   %s" % (loc[0], loc[1], n / frames, ops))


if __name__ == "__main__":
    main()
