#!/usr/bin/env python
"""Static SASS instruction count per source line of one kernel (nvdisasm -g line info).

    python profiles/sass_lines.py <kernel-substring> [lib.so]

Every frame of the fused kernels is straight-line code (all loops over lanes' registers
are unrolled), so the static count of the lines inside the per-frame device functions is
the number of warp-instructions one frame costs.
"""
import collections
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    pat = sys.argv[1]
    lib = os.path.abspath(sys.argv[2]) if len(sys.argv) > 2 else os.path.join(ROOT, "mfcc_b200", "libmfcc_b200.so")
    with tempfile.TemporaryDirectory() as td:
        subprocess.check_call(["cuobjdump", "-xelf", "all", lib], cwd=td, stdout=subprocess.DEVNULL)
        dis = ""
        for cub in sorted(f for f in os.listdir(td) if f.endswith(".cubin")):
            dis += subprocess.run(["nvdisasm", "-g", "-c", os.path.join(td, cub)], capture_output=True,
                                  text=True).stdout
    sec, cur = None, None
    per_line = collections.Counter()
    per_op = collections.defaultdict(collections.Counter)
    total = 0
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
        if m and cur:
            op = m.group(1).split(".")[0]
            per_line[cur] += 1
            per_op[cur][op] += 1
            total += 1
    print("kernel ~ %s: %d instructions" % (pat, total))
    by_file = collections.Counter()
    for (f, l), n in per_line.items():
        by_file[f] += n
    print(dict(by_file))
    for (f, l), n in sorted(per_line.items()):
        if n >= int(os.environ.get("MIN", "8")):
            ops = " ".join("%s:%d" % kv for kv in per_op[(f, l)].most_common(6))
            print("%-18s %4d  %5d   %s" % (f, l, n, ops))


if __name__ == "__main__":
    main()
