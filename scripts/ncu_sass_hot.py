#!/usr/bin/env python
"""Hottest SASS instructions of one kernel in an .ncu-rep (source page): stall samples, executions, L2 sectors.
Usage: ncu_sass_hot.py file.ncu-rep kernel-regex [top N]"""
import csv, subprocess, sys
rep, pat = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + pat], capture_output=True, text=True).stdout
lines = raw.splitlines()
# the first launch only
start = [i for i, l in enumerate(lines) if l.startswith('"Kernel Name"')]
seg = lines[start[0] + 1:(start[1] if len(start) > 1 else len(lines))]
rows = list(csv.reader(seg))
hdr, data = rows[0], rows[1:]
ci = {n: hdr.index(n) for n in ("Source", "Warp Stall Sampling (All Samples)", "Instructions Executed", "L2 Theoretical Sectors Global")}
tot = sum(int(r[ci["Warp Stall Sampling (All Samples)"]] or 0) for r in data)
print(lines[start[0]][:120], "total samples", tot)
order = sorted(range(len(data)), key=lambda i: -int(data[i][ci["Warp Stall Sampling (All Samples)"]] or 0))[:top]
for i in sorted(order):
    r = data[i]
    print("%5d %5.1f%% exec %9s l2sect %10s  %s" % (i, 100.0 * int(r[ci["Warp Stall Sampling (All Samples)"]] or 0) / max(tot, 1),
          r[ci["Instructions Executed"]], r[ci["L2 Theoretical Sectors Global"]], r[ci["Source"]].strip()[:110]))
