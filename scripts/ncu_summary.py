#!/usr/bin/env python
"""Summarise an .ncu-rep (raw page) into one line per kernel launch.  Usage: ncu_summary.py file.ncu-rep [more metrics]"""
import csv, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units, data = rows[0], rows[1], rows[2:]
cols = [("Kernel Name", "kernel"), ("gpu__time_duration.sum", "us"), ("dram__bytes_read.sum", "rd"), ("dram__bytes_write.sum", "wr"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"), ("lts__t_sector_hit_rate.pct", "L2hit%"),
        ("l1tex__t_sector_hit_rate.pct", "L1hit%"), ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"), ("launch__registers_per_thread", "regs"),
        ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "stall_lsb"),
        ("smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", "stall_lg")] + [(m, m) for m in sys.argv[2:]]
idx = [(hdr.index(c), n) for c, n in cols if c in hdr]
def fmt(i, v):
    try:
        x = float(v)
    except ValueError:
        return v.split("(")[0][:28]
    u = units[i]
    if u == "Gbyte": return "%.0fMB" % (x * 1000)
    if u == "Mbyte": return "%.0fMB" % x
    if u == "Kbyte": return "%.1fMB" % (x / 1000)
    if u == "byte": return "%.2fMB" % (x / 1e6)
    if u == "ms": return "%.1f" % (x * 1000)
    return "%.1f" % x
print(" | ".join(n for _, n in idx))
for r in data:
    print(" | ".join(fmt(i, r[i]) for i, _ in idx))
