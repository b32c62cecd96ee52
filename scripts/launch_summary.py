#!/usr/bin/env python
"""Per-kernel summary of an `ncu --metrics gpu__time_duration.sum --csv` launch list."""
import csv, sys, collections
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10 and r[0].isdigit()]
agg = collections.OrderedDict()
for r in rows:
    name = r[4].split("(")[0].replace("void ", "")
    agg.setdefault(name, []).append(float(r[-1]) / 1e3)
tot = sum(sum(v) for v in agg.values())
print("%-34s %6s %10s %10s %7s" % ("kernel", "count", "mean_us", "total_us", "share"))
for k, v in agg.items():
    print("%-34s %6d %10.1f %10.1f %6.1f%%" % (k, len(v), sum(v) / len(v), sum(v), 100 * sum(v) / tot))
print("launches %d, total %.1f us (cold-cache, serialised: compare shares, not absolutes)" % (len(rows), tot))
