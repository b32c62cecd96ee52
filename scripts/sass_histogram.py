#!/usr/bin/env python
"""Opcode histogram of selected kernels in the built library (cuobjdump -sass): what the memory path really is.
Usage: sass_histogram.py [kernel-substring ...]"""
import collections, os, re, subprocess, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vulkanengine_b200 import build
pats = sys.argv[1:] or ["k_insert_recordsILi0ELb0", "k_write_vertices_recILi0ELb0", "k_write_indices_recILb0"]
sass = subprocess.run(["cuobjdump", "-sass", build.LIB_PATH], capture_output=True, text=True).stdout
fn, hist = None, collections.defaultdict(collections.Counter)
for line in sass.splitlines():
    if "Function :" in line:
        fn = line.split("Function :")[1].strip()
        continue
    m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and fn and any(p in fn for p in pats):
        hist[fn][m.group(1)] += 1
mem = ("LDG", "STG", "ATOMG", "REDG", "LDS", "STS", "LDL", "STL", "LDGSTS", "UBLKCP", "UTMALDG", "UTMASTG", "CCTL", "MATCH", "SHFL", "REDUX", "VOTE")
for f, h in hist.items():
    print("==", f, "instructions", sum(h.values()))
    print("   memory / warp-level:", {k: v for k, v in sorted(h.items()) if k.split(".")[0] in mem})
    top = collections.Counter()
    for k, v in h.items():
        top[k.split(".")[0]] += v
    print("   by base opcode:", dict(top.most_common(14)))
tma = [f for f, h in hist.items() if any(k.startswith(("UBLKCP", "UTMALDG", "UTMASTG")) for k in h)]
print("bulk-copy / TMA instructions (UBLKCP, UTMALDG, UTMASTG):", tma or "none -- these are gather / scatter kernels; a bulk-copy ring on the index stream measured slower (see vkm_records.cuh)")
