#!/usr/bin/env python
"""Runs a few device-resident imports of one synthetic workload (no torch): the command ncu wraps.

    python scripts/profile_import.py [--workload highdup16m] [--imports 3]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from bench import WORKLOADS  # noqa: E402
from vulkanengine_b200 import capi, synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="highdup16m", choices=sorted(WORKLOADS))
ap.add_argument("--imports", type=int, default=3)
args = ap.parse_args()

n, perm, _ = WORKLOADS[args.workload]
m = synth.grid(n, permute_seed=perm)
dev = 0
d = [capi.DeviceArray(dev, a) for a in (m.pos, m.nrm, m.uv, m.idx)]
na = max(len(m.pos), len(m.nrm), len(m.uv))
plan = capi.Plan(dev, m.corners, max_attrib=na, max_vertices=na + na // 8 + 1024)
plan.set_timing(True)
for i in range(args.imports):
    plan.import_async(d[0].ptr, len(m.pos), d[1].ptr, len(m.nrm), d[2].ptr, len(m.uv), d[3].ptr, m.shape_offsets)
    r = plan.finish()
print("triangles", m.triangles, "unique", r.total_vertices, "phase_ms",
      dict(zip(plan.phase_names(), [round(x, 4) for x in r.phase_ms])))
