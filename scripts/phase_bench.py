#!/usr/bin/env python
"""Phase times of device-resident imports under several settings of the library's environment knobs, one mesh
upload for all of them (no torch).  A measurement helper for kernel work:

    python scripts/phase_bench.py --workload highdup16m --steps 30 --cfg VKM_RECORDS=0 --cfg VKM_RECORDS=1,VKM_M2_PCT=200

Every --cfg is a comma-separated list of NAME=VALUE pairs set before the plan is created (the plan reads them
once).  Prints one JSON line per configuration: phase times (ms, mean over the timed imports) and their sum.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from bench import WORKLOADS  # noqa: E402
from vulkanengine_b200 import capi, synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="highdup16m", choices=sorted(WORKLOADS))
ap.add_argument("--steps", type=int, default=30)
ap.add_argument("--warmup", type=int, default=3)
ap.add_argument("--cfg", action="append", default=[])
ap.add_argument("--dirty", action="store_true", help="one duplicated position record: the mesh needs the general schedule")
args = ap.parse_args()

n, perm, _ = WORKLOADS[args.workload]
if n is None:
    from vulkanengine_b200 import sharding

    ns, tris = sharding.batch_sizes(10000)
    m = sharding.build_shard(ns, sharding.lpt_partition(tris, 1)[0])
elif args.workload == "sphere1m":
    m = synth.uv_sphere(n, n)
elif args.workload == "adversarial64m":
    m = synth.adversarial(n, frac=2e-4, nan_frac=2e-5)
else:
    m = synth.grid(n, permute_seed=perm)
if args.dirty:
    m = synth.with_duplicate_record(m)
dev = 0
d = [capi.DeviceArray(dev, a) for a in (m.pos, m.nrm, m.uv, m.idx)]
na = max(len(m.pos), len(m.nrm), len(m.uv))
nshapes = len(m.shape_offsets) - 1
for cfg in args.cfg or [""]:
    pairs = [kv.split("=", 1) for kv in cfg.split(",") if kv]
    for k, v in pairs:
        os.environ[k] = v
    plan = capi.Plan(dev, m.corners, max_attrib=na, max_vertices=na + na // 8 + 1024, max_shapes=nshapes)
    first_stats = None
    for _ in range(args.warmup):
        plan.import_async(d[0].ptr, len(m.pos), d[1].ptr, len(m.nrm), d[2].ptr, len(m.uv), d[3].ptr, m.shape_offsets)
        r = plan.finish()
        first_stats = first_stats or plan.stats()
    plan.set_timing(True)
    for _ in range(args.steps):
        plan.import_async(d[0].ptr, len(m.pos), d[1].ptr, len(m.nrm), d[2].ptr, len(m.uv), d[3].ptr, m.shape_offsets)
    r = plan.finish()
    ph = dict(zip(plan.phase_names(), [round(float(x), 4) for x in r.phase_ms]))
    total = sum(ph.values())
    b_alg = m.algorithmic_bytes(int(r.total_vertices))
    print(json.dumps({"cfg": cfg, "workload": args.workload, "triangles": m.triangles, "unique": int(r.total_vertices),
                      "phase_ms": ph, "sum_ms": round(total, 4), "gtris_s": round(m.triangles / total / 1e6, 3),
                      "b_alg_gbs": round(b_alg / total / 1e6, 1), "stats": plan.stats(), "first_import_stats": first_stats}), flush=True)
    del plan
    for k, _ in pairs:
        os.environ.pop(k, None)
