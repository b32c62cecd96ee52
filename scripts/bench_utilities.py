#!/usr/bin/env python
"""Stand-alone geometry utilities (SURVEY.md 8(a) rows A5/A6, 8(d) "stand-alone utilities"): device-resident
vkm_generate_tangents_dev / vkm_generate_normals_dev on the finished buffers of a synthetic mesh, CUDA-event timed,
against the HBM roofline with the survey's useful-byte counts (tangents 12F + 36U, normals 12F + 24U) and the
sector-granular floor beside them; the reference's own Geometry::GenerateTangents / GenerateNormals (oracle/_ref)
timed on a bounded sample of the same family.  Prints one JSON line.

    python scripts/bench_utilities.py [--workload highdup16m] [--steps 50]
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402
import torch  # noqa: E402

from bench import WORKLOADS, hbm_peak  # noqa: E402
from oracle import port, refcpu  # noqa: E402
from vulkanengine_b200 import capi, synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="highdup16m", choices=sorted(WORKLOADS))
ap.add_argument("--steps", type=int, default=50)
ap.add_argument("--warmup", type=int, default=5)
ap.add_argument("--cpu-grid", type=int, default=708, help="grid size of the CPU reference sample (708 = 1.0 M triangles)")
args = ap.parse_args()

dev = 0
torch.cuda.set_device(dev)
n, perm, desc = WORKLOADS[args.workload]
m = synth.grid(n, permute_seed=perm)
d = [capi.DeviceArray(dev, a) for a in (m.pos, m.nrm, m.uv, m.idx)]
na = max(len(m.pos), len(m.nrm), len(m.uv))
plan = capi.Plan(dev, m.corners, max_attrib=na, max_vertices=na + na // 8 + 1024)
plan.import_async(d[0].ptr, len(m.pos), d[1].ptr, len(m.nrm), d[2].ptr, len(m.uv), d[3].ptr, m.shape_offsets)
r = plan.finish()
U, C = int(r.total_vertices), int(r.total_indices)
F = C // 3
scratch = capi.DeviceArray(dev, nbytes=4 * U)
L = capi.lib()
stream = torch.cuda.Stream(device=dev)
peak, peak_src = hbm_peak()


def timed(fn):
    for _ in range(args.warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        e0.record(stream)
        for _ in range(args.steps):
            fn()
        e1.record(stream)
    e1.synchronize()
    return e0.elapsed_time(e1) / args.steps


def call(f):
    def run():
        rc = f(U, r.d_vertices, C, r.d_indices, scratch.ptr, stream.cuda_stream)
        if rc:
            raise SystemExit(L.vkm_last_error().decode())
    return run


out = {"workload": args.workload, "description": desc, "triangles": F, "vertices": U, "steps": args.steps,
       "peak_GBps": peak, "peak_source": peak_src,
       "l2": "vertex + index buffers %.0f MB exceed the 126 MB L2" % ((64 * U + 4 * C) / 1e6) if 64 * U + 4 * C > 200e6
             else "working set below L2 size: L2-warm"}
for name, f, useful, floor in (
        ("generate_tangents", L.vkm_generate_tangents_dev, 12 * F + 36 * U, 12 * F + 64 * U + 32 * U),
        ("generate_normals", L.vkm_generate_normals_dev, 12 * F + 24 * U, 12 * F + 32 * U + 32 * U)):
    ms = timed(call(f))
    out[name] = {"ms": ms, "triangles_per_s": F / (ms * 1e-3), "useful_bytes": useful,
                 "achieved_GBps": useful / (ms * 1e-3) / 1e9, "frac": useful / (ms * 1e-3) / 1e9 / peak,
                 "sector_floor_bytes": floor, "frac_sector_floor": floor / (ms * 1e-3) / 1e9 / peak,
                 "launches_per_call": "1 memset + 2 kernels"}

# ---- the same utilities through the HOST entry points (what the engine's Geometry:: statics call): caller arrays in
# pageable memory, upload + kernels + download of the changed column inside the timed region
shapes_h = plan.fetch(r, m.shape_offsets)
hv, hi = shapes_h[0].vertices.copy(), np.ascontiguousarray(shapes_h[0].indices)
capi.init([dev])


def host_timed(fn, reps=3):
    fn()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    return (time.perf_counter() - t0) / reps * 1e3


for name, f in (("generate_tangents", L.vkm_generate_tangents), ("generate_normals", L.vkm_generate_normals)):
    ms = host_timed(lambda: capi._check(f(len(hv), hv.ctypes.data, len(hi), hi.ctypes.data)))
    out[name]["host_e2e"] = {"ms": ms, "triangles_per_s": F / (ms * 1e-3), "h2d_bytes": int(hv.nbytes + hi.nbytes),
                             "d2h_bytes": int(len(hv) * (16 if name == "generate_tangents" else 12)),
                             "api": "vkm_" + name + " (pageable caller arrays; only the changed column is downloaded)"}
# Geometry::DeduplicateVertices: the first 2^20 corners of the mesh expanded to one vertex per corner (about 6 copies
# of every vertex), indices 0 .. n-1
nd = min(1 << 20, C)
dv = np.ascontiguousarray(hv[hi[:nd]])
di = np.arange(nd, dtype=np.uint32)


def dedup_once():
    ix = di.copy()
    outp = C_.c_void_p()
    n_u = L.vkm_deduplicate_vertices(nd, dv.ctypes.data, nd, ix.ctypes.data, C_.byref(outp))
    assert n_u != 0xFFFFFFFF, L.vkm_last_error()
    C_.CDLL(None).free(outp)
    return n_u


import ctypes as C_  # noqa: E402

n_unique = dedup_once()
ms = host_timed(dedup_once)
out["deduplicate_vertices"] = {"ms": ms, "vertices_in": nd, "vertices_out": int(n_unique), "vertices_per_s": nd / (ms * 1e-3),
                               "api": "vkm_deduplicate_vertices (host arrays in and out; cell hash + one-warp recurrence + remap)"}
nsmall = 20000  # the reference's loop is O(V * U): a bounded sample
if refcpu.available():
    sv, si = np.ascontiguousarray(dv[:nsmall]), np.arange(nsmall, dtype=np.uint32)
    t0 = time.perf_counter()
    refcpu.deduplicate_vertices(sv, si)
    out["deduplicate_vertices"]["cpu_reference"] = {"vertices_in": nsmall, "seconds": time.perf_counter() - t0,
                                                    "note": "Geometry::DeduplicateVertices is quadratic; 1 thread"}

# the reference's own functions on this host (1 thread: they are plain loops), bounded sample
mc = synth.grid(args.cpu_grid)
shapes = port.import_arrays(mc.pos, mc.nrm, mc.uv, mc.idx, mc.shape_offsets)
v, ix = shapes[0].vertices.copy(), shapes[0].indices
if refcpu.available():
    t0 = time.perf_counter(); refcpu.generate_tangents(v, ix); dt_t = time.perf_counter() - t0
    t0 = time.perf_counter(); refcpu.generate_normals(v, ix); dt_n = time.perf_counter() - t0
    out["cpu_reference"] = {"kind": "reference", "cores": 1, "sample": f"grid n={args.cpu_grid} ({len(ix) // 3} triangles)",
                            "generate_tangents_triangles_per_s": len(ix) // 3 / dt_t,
                            "generate_normals_triangles_per_s": len(ix) // 3 / dt_n}
print(json.dumps(out))
