#!/usr/bin/env python
"""Config 4 through the native batch entry point: N separate meshes (own attribute arrays each, ordinary pageable
memory) handed to ONE vkm_import_batch call, host buffers in, host buffers out, wall clock around the C call.
Prints one JSON line.

    python scripts/bench_batch_native.py [--meshes 2000] [--result-memory 1]
"""
import argparse, ctypes as C, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from vulkanengine_b200 import capi, synth

ap = argparse.ArgumentParser()
ap.add_argument("--meshes", type=int, default=2000)
ap.add_argument("--result-memory", type=int, default=1)
ap.add_argument("--repeats", type=int, default=3)
ap.add_argument("--gpus", type=int, default=1, help="GPUs of this ONE process (vkm_init([0..N-1]); the batch is LPT-partitioned over them)")
args = ap.parse_args()

meshes = synth.batch(args.meshes)
L = capi.lib()
n = len(meshes)
ins = (capi.MeshIn * n)()
outs = (C.POINTER(capi.ShapeOut) * n)()
keep = []
for i, m in enumerate(meshes):
    so = np.ascontiguousarray(m.shape_offsets, np.uint64)
    arr = (capi.ShapeOut * (len(so) - 1))()
    keep += [so, arr]
    ins[i] = capi.MeshIn(m.pos.ctypes.data, len(m.pos), m.nrm.ctypes.data, len(m.nrm), m.uv.ctypes.data, len(m.uv),
                         m.idx.ctypes.data, so.ctypes.data, len(so) - 1)
    outs[i] = C.cast(arr, C.POINTER(capi.ShapeOut))
capi.init(list(range(args.gpus)))
capi.set_result_memory(args.result_memory)
F = sum(m.triangles for m in meshes)
bytes_in = sum(m.input_bytes() for m in meshes)
best = 1e9
for r in range(args.repeats + 1):
    t0 = time.perf_counter()
    rc = L.vkm_import_batch(ins, n, outs, None)
    dt = time.perf_counter() - t0
    assert rc == 0, L.vkm_last_error()
    nv = sum(outs[i][0].vertex_count for i in range(n))
    for i in range(n):
        L.vkm_free_shape(C.byref(outs[i][0]))
    if r:
        best = min(best, dt)
print(json.dumps({"gpus": args.gpus, "meshes": n, "triangles": F, "unique_vertices": nv, "input_MB": bytes_in / 1e6,
                  "output_MB": (64 * nv + 12 * F) / 1e6, "result_memory": args.result_memory, "seconds": best,
                  "triangles_per_s": F / best, "us_per_mesh": best / n * 1e6}))
