#!/usr/bin/env python
"""OBJ text -> engine buffers: the reference's LoadOBJ (tinyobj parse + geometry stage, CPU, 1 thread) beside
vkm_import_obj_text (everything on the GPU) on the same 1M-triangle file.  Prints one JSON line."""
import json, os, sys, tempfile, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from oracle import refcpu
from vulkanengine_b200 import capi, synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 708
m = synth.grid(n)
t0 = time.perf_counter()
text = synth.to_obj_text(m).encode()
gen = time.perf_counter() - t0
d = tempfile.mkdtemp()
path = os.path.join(d, "grid.obj")
open(path, "wb").write(text)
out = {"triangles": m.triangles, "text_MB": len(text) / 1e6, "textgen_s": gen}
if refcpu.available():
    t0 = time.perf_counter(); refcpu.parse_obj(path); out["tinyobj_parse_s"] = time.perf_counter() - t0
    t0 = time.perf_counter(); ref = refcpu.load_obj(path, "grid"); out["reference_LoadOBJ_s"] = time.perf_counter() - t0
capi.parse_obj(text[:100000].rsplit(b"\n", 1)[0] + b"\n")  # warm-up (context, module load)
best_p = best_i = 1e9
for _ in range(3):
    t0 = time.perf_counter(); s = capi.parse_obj(text); best_p = min(best_p, time.perf_counter() - t0)
    t0 = time.perf_counter(); shapes, names, mats = capi.import_obj_text(text); best_i = min(best_i, time.perf_counter() - t0)
out["gpu_parse_s"] = best_p
out["gpu_import_obj_text_s"] = best_i
# the C call alone (what an engine pays): no numpy copies of the results
import ctypes as C_
L = capi.lib()
for mode, key in ((0, "c_call_import_obj_text_s"), (1, "c_call_import_obj_text_pooled_s")):
    capi.set_result_memory(mode)
    best = 1e9
    for _ in range(4):
        r = capi.ObjImport()
        t0 = time.perf_counter()
        rc = L.vkm_import_obj_text(0, text, len(text), C_.byref(r))
        dt = time.perf_counter() - t0
        assert rc == 0, L.vkm_last_error()
        L.vkm_free_obj_import(C_.byref(r))
        best = min(best, dt)
    out[key] = best
capi.set_result_memory(0)
out["c_call_text_MB_per_s"] = len(text) / 1e6 / out["c_call_import_obj_text_pooled_s"]
out["gpu_text_MB_per_s"] = len(text) / 1e6 / best_p
if refcpu.available():
    out["parse_speedup"] = out["tinyobj_parse_s"] / best_p
    out["loadobj_speedup"] = out["reference_LoadOBJ_s"] / best_i
    out["loadobj_speedup_c_call"] = out["reference_LoadOBJ_s"] / out["c_call_import_obj_text_s"]
    out["loadobj_speedup_c_call_pooled"] = out["reference_LoadOBJ_s"] / out["c_call_import_obj_text_pooled_s"]
    same = (np.array_equal(shapes[0].indices, ref[0].indices) and shapes[0].vertices.tobytes() == ref[0].vertices.tobytes())
    out["identical_to_reference"] = bool(same)
print(json.dumps(out))
