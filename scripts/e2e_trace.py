#!/usr/bin/env python
"""Host-side time line of vkm_import_shapes on the 16M mesh (VKM_TRACE=1).  `--pageable`: inputs in ordinary numpy
memory (the drop-in default: what integration/engine_shims.cpp passes)."""
import ctypes as C, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from vulkanengine_b200 import capi, synth
os.environ["VKM_TRACE"] = "1"
m = synth.grid(2896)
L = capi.lib()
def pin(a):
    p = L.vkm_host_alloc(a.nbytes); C.memmove(p, a.ctypes.data, a.nbytes); return p
if "--pageable" in sys.argv:
    hp, hn, hu, hi = (a.ctypes.data for a in (m.pos, m.nrm, m.uv, m.idx))
else:
    hp, hn, hu, hi = (pin(a) for a in (m.pos, m.nrm, m.uv, m.idx))
so = np.ascontiguousarray(m.shape_offsets, np.uint64)
out = (capi.ShapeOut * 1)()
for it in range(3):
    t0 = time.perf_counter()
    rc = L.vkm_import_shapes(hp, len(m.pos), hn, len(m.nrm), hu, len(m.uv), hi, so.ctypes.data, 1, out)
    t1 = time.perf_counter()
    L.vkm_free_shape(C.byref(out[0]))
    t2 = time.perf_counter()
    print("call %.2f ms, free %.2f ms rc=%d" % ((t1 - t0) * 1e3, (t2 - t1) * 1e3, rc), file=sys.stderr)
