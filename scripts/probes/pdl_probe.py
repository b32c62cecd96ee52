"""Measurement only: back-to-back imports of grid(n) with / without programmatic dependent launch (VKM_PDL=0|1).
    VKM_PDL=0 python scripts/probes/pdl_probe.py 708"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from vulkanengine_b200 import capi, synth
n = int(sys.argv[1])
m = synth.grid(n)
d = [capi.DeviceArray(0, a) for a in (m.pos, m.nrm, m.uv, m.idx)]
na = max(len(m.pos), len(m.nrm), len(m.uv))
plan = capi.Plan(0, m.corners, max_attrib=na, max_vertices=na + na // 8 + 1024)
st = torch.cuda.Stream()
def step():
    plan.import_async(d[0].ptr, len(m.pos), d[1].ptr, len(m.nrm), d[2].ptr, len(m.uv), d[3].ptr, m.shape_offsets, stream=st.cuda_stream)
for _ in range(5): step(); plan.finish()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
K = 100
e0.record(st)
for _ in range(K): step()
e1.record(st)
plan.finish(); torch.cuda.synchronize()
print(os.environ.get("VKM_PDL"), n, e0.elapsed_time(e1) / K)
