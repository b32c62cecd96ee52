#!/usr/bin/env python
"""Device-resident throughput with ONE import in flight (what bench.py times) against TWO independent imports in
flight on two streams / two plans (what the two lanes of vkm_import_batch give the GPU).  Prints one JSON line."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from bench import WORKLOADS
from vulkanengine_b200 import capi, synth

wl = sys.argv[1] if len(sys.argv) > 1 else "highdup16m"
K = int(sys.argv[2]) if len(sys.argv) > 2 else 100
n, perm, _ = WORKLOADS[wl]
dev = 0
torch.cuda.set_device(dev)
meshes = [synth.grid(n, permute_seed=perm), synth.grid(n, permute_seed=perm, offset=(1.0, -1.0, 0.5), scale=1.25)]
d = [[capi.DeviceArray(dev, a) for a in (m.pos, m.nrm, m.uv, m.idx)] for m in meshes]
na = max(len(meshes[0].pos), len(meshes[0].nrm), len(meshes[0].uv))
plans = [capi.Plan(dev, m.corners, max_attrib=na, max_vertices=na + na // 8 + 1024) for m in meshes]
streams = [torch.cuda.Stream(device=dev) for _ in meshes]


def step(i):
    m, dd = meshes[i], d[i]
    plans[i].import_async(dd[0].ptr, len(m.pos), dd[1].ptr, len(m.nrm), dd[2].ptr, len(m.uv), dd[3].ptr, m.shape_offsets,
                          stream=streams[i].cuda_stream)


def run(nstreams, k):
    for i in range(nstreams):
        for _ in range(3):
            step(i)
        plans[i].finish()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(torch.cuda.current_stream())
    for s in streams[:nstreams]:
        s.wait_event(e0)
    for j in range(k):
        step(j % nstreams)
    for s in streams[:nstreams]:
        torch.cuda.current_stream().wait_stream(s)
    e1.record(torch.cuda.current_stream())
    e1.synchronize()
    for i in range(nstreams):
        plans[i].finish()
    return e0.elapsed_time(e1) / k


one = run(1, K)
two = run(2, K)
F = meshes[0].triangles
print(json.dumps({"workload": wl, "steps": K, "ms_per_import_one_in_flight": one, "ms_per_import_two_in_flight": two,
                  "triangles_per_s_one": F / one * 1e3, "triangles_per_s_two": F / two * 1e3, "gain": one / two}))
