#!/usr/bin/env python
"""The host<->device copy ceiling that bounds the end-to-end numbers: every rank moves one import's bytes (872 MB up,
738 MB down for the 16.8 M-triangle mesh) between page-locked host memory and its GPU, no kernels at all.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 scripts/dma_ceiling.py [--steps 10]

Prints one JSON line (rank 0): aggregate GB/s over the N ranks for upload only, download only and both directions at
once, plus the triangles/s an import could reach at most if its copies were the only cost.
"""
import argparse
import json
import os

import torch
import torch.distributed as dist

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=10)
ap.add_argument("--h2d-bytes", type=int, default=872414240)
ap.add_argument("--d2h-bytes", type=int, default=738410560)
ap.add_argument("--triangles", type=int, default=16773632)
args = ap.parse_args()

rank = int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))

h_up = torch.empty(args.h2d_bytes, dtype=torch.uint8).pin_memory()
h_dn = torch.empty(args.d2h_bytes, dtype=torch.uint8).pin_memory()
d_up = torch.empty(args.h2d_bytes, dtype=torch.uint8, device="cuda")
d_dn = torch.empty(args.d2h_bytes, dtype=torch.uint8, device="cuda")
s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()


def barrier():
    if world > 1:
        t = torch.zeros(1, device="cuda")
        dist.all_reduce(t)
    torch.cuda.synchronize()


def run(up: bool, down: bool) -> float:
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    s_up.wait_stream(torch.cuda.current_stream())
    s_dn.wait_stream(torch.cuda.current_stream())
    for _ in range(args.steps):
        if up:
            with torch.cuda.stream(s_up):
                d_up.copy_(h_up, non_blocking=True)
        if down:
            with torch.cuda.stream(s_dn):
                h_dn.copy_(d_dn, non_blocking=True)
    torch.cuda.current_stream().wait_stream(s_up)
    torch.cuda.current_stream().wait_stream(s_dn)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    return ms


run(True, True)  # warm-up
ms_up, ms_dn, ms_both = run(True, False), run(False, True), run(True, True)
if rank == 0:
    gb = lambda b, ms: b * args.steps * world / (ms * 1e-3) / 1e9
    print(json.dumps({
        "n_gpus": world, "steps": args.steps, "h2d_bytes": args.h2d_bytes, "d2h_bytes": args.d2h_bytes,
        "h2d_only_gbs": gb(args.h2d_bytes, ms_up), "d2h_only_gbs": gb(args.d2h_bytes, ms_dn),
        "both_gbs": gb(args.h2d_bytes + args.d2h_bytes, ms_both), "both_ms_per_step": ms_both / args.steps,
        "import_ceiling_triangles_per_s": args.triangles * args.steps * world / (ms_both * 1e-3),
        "how": "page-locked host memory, cudaMemcpyAsync on two streams per rank (one per direction), CUDA events, max over ranks; no kernels"}))
if world > 1:
    dist.destroy_process_group()
