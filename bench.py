#!/usr/bin/env python
"""Benchmark of the mesh-import geometry path (BASELINE.json metric: mesh triangles/s, % HBM peak).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl reference]

One "step" = one pass of the hot path (face expansion -> first-occurrence vertex dedup -> bounding box
-> tangents) over one synthetic mesh per GPU, inputs resident in HBM, outputs left in HBM.  Multi-GPU
(torchrun, one rank per GPU): every rank imports its own independent mesh, no data-path collective
("weak" scaling); the only communication is the timing barrier / max over ranks.

The JSON line carries, besides the base contract: `roofline` (dominant kernel, algorithmic bytes over
CUDA-event time against the measured HBM copy bandwidth), `pipeline_roofline` (the whole step against
SURVEY.md 8(d)'s B_alg), `cpu_baseline` (the reference's own code, oracle/_ref, timed on this host on
a bounded sample), `e2e` (the same import through the host C-ABI entry vkm_import_shapes, H2D and D2H
inside the timed region), `gpu_launches` and `clocks`.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "mesh triangles/sec (dedup+normals+tangents)"
UNIT = "triangles/s"
def general_launches(mesh, stats) -> int:
    """Launches of one general-schedule import (vkmesh.cu plan_enqueue): 13, plus k_sample_box and k_cell_map when the
    value table is position-ordered (single shape, table of 2 x (1.25 x records + 1024) + 64 words beyond 32 MB, hash
    generation 0)."""
    na = max(len(mesh.pos), len(mesh.nrm), len(mesh.uv))
    m2 = 2 * (na + na // 4 + 1024) + 64
    ordered = len(mesh.shape_offsets) == 2 and 4 * m2 >= (32 << 20) and int(stats.get("hash_seed", 0)) == 0
    return 13 + (2 if ordered else 0)


KERNELS_PER_STEP = 8  # our kernels per clean import (clear, insert, flags, scan x2, ranks, indices, vertices); vkmesh.cu plan_enqueue
FALLBACK_HBM_GBS = 6650.0

WORKLOADS = {
    # name: (grid n, permute seed, description)
    "highdup16m": (2896, None, "synthetic 16.8M-triangle high-duplication grid mesh (valence 6), configs[2]"),
    "grid1m": (708, None, "synthetic 1.0M-triangle grid mesh, configs[1]"),
    "highdup16m_perm": (2896, 0x5EED0003, "16.8M-triangle grid, seeded random face order (locality defeated)"),
    "grid64m": (5793, None, "67.1M-triangle grid (configs[4] base, without injections)"),
    # configs[1]'s other shape: lat/long sphere with a texture seam and collapsed poles (key attribute = uv, seam variants
    # in the side table, degenerate pole triangles with NaN tangents)
    "sphere1m": (708, None, "synthetic 1.0M-triangle UV sphere (708 rings x 708 segments, seam + poles), configs[1]"),
    # strong scaling: a FIXED batch of independent meshes, LPT-partitioned over the ranks (configs[3])
    "batch10k": (None, None, "batch of 10k synthetic meshes (1k-100k triangles each, log-uniform), configs[3]"),
    # config 5: degenerate triangles, duplicated records, +-0 / 1-ulp twins, NaN -- the general (full) schedule
    "adversarial64m": (5793, None, "67.1M-triangle adversarial mesh (degenerate triangles, duplicated records, +-0 and 1-ulp twins, NaN), configs[4]"),
}


def hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """SM clock and throttle reasons sampled (NVML, every ~5 ms, in a thread) while the timed region runs."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.samples = []  # (time, sm_mhz, reasons bitmask)
        self.t0 = self.t1 = None
        self._stop = threading.Event()
        self._thread = None
        self.max_mhz = None
        self.err = None

    def _run(self):
        try:
            import pynvml

            pynvml.nvmlInit()
            visible = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = self.gpu
            if visible:
                try:
                    idx = int(visible.split(",")[self.gpu])
                except (ValueError, IndexError):
                    idx = self.gpu
            h = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
            get_reasons = getattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
                pynvml.nvmlDeviceGetCurrentClocksThrottleReasons
            while not self._stop.is_set():
                try:
                    clk = float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM))
                    try:
                        why = int(get_reasons(h))
                    except Exception as e:  # noqa: BLE001 - reasons are optional, the clock is not
                        why, self.err = 0, repr(e)
                    self.samples.append((time.time(), clk, why))
                except Exception as e:  # noqa: BLE001 - one failed query must not end the sampling
                    self.err = repr(e)
                time.sleep(0.005)
        except Exception as e:  # noqa: BLE001 - sampling must never break the benchmark
            self.err = repr(e)

    def start(self):
        self._thread = threading.Thread(target=self._run, daemon=True)
        self._thread.start()

    def mark(self, which):
        setattr(self, which, time.time())

    def stop(self):
        self._stop.set()
        if self._thread:
            self._thread.join(timeout=2)
        inside = [x for x in self.samples if self.t0 and self.t1 and self.t0 <= x[0] <= self.t1]
        use = inside if len(inside) >= 2 else self.samples
        if not use:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "samples": 0, "error": self.err}
        mask = 0
        for x in use:
            mask |= x[2]
        return {"sm_mhz": float(np.median([x[1] for x in use])), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(n for bit, n in self.REASONS.items() if mask & bit), "samples": len(use),
                "in_timed_region": len(inside) >= 2, "error": self.err}


# ------------------------------------------------------------------------------------------ workloads
def build_workload(args, rank: int = 0, world: int = 1):
    """The synthetic input of `args.workload` for this rank: (mesh streams, description, scaling)."""
    from vulkanengine_b200 import synth

    n, perm, desc = WORKLOADS[args.workload]
    if args.workload == "batch10k":
        from vulkanengine_b200 import sharding

        ns, tris = sharding.batch_sizes(args.batch_meshes)
        parts = sharding.lpt_partition(tris, world)
        mesh = sharding.build_shard(ns, parts[rank])
        desc += f"; {args.batch_meshes} meshes, LPT partition, imbalance {sharding.imbalance(tris, parts):.4f}"
        return mesh, desc, "strong"
    if args.workload == "sphere1m":
        return synth.uv_sphere(n, n), desc, "weak"
    if args.workload == "adversarial64m":
        return synth.adversarial(n, frac=2e-4, nan_frac=2e-5), desc, "weak"
    # independent meshes per rank: same shape family, different placement
    mesh = synth.grid(n, permute_seed=perm, offset=(float(rank), -float(rank), 0.5 * rank), scale=1.0 + 0.25 * rank)
    return mesh, desc, "weak"


# ------------------------------------------------------------------------------------------ reference arm
def host_cores() -> int:
    return len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)


def reference_threads(mesh, want: int) -> int:
    """Threads for concurrent reference imports of `mesh`: every core, unless memory binds (one import keeps an
    unordered_map node, a vector element and the result copies per unique vertex: about 50 B per corner on the grid
    family, budgeted at 100 B)."""
    cores = host_cores()
    t = cores if want <= 0 else min(want, cores)
    try:
        avail = os.sysconf("SC_AVPHYS_PAGES") * os.sysconf("SC_PAGE_SIZE")
        t = min(t, max(1, int(0.6 * avail // max(1, 100 * mesh.corners))))
    except (ValueError, OSError):
        pass
    return max(1, t)


def split_shapes(mesh):
    """A multi-shape stream as a list of independent single-mesh streams (views; the batch10k workload)."""
    from vulkanengine_b200.synth import MeshStreams

    so = np.asarray(mesh.shape_offsets, np.int64)
    if len(so) == 2:
        return [mesh]
    out = []
    for s in range(len(so) - 1):
        ix = mesh.idx[so[s]:so[s + 1]]
        out.append(MeshStreams(mesh.pos, mesh.nrm, mesh.uv, ix, np.array([0, len(ix)], np.uint64)))
    return out


def cpu_reference_run(mesh, imports: int, threads: int):
    """Times `imports` passes of the reference's own import loop (oracle/_ref, else the plain-C port) over `mesh` on a
    pool of `threads` host threads.  The reference imports ONE mesh on ONE thread (MeshLoader::LoadOBJ has no parallel
    section), so independent imports are the only parallelism its design admits: a pass over a single mesh is one
    task, a pass over a batch is one task per mesh.  Returns (triangles per second, kind, seconds)."""
    from concurrent.futures import ThreadPoolExecutor

    from oracle import port, refcpu

    if refcpu.available():
        fn, kind = refcpu.import_arrays, "reference"
    else:
        fn, kind = port.import_arrays, "port"
    parts = split_shapes(mesh)

    def work(m):
        fn(m.pos, m.nrm, m.uv, m.idx, m.shape_offsets, collect=False)  # ctypes releases the GIL

    t0 = time.perf_counter()
    if threads <= 1:
        for _ in range(imports):
            for m in parts:
                work(m)
    else:
        with ThreadPoolExecutor(threads) as ex:
            list(ex.map(work, [m for _ in range(imports) for m in parts]))
    dt = time.perf_counter() - t0
    return mesh.triangles * imports / dt, kind, dt


def run_reference(args, rank, world, emit):
    if rank != 0:
        return
    # The SAME mesh the GPU arm imports per step (rank 0's), K + W times: a step is one import of it by the reference's
    # own loop; the imports are independent, so they run on a pool of host threads (as many as cores and memory allow).
    mesh, desc, scaling = build_workload(args, 0, 1)
    threads = reference_threads(mesh, args.ref_threads)
    if len(mesh.shape_offsets) == 2:
        # whole waves: K imports on T threads take ceil(K / T) rounds; use the fewest threads that keep that round count
        # (20 steps on 16 cores: two rounds of 10 rather than 16 + 4)
        rounds = -(-max(args.steps, 1) // threads)
        threads = min(threads, -(-max(args.steps, 1) // rounds))
    if args.warmup:
        cpu_reference_run(mesh, args.warmup, min(threads, args.warmup) if len(mesh.shape_offsets) == 2 else threads)
    value, kind, dt = cpu_reference_run(mesh, args.steps, threads)
    sample = (f"the workload's own mesh ({mesh.triangles} triangles, {len(mesh.shape_offsets) - 1} mesh(es)), imported "
              f"{args.steps} times (= steps) by the reference's loop on a pool of {threads} host thread(s) "
              f"({host_cores()} cores; one import is single-threaded by design); g++ -O2 -ffp-contract=off")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": scaling,
        "vs_baseline": None, "dtype": "u32+f32", "data": "synthetic",
        "config": workload_config(args, mesh, desc, None, world),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def workload_config(args, mesh, desc, unique, world):
    """`config` of the JSON line: identical for both arms (the driver compares them)."""
    c = {"workload": args.workload, "description": desc, "triangles_per_gpu": mesh.triangles,
         "corners_per_gpu": mesh.corners, "meshes_per_gpu": len(mesh.shape_offsets) - 1,
         "parallelism": f"{world} rank(s), independent meshes sharded over GPUs, no collective",
         "l2": (f"inputs {mesh.input_bytes() / 1e6:.0f} MB per step exceed the 126 MB L2 (no flush needed)"
                if mesh.input_bytes() > 200e6 else "working set below L2 size: numbers are L2-warm")}
    return c


# ------------------------------------------------------------------------------------------ our arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--workload", default="highdup16m", choices=sorted(WORKLOADS))  # (reference arm: same names)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--e2e-steps", type=int, default=4)
    ap.add_argument("--batch-meshes", type=int, default=10000, help="mesh count of the batch10k workload")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--ref-threads", type=int, default=0, help="threads of the CPU reference runs (0 = every core)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world == 1 and args.gpus > 1:
        # launched without torchrun: re-exec under torchrun so that there is one process per GPU
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", str(29500 + os.getpid() % 2000), os.path.abspath(__file__),
               *sys.argv[1:]]
        sys.exit(subprocess.call(cmd))

    # stdout carries exactly ONE line (the JSON): anything a library prints there (NCCL's version banner, ...) is sent
    # to stderr by pointing fd 1 at fd 2 for the duration of the run; emit() writes the line to the real stdout.
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)

    def emit(line: dict) -> None:
        sys.stdout.flush()
        os.write(real_stdout, (json.dumps(line) + "\n").encode())

    if args.impl == "reference":
        run_reference(args, rank, world, emit)
        return

    import torch
    import torch.distributed as dist

    from vulkanengine_b200 import capi, synth

    if not torch.cuda.is_available() or capi.lib().vkm_device_count() == 0:
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = local_rank
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if world > 1:
            t = torch.zeros(1, device="cuda")
            dist.all_reduce(t)
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    mesh, desc, scaling = build_workload(args, rank, world)
    nshapes = len(mesh.shape_offsets) - 1
    F, C = mesh.triangles, mesh.corners
    max_attrib = max(len(mesh.pos), len(mesh.nrm), len(mesh.uv))

    d_pos, d_nrm, d_uv, d_idx = (capi.DeviceArray(dev, a) for a in (mesh.pos, mesh.nrm, mesh.uv, mesh.idx))
    plan = capi.Plan(dev, C, max_attrib=max_attrib, max_vertices=max_attrib + max_attrib // 8 + 1024,
                     max_shapes=nshapes)
    stream = torch.cuda.Stream(device=dev)

    def step():
        plan.import_async(d_pos.ptr, len(mesh.pos), d_nrm.ptr, len(mesh.nrm), d_uv.ptr, len(mesh.uv), d_idx.ptr,
                          mesh.shape_offsets, stream=stream.cuda_stream)

    sampler = ClockSampler(dev)
    sampler.start()
    for _ in range(args.warmup):
        step()
        r = plan.finish()
    U = int(r.total_vertices)

    # ---- timed region: K steps, device time (CUDA events on the launching stream), max over ranks
    # The K imports are queued back to back on one stream (no host sync between steps), exactly as the library launches
    # them for a caller: no events between the kernels of an import.
    barrier()
    sampler.mark("t0")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    r = plan.finish()  # waits for the last import, checks the device status word
    barrier()
    sampler.mark("t1")
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    clocks = sampler.stop()
    total_tris = sum_over_ranks(float(F)) * args.steps
    value = total_tris / (ms_total * 1e-3)
    ms_per_step = ms_total / args.steps
    # ---- the same steps once more with the plan's phase events on (a CUDA event between the kernels of every import: a
    # few microseconds per import, and no programmatic dependent launch): the per-kernel times behind `roofline`
    plan.set_timing(True)
    k2 = max(3, min(args.steps, 50))
    barrier()
    i0, i1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    i0.record(stream)
    for _ in range(k2):
        step()
    i1.record(stream)
    r = plan.finish()
    barrier()
    ms_instr = max_over_ranks(i0.elapsed_time(i1)) / k2
    plan.set_timing(False)
    phase_ms = np.array(list(r.phase_ms))

    # ---- what a caller gets besides the steady state above (device time, CUDA events around import + finish)
    def timed_import(pl, arrays, m, reps=1):
        best = []
        for _ in range(reps):
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a0.record(stream)
            pl.import_async(arrays[0].ptr, len(m.pos), arrays[1].ptr, len(m.nrm), arrays[2].ptr, len(m.uv), arrays[3].ptr,
                            m.shape_offsets, stream=stream.cuda_stream)
            pl.finish()  # a discarded speculation is re-run in here, on the same stream
            a1.record(stream)
            torch.cuda.synchronize()
            best.append(a0.elapsed_time(a1))
        return float(np.median(best)), pl.stats()

    schedules = {"steady_state": {"ms": ms_per_step, "ms_with_phase_events": ms_instr, "stats": plan.stats(),
                                  "what": "the timed region above: imports queued back to back on a warm plan whose "
                                          "speculation (clean mesh or not) matches the mesh"}}
    clean_arrays = (d_pos, d_nrm, d_uv, d_idx)
    fresh = capi.Plan(dev, C, max_attrib=max_attrib, max_vertices=max_attrib + max_attrib // 8 + 1024, max_shapes=nshapes)
    ms, st = timed_import(fresh, clean_arrays, mesh)
    schedules["first_import"] = {"ms": ms, "stats": st, "what": "first import of a newly created plan (cold tables; one import, finish included)"}
    if nshapes == 1 and not mesh.meta.get("kind") == "adversarial":
        dirty = synth.with_duplicate_record(mesh)
        d_dpos = capi.DeviceArray(dev, dirty.pos)
        d_didx = capi.DeviceArray(dev, dirty.idx)
        dirty_arrays = (d_dpos, d_nrm, d_uv, d_didx)
        ms, st = timed_import(fresh, dirty_arrays, dirty)  # the plan expected a clean mesh: record run + full re-run
        schedules["mis_speculated"] = {"ms": ms, "stats": st, "what": "the same mesh with ONE duplicated position record, after a clean import: the clean-speculating run is discarded and the general schedule runs"}
        ms, st = timed_import(fresh, dirty_arrays, dirty, reps=5)
        schedules["full_schedule"] = {"ms": ms, "stats": st, "what": "that not-clean mesh again (the plan now expects it): general schedule only, median of 5"}
        ms, st = timed_import(fresh, clean_arrays, mesh)
        schedules["clean_after_dirty"] = {"ms": ms, "stats": st, "what": "the clean mesh after a not-clean one: general schedule (the plan follows the last mesh)"}
        d_dpos.free()
        d_didx.free()
    fresh.destroy()

    # ---- roofline
    peak, peak_src = hbm_peak()
    b_alg = mesh.algorithmic_bytes(U)
    names = plan.phase_names()
    # algorithmic bytes each phase is the first to read / the one to write (they sum to B_alg)
    # (timed imports run the short schedule: the value-level proof, and with it the first read of the attribute
    # arrays, rides along in the vertex pass; `resolve_values` only sets the flag bits)
    attr_bytes = mesh.pos.nbytes + mesh.nrm.nbytes + mesh.uv.nbytes + 36 * nshapes  # attributes + extents
    full = schedules["steady_state"]["stats"].get("schedule") == 2  # general schedule: k_resolve_slots reads the attributes first
    share = {
        "insert_triplets": 36 * F,                                                  # the (v,n,t) index stream
        "resolve_values": attr_bytes if full else 0,
        "indices": 12 * F,                                                          # u32 index buffer
        "vertices": 64 * U + (0 if full else attr_bytes),                           # Vertex3D (+ attributes)
        "tangents": 0, "clear": 0, "mark_scan": 0,
    }
    dom = int(np.argmax(phase_ms[: len(names)]))
    dom_name = names[dom]
    kernels = {nm: {"ms": float(phase_ms[i]), "algorithmic_bytes": int(share.get(nm, 0))} for i, nm in enumerate(names)}
    tmap = {}
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        try:
            tmap = {k: v for k, v in json.load(open(tpath)).get(args.workload, {}).items() if not k.startswith("_")}
        except Exception:
            tmap = {}
    dom_bytes = share.get(dom_name, 0)
    dom_ach = dom_bytes / (phase_ms[dom] * 1e-3) / 1e9 if phase_ms[dom] > 0 else 0.0
    dominant = {"kernel": dom_name, "achieved": dom_ach, "peak": peak, "unit": "GB/s", "frac": dom_ach / peak,
                "traffic": tmap.get(dom_name), "algorithmic_bytes_per_launch": int(dom_bytes),
                "ms_per_launch": float(phase_ms[dom]), "share_of_step": float(phase_ms[dom] / max(phase_ms.sum(), 1e-9)),
                "note": "algorithmic bytes this phase is the first to read / the one to write (the phases' shares sum to B_alg)"}
    # roofline.achieved = B_alg / t_pipeline (the step is one import = 8 launches); the dominant
    # launch is broken out beside it, and every phase with its time and owned bytes under pipeline_roofline.
    pipe_ach = b_alg / (ms_per_step * 1e-3) / 1e9
    roofline = {"bound": "hbm", "achieved": pipe_ach, "peak": peak, "unit": "GB/s", "frac": pipe_ach / peak,
                "traffic": (sum(tmap.values()) if tmap else None), "peak_source": peak_src,
                "scope": "whole step = one import (B_alg / t_step over the timed region, SURVEY.md 8(d)); traffic = ncu DRAM bytes of its launches; per-kernel times (dominant_kernel, pipeline_roofline.kernels) come from the instrumented repeat of the same steps (schedules.steady_state.ms_with_phase_events)",
                "algorithmic_bytes_per_step": int(b_alg), "dominant_kernel": dominant}
    pipeline_roofline = {"bound": "hbm", "achieved": pipe_ach, "peak": peak, "unit": "GB/s", "frac": pipe_ach / peak,
                         "algorithmic_bytes_per_step": int(b_alg), "bytes_per_triangle": b_alg / F, "kernels": kernels}

    # ---- end to end through the host C-ABI entry point (pinned host inputs, H2D + kernels + D2H per step)
    import ctypes as C_

    L = capi.lib()
    plan.destroy()
    for d in (d_pos, d_nrm, d_uv, d_idx):
        d.free()

    def pinned_copy(a):
        p = L.vkm_host_alloc(a.nbytes)
        if not p:
            raise SystemExit("pinned host allocation failed")
        C_.memmove(p, a.ctypes.data, a.nbytes)
        return p

    h_pos, h_nrm, h_uv, h_idx = (pinned_copy(a) for a in (mesh.pos, mesh.nrm, mesh.uv, mesh.idx))
    so = np.ascontiguousarray(mesh.shape_offsets, np.uint64)
    out = (capi.ShapeOut * nshapes)()

    def e2e_step():
        rc = L.vkm_import_shapes(h_pos, len(mesh.pos), h_nrm, len(mesh.nrm), h_uv, len(mesh.uv), h_idx, so.ctypes.data,
                                 nshapes, out)
        if rc:
            raise SystemExit("vkm_import_shapes failed: " + L.vkm_last_error().decode())
        nv = sum(out[i].vertex_count for i in range(nshapes))
        ni = sum(out[i].index_count for i in range(nshapes))
        for i in range(nshapes):
            L.vkm_free_shape(C_.byref(out[i]))
        return nv, ni

    capi.init([dev])
    e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        nv, ni = e2e_step()
    torch.cuda.synchronize()
    single_s = max_over_ranks(time.perf_counter() - t0)
    barrier()
    capi.set_result_memory(1)
    e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        e2e_step()
    torch.cuda.synchronize()
    single_pooled_s = max_over_ranks(time.perf_counter() - t0)

    # the same call on PAGEABLE inputs (numpy arrays: what a drop-in caller holding tinyobj's vectors has)
    def pageable_step():
        rc = L.vkm_import_shapes(mesh.pos.ctypes.data, len(mesh.pos), mesh.nrm.ctypes.data, len(mesh.nrm),
                                 mesh.uv.ctypes.data, len(mesh.uv), mesh.idx.ctypes.data, so.ctypes.data, nshapes, out)
        if rc:
            raise SystemExit("vkm_import_shapes failed: " + L.vkm_last_error().decode())
        for i in range(nshapes):
            L.vkm_free_shape(C_.byref(out[i]))

    pageable_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        pageable_step()
    torch.cuda.synchronize()
    pageable_s = max_over_ranks(time.perf_counter() - t0)
    capi.set_result_memory(0)
    barrier()
    # exactly what integration/engine_shims.cpp does: pageable inputs, malloc()ed results, one synchronous call
    pageable_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        pageable_step()
    torch.cuda.synchronize()
    dropin_s = max_over_ranks(time.perf_counter() - t0)
    barrier()

    # The same steps through the batch entry point: every step is one mesh of the batch, its inputs are copied
    # host->device and its results device->host inside the timed region; two lanes per GPU let the copies of one
    # step overlap the kernels / copies of the next.
    # steps of the batch: 2 x e2e-steps, fewer when one step's host buffers are large (results of all steps are
    # outstanding until the batch returns)
    step_bytes = mesh.input_bytes() + 64 * U + 4 * C
    nb = int(max(2, min(2 * args.e2e_steps, 12e9 // max(step_bytes, 1))))
    ins = (capi.MeshIn * nb)()
    outs_b = (C_.POINTER(capi.ShapeOut) * nb)()
    arrs = []
    for i in range(nb):
        ins[i] = capi.MeshIn(h_pos, len(mesh.pos), h_nrm, len(mesh.nrm), h_uv, len(mesh.uv), h_idx, so.ctypes.data, nshapes)
        arr = (capi.ShapeOut * nshapes)()
        arrs.append(arr)
        outs_b[i] = C_.cast(arr, C_.POINTER(capi.ShapeOut))

    def batch_step():
        rc = L.vkm_import_batch(ins, nb, outs_b, None)
        if rc:
            raise SystemExit("vkm_import_batch failed: " + L.vkm_last_error().decode())
        for arr in arrs:
            for i in range(nshapes):
                L.vkm_free_shape(C_.byref(arr[i]))

    def timed_batch(result_mode):
        capi.set_result_memory(result_mode)
        batch_step()  # warm-up: lanes, plans, pinned staging / result blocks
        barrier()
        t0 = time.perf_counter()
        batch_step()
        torch.cuda.synchronize()
        dt = max_over_ranks(time.perf_counter() - t0)
        barrier()
        return dt

    malloc_s = timed_batch(0)
    e2e_s = timed_batch(1)
    capi.set_result_memory(0)
    e2e_value = sum_over_ranks(float(F)) * nb / e2e_s
    e2e = {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(mesh.input_bytes()),
           "d2h_bytes_per_step": int(64 * nv + 4 * ni), "ms_per_step": e2e_s / nb * 1e3, "steps": nb,
           "api": "vkm_import_batch (host pointers; inputs page-locked; results in the library's recycled page-locked "
                  "blocks, vkm_set_result_memory(1), released with vkm_free_shape inside the timed region; 2 lanes per "
                  "GPU take turns per PCIe direction so one step's upload runs against another step's download)",
           "needs": "the headline mode asks two things of the engine: page-locked input arrays (vkm_host_alloc; tinyobj's "
                    "std::vectors are pageable) and vkm_set_result_memory(1) with vkm_free_shape() in ~MeshResource "
                    "(INTEGRATION.md section 2c); `dropin_default` is the unmodified engine + integration/engine_shims.cpp",
           "dropin_default": {"value": sum_over_ranks(float(F)) * args.e2e_steps / dropin_s, "unit": UNIT,
                              "ms_per_step": dropin_s / args.e2e_steps * 1e3,
                              "api": "vkm_import_shapes, one synchronous call per step, inputs in ordinary pageable memory, "
                                     "results malloc()ed (what integration/engine_shims.cpp calls; no engine change)"},
           "malloc_results": {"value": sum_over_ranks(float(F)) * nb / malloc_s, "unit": UNIT,
                              "ms_per_step": malloc_s / nb * 1e3,
                              "api": "vkm_import_batch, results malloc()ed per shape as the reference's GeometryConfig "
                                     "owns them (default mode; one extra host copy out of page-locked staging)"},
           "single_call": {"value": sum_over_ranks(float(F)) * args.e2e_steps / single_s, "unit": UNIT,
                           "ms_per_step": single_s / args.e2e_steps * 1e3,
                           "api": "vkm_import_shapes, one synchronous call per step, malloc()ed results"},
           "single_call_pooled": {"value": sum_over_ranks(float(F)) * args.e2e_steps / single_pooled_s, "unit": UNIT,
                                  "ms_per_step": single_pooled_s / args.e2e_steps * 1e3,
                                  "api": "vkm_import_shapes, one synchronous call per step, page-locked result blocks"},
           "single_call_pageable_inputs": {"value": sum_over_ranks(float(F)) * args.e2e_steps / pageable_s, "unit": UNIT,
                                           "ms_per_step": pageable_s / args.e2e_steps * 1e3,
                                           "api": "as single_call_pooled, inputs in ordinary pageable memory (staged "
                                                  "through the library's page-locked ring by copy threads)"}}
    for p in (h_pos, h_nrm, h_uv, h_idx):
        L.vkm_host_free(p)

    # ---- CPU baseline (rank 0, N = 1 only): the reference's own code on this host, on the workload's own mesh
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        nthr = reference_threads(mesh, args.ref_threads)
        single = len(mesh.shape_offsets) == 2
        v1, kind, dt1 = cpu_reference_run(mesh, 1, 1) if single else (None, None, None)
        v, kind, dt = cpu_reference_run(mesh, nthr if single else 1, nthr)
        cpu_baseline = {"value": v, "unit": UNIT, "cores": nthr, "kind": kind,
                        "sample": f"the workload's own mesh ({F} triangles, {nshapes} mesh(es)): "
                                  + (f"{nthr} concurrent imports, one per thread" if single else f"one pass over the batch, one mesh per task on {nthr} threads")
                                  + f", {dt:.1f} s (the reference imports a mesh on one thread, so independent meshes are its only "
                                    f"parallelism); g++ -O2 -ffp-contract=off",
                        "single_thread": ({"value": v1, "unit": UNIT, "seconds": dt1} if single else None)}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": scaling, "vs_baseline": None,
            "dtype": "u32+f32", "data": "synthetic",
            "config": workload_config(args, mesh, desc, U, world), "unique_vertices_per_gpu": U,
            "roofline": roofline, "pipeline_roofline": pipeline_roofline, "schedules": schedules, "cpu_baseline": cpu_baseline, "e2e": e2e,
            "gpu_launches": (general_launches(mesh, schedules["steady_state"]["stats"]) if full else KERNELS_PER_STEP) * args.steps,
            "clocks": clocks,
        }
        emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
