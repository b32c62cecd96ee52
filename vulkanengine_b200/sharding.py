"""Host-side sharding of mesh batches across GPUs.

A single mesh is one dedup domain and stays on one GPU; a *batch* of independent meshes is the natural shard.
One process per GPU: every rank derives the same partition from the mesh sizes (no communication), imports its
own sub-batch as ONE multi-shape stream (segmented kernels, one launch sequence per rank) and the per-mesh
summaries are gathered on the host at the end.  There is no collective on the data path.
"""
from __future__ import annotations

import zlib

import numpy as np

from . import synth


def lpt_partition(sizes, nranks: int) -> list:
    """Longest-processing-time greedy: meshes sorted by size descending, each to the least-loaded rank.
    Deterministic (ties by mesh id, then by rank id).  Returns a list of sorted mesh-id lists, one per rank."""
    sizes = np.asarray(sizes, dtype=np.int64)
    order = np.lexsort((np.arange(len(sizes)), -sizes))
    loads = [0] * nranks
    parts = [[] for _ in range(nranks)]
    for i in order:
        r = min(range(nranks), key=lambda k: (loads[k], k))
        parts[r].append(int(i))
        loads[r] += int(sizes[i])
    return [sorted(p) for p in parts]


def imbalance(sizes, parts) -> float:
    sizes = np.asarray(sizes, dtype=np.int64)
    loads = [int(sizes[p].sum()) if p else 0 for p in parts]
    mean = sum(loads) / max(len(loads), 1)
    return max(loads) / mean if mean else 1.0


def batch_sizes(count: int, seed: int = 0x5EED0004, min_tris: int = 1000, max_tris: int = 100_000):
    """Grid sizes n_i (and triangle counts 2 n_i^2) of the config-4 batch, without building the meshes."""
    rng = np.random.Generator(np.random.PCG64(seed))
    logs = rng.uniform(np.log(min_tris), np.log(max_tris), size=count)
    ns = np.maximum(2, np.rint(np.sqrt(np.exp(logs) / 2.0)).astype(np.int64))
    return ns, 2 * ns * ns


def build_shard(ns, mesh_ids, seed: int = 0x5EED0004) -> synth.MeshStreams:
    """Concatenated multi-shape stream of the meshes `mesh_ids` (each a grid patch with its own placement)."""
    meshes = []
    for i in mesh_ids:
        rng = np.random.Generator(np.random.PCG64([seed, int(i)]))
        off = rng.uniform(-10.0, 10.0, size=3)
        sc = float(rng.uniform(0.5, 4.0))
        meshes.append(synth.grid(int(ns[i]), offset=tuple(float(v) for v in off), scale=sc))
    return synth.concat(meshes)


def summarize(shapes) -> list:
    """Per-mesh (vertex_count, index_count, crc32(indices), crc32(vertex bytes with NaN canonicalised))."""
    out = []
    for s in shapes:
        v = np.ascontiguousarray(s.vertices).view(np.uint32).reshape(-1, 16).copy()
        f = v.view(np.float32)
        v[np.isnan(f)] = 0x7FC00000  # NaN payloads differ between CPU and GPU
        out.append((int(len(s.vertices)), int(len(s.indices)), zlib.crc32(np.ascontiguousarray(s.indices).tobytes()),
                    zlib.crc32(v.tobytes())))
    return out


def run_sharded(ns, rank: int, world: int, import_fn, gather_fn=None, seed: int = 0x5EED0004):
    """Imports this rank's share of the batch with `import_fn(pos, nrm, uv, idx, shape_offsets) -> shapes` and
    gathers {mesh id: summary} over all ranks with `gather_fn(obj) -> list of objs` (e.g. all_gather_object)."""
    tris = 2 * np.asarray(ns, dtype=np.int64) ** 2
    parts = lpt_partition(tris, world)
    mine = parts[rank]
    local = {}
    if mine:
        m = build_shard(ns, mine, seed)
        shapes = import_fn(m.pos, m.nrm, m.uv, m.idx, m.shape_offsets)
        local = dict(zip(mine, summarize(shapes)))
    gathered = gather_fn(local) if gather_fn else [local]
    merged = {}
    for part in gathered:
        for k, val in part.items():
            assert k not in merged, f"mesh {k} imported twice"
            merged[k] = val
    return merged
