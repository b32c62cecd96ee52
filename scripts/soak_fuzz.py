#!/usr/bin/env python
"""Soak: mid-size random meshes (50 k .. 1.5 M triangles) through vkm_import_shapes against the reference itself
(oracle/_ref on a thread pool), bit for bit.  The seeded fuzz in tests/test_gpu_fuzz.py uses meshes of a few thousand
corners -- one or two CTAs; this one is large enough for every kernel to run with the whole grid resident, so that the
claim / publish / atomicMin windows of the insert pass and the bucket claims of the fingerprint table are exercised
under real concurrency, on clean and not-clean meshes, single and multi shape, with and without locality.

    python scripts/soak_fuzz.py --meshes 80 --seed 1            (GPU box; ~1 s per mesh)

Prints one JSON line: meshes, triangles, the schedules taken, mismatches (exit code 1 if any).
"""
import argparse
import json
import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))

from conftest import assert_shapes_equal  # noqa: E402
from oracle import refcpu  # noqa: E402
from vulkanengine_b200 import capi, synth  # noqa: E402


def cut_shapes(rng, ntri, most):
    k = int(rng.integers(1, most + 1))
    cuts = np.sort(rng.integers(0, ntri + 1, k - 1)) * 3
    return np.concatenate([[0], cuts, [3 * ntri]]).astype(np.uint64)


def windowed_permutation(rng, idx, window):
    """Faces shuffled inside consecutive windows of `window` triangles: locality from perfect to none."""
    f = idx.reshape(-1, 3, 3)
    n = len(f)
    if window <= 1:
        return idx
    order = np.arange(n)
    key = (order // window).astype(np.float64) + rng.random(n)
    return f[np.argsort(key, kind="stable")].reshape(-1, 3)


def soup(rng):
    """Arbitrary (v, vn, vt) combinations over pools of different size, with locality in v: many variants per key
    record (side table, its growth, the fall-back to the general schedule), any attribute as the key."""
    npos, nnrm, nuv = (int(rng.integers(lo, hi)) for lo, hi in ((2000, 300000), (50, 300000), (2000, 400000)))
    ntri = int(rng.integers(50000, 900000))
    pos = (rng.permutation(npos * 3).reshape(npos, 3) * 0.25 + 1.0).astype(np.float32)
    nrm = (rng.permutation(nnrm * 3).reshape(nnrm, 3) * 0.5 - 7.0).astype(np.float32)
    uv = (rng.permutation(nuv * 2).reshape(nuv, 2) * 0.125 + 0.5).astype(np.float32)
    c = 3 * ntri
    base = (np.arange(c) // int(rng.integers(2, 7))) + rng.integers(0, int(rng.integers(2, 40)), c)
    v = base % npos
    variants = int(rng.integers(1, 5))
    n = (v * 7 + rng.integers(0, variants, c)) % nnrm
    t = (v + (rng.random(c) < rng.random() * 0.3) * rng.integers(1, 50, c)) % nuv
    idx = np.stack([v, n, t], 1).astype(np.int32)
    return synth.MeshStreams(pos, nrm, uv, idx, np.array([0, c], np.uint64))


def make_dirty(rng, m):
    """Value duplicates (records copied under new indices and referenced), +-0 twins, a few NaN / Inf records."""
    pos, nrm, uv, idx = m.pos.copy(), m.nrm.copy(), m.uv.copy(), m.idx.copy()
    c = len(idx)
    k = int(rng.integers(1, 200))
    src = rng.integers(0, len(pos), k)
    pos = np.concatenate([pos, pos[src]])
    where = rng.integers(0, c, 4 * k)
    hit = np.isin(idx[where, 0], src)
    remap = {int(s): len(m.pos) + j for j, s in enumerate(src)}
    for w in where[hit][: 2 * k]:
        idx[w, 0] = remap.get(int(idx[w, 0]), idx[w, 0])  # (a corner drawn twice is already redirected)
    # corners redirected to the copies wherever the sample did not hit: take corners that use src directly
    for j, s in enumerate(src[:50]):
        users = np.nonzero(idx[:, 0] == s)[0]
        if len(users) > 1:
            idx[users[-1], 0] = len(m.pos) + j
    if rng.random() < 0.5:
        z = rng.integers(0, len(uv), 20)
        uv[z, 0] = np.float32(0.0)
        uv[z[::2], 0] = np.float32(-0.0)
    if rng.random() < 0.4:
        bad = rng.integers(0, len(nrm), 5)
        nrm[bad, int(rng.integers(0, 3))] = np.float32(np.nan) if rng.random() < 0.5 else np.float32(np.inf)
    return synth.MeshStreams(pos, nrm, uv, idx, m.shape_offsets, dict(m.meta))


def random_mesh(rng, big=False):
    kind = int(rng.integers(0, 4))
    if big:
        # large enough (> 3.3 M records) for the position-ordered value tables of both schedules to be in use
        kind = 0
        m = synth.grid(int(rng.integers(1850, 2300)), offset=tuple(float(x) for x in rng.uniform(-3, 3, 3)),
                       scale=float(rng.uniform(0.3, 5.0)))
    elif kind == 0:
        m = synth.grid(int(rng.integers(150, 860)))
    elif kind == 1:
        m = synth.uv_sphere(int(rng.integers(100, 700)), int(rng.integers(100, 900)))
    else:
        m = soup(rng)
    idx = windowed_permutation(rng, m.idx, int(rng.choice([1, 1, 64, 4096, 1 << 30])))
    ntri = len(idx) // 3
    so = cut_shapes(rng, ntri, int(rng.choice([1, 1, 4, 60])))
    m = synth.MeshStreams(m.pos, m.nrm, m.uv, np.ascontiguousarray(idx), so, dict(m.meta, kind=kind))
    if rng.random() < (0.5 if big else 0.35):
        m = make_dirty(rng, m)
        m.meta["dirty"] = True
    return m


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--meshes", type=int, default=60)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--big", action="store_true", help="4 M .. 10 M-triangle grids (position-ordered tables in use)")
    ap.add_argument("--fresh-every", type=int, default=7, help="vkm_shutdown() every N meshes: new plan, cold tables")
    args = ap.parse_args()
    assert refcpu.available(), "oracle/_ref is not built"
    rng = np.random.default_rng(0x50AC0000 + args.seed)
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    sched = {0: 0, 1: 0, 2: 0}
    stats = {"mis_speculated": 0, "table_overflows": 0, "reseeds": 0, "dirty": 0, "multi_shape": 0}
    tris = 0
    bad = []
    t0 = time.time()
    with ThreadPoolExecutor(max(1, cores - 1)) as ex:
        pending = []
        for k in range(args.meshes):
            m = random_mesh(rng, args.big)
            fut = ex.submit(refcpu.import_arrays, m.pos, m.nrm, m.uv, m.idx, m.shape_offsets)
            if k % args.fresh_every == 0:
                capi.shutdown()
            got = capi.import_shapes(m.pos, m.nrm, m.uv, m.idx, m.shape_offsets)
            st = capi.last_import_stats()
            sched[int(st["schedule"])] += 1
            for key in ("mis_speculated", "table_overflows", "reseeds"):
                stats[key] += int(st[key])
            stats["dirty"] += bool(m.meta.get("dirty"))
            stats["multi_shape"] += len(m.shape_offsets) > 2
            tris += m.triangles
            pending.append((k, m.meta, got, fut))
            while pending and (len(pending) > cores or k == args.meshes - 1):
                kk, meta, g, f = pending.pop(0)
                try:
                    assert_shapes_equal(g, f.result())
                except AssertionError as e:
                    bad.append({"mesh": kk, "meta": {a: str(b) for a, b in meta.items()}, "error": str(e)[:300]})
    print(json.dumps({"seed": args.seed, "meshes": args.meshes, "triangles": tris, "schedules": sched, **stats,
                      "mismatches": len(bad), "first_bad": bad[:3], "seconds": round(time.time() - t0, 1)}))
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
