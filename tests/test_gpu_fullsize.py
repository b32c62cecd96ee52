"""Bit-exact parity at the FULL sizes BASELINE.json names, against the reference itself (oracle/_ref = the reference's
own sources compiled unmodified; it travels to the GPU box), not against properties or samples:

    config 3   16.8 M-triangle high-duplication grid, and its seeded face permutation (seed 0x5EED0003)
    config 4   every mesh of the 10 000-mesh batch, through vkm_import_batch
    config 5   67 M-triangle adversarial mesh (degenerate triangles, duplicated records, +-0 twins, 1-ulp twins, NaN,
               and key clusters engineered against the build's own table hashes)

Bars as everywhere: index buffers bit-exact, vertex floats bit-exact (NaN equals NaN), extents equal by value.
The reference loop runs at ~1.2 M triangles/s per core: ~15 s for 16.8 M, about a minute for 67 M, the batch on all
cores.  `VKM_SKIP_HUGE=1` skips this file (the only opt-out).
"""
import os
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import pytest
from conftest import assert_shapes_equal

from oracle import refcpu
from vulkanengine_b200 import capi, hash_collide, sharding, synth

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(os.environ.get("VKM_SKIP_HUGE") == "1", reason="VKM_SKIP_HUGE=1"),
              pytest.mark.skipif(not refcpu.available(), reason="oracle/_ref not built")]


def _ref(m):
    return refcpu.import_arrays(m.pos, m.nrm, m.uv, m.idx, m.shape_offsets)


def _gpu(m):
    return capi.import_shapes(m.pos, m.nrm, m.uv, m.idx, m.shape_offsets)


def _both(m):
    """The reference on a host thread while the GPU import runs (ctypes releases the GIL in both)."""
    with ThreadPoolExecutor(1) as ex:
        fut = ex.submit(_ref, m)
        got = _gpu(m)
        stats = capi.last_import_stats()
        want = fut.result()
    return got, want, stats


@pytest.mark.parametrize("perm", [None, 0x5EED0003], ids=["config3", "config3_permuted"])
def test_config3_16m_equals_reference(perm):
    m = synth.grid(2896, permute_seed=perm)
    assert m.triangles == 16_773_632
    capi.shutdown()  # a fresh host plan: it starts with the clean-mesh speculation whatever ran before
    got, want, stats = _both(m)
    assert len(want[0].vertices) == 2897 * 2897
    assert_shapes_equal(got, want)
    assert stats["schedule"] == 0 and stats["runs"] == 1, stats  # the clean speculation stood


def test_config2_uv_sphere_1m_equals_reference():
    """SURVEY 8(d) config 2's second shape at its full size: 708 x 708 lat/long sphere = 1 002 528 triangles, texture
    seam (position records shared by two uv records: the side table of the record schedule), collapsed poles (zero-area
    triangles, NaN tangents), key attribute = vt."""
    m = synth.uv_sphere(708, 708)
    assert m.triangles == 1_002_528 and len(m.uv) > len(m.pos)
    capi.shutdown()
    got, want, stats = _both(m)
    assert_shapes_equal(got, want)
    assert stats["schedule"] == 0, stats  # clean: distinct triplets carry distinct values
    assert np.isnan(want[0].vertices["tangent"]).any()  # the pole triangles are degenerate


def test_config5_adversarial_64m_equals_reference():
    m = synth.adversarial(5793, frac=2e-4, nan_frac=2e-5)
    # + keys that cluster in the build's own hashes (at this table size a 2048-probe overflow would need > 1e9
    # candidates; the overflow itself is forced at 1 M triangles in test_gpu_collisions.py)
    m = hash_collide.inject(m, value_keys=48, triplet_keys=256, window_slots=64)
    assert m.triangles > 67_000_000
    t0 = time.time()
    got, want, stats = _both(m)
    print(f"config 5: {m.triangles} triangles, {len(want[0].vertices)} vertices, both arms {time.time() - t0:.1f} s, {stats}")
    assert len(want[0].vertices) < (1 << 26)  # the reference's own u32 byte-size limit (mesh_loader.cpp:301)
    assert_shapes_equal(got, want)
    assert stats["schedule"] == 2, stats  # NaN keys and duplicated records: the general schedule decided


def test_config4_batch_10k_every_mesh_equals_reference():
    """All 10 000 meshes, in slices of 1000 (bounded host memory): vkm_import_batch against the reference run on a
    thread pool, one mesh per task."""
    ns, tris = sharding.batch_sizes(10000)
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    total = 0
    t0 = time.time()
    with ThreadPoolExecutor(max(1, cores)) as ex:
        for lo in range(0, len(ns), 1000):
            ids = range(lo, min(lo + 1000, len(ns)))
            meshes = []
            for i in ids:
                rng = np.random.Generator(np.random.PCG64([0x5EED0004, int(i)]))
                off = rng.uniform(-10.0, 10.0, size=3)
                sc = float(rng.uniform(0.5, 4.0))
                meshes.append(synth.grid(int(ns[i]), offset=tuple(float(v) for v in off), scale=sc))
            futs = [ex.submit(_ref, mm) for mm in meshes]
            got = capi.import_batch(meshes)
            for i, g, f in zip(ids, got, futs):
                w = f.result()
                try:
                    assert_shapes_equal(g, w)
                except AssertionError as e:
                    raise AssertionError(f"mesh {i} (n={int(ns[i])}): {e}") from None
                total += len(w[0].indices) // 3
    assert total == int(tris.sum())
    print(f"config 4: {total} triangles in 10000 meshes, both arms {time.time() - t0:.1f} s on {cores} cores")
