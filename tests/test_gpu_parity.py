"""GPU parity tests (pytest -m gpu, on the B200 box): the CUDA path, called through the C ABI, against
the oracle (oracle/oracle_port.c, pinned by tests/test_oracle.py) and the committed golden fixtures.

Bars: index buffers bit-exact; vertex position/normal/uv/color bit-exact; tangents bit-exact (the
kernels use round-to-nearest intrinsics without FMA contraction), NaN matches NaN regardless of
sign/payload; extents equal by value.
"""
import numpy as np
import pytest
from conftest import assert_shapes_equal, load_golden, vertices_equal

from oracle import port
from oracle.refcpu import VERTEX_DTYPE, ShapeResult
from vulkanengine_b200 import capi, synth

pytestmark = pytest.mark.gpu


def golden_shapes(g):
    return [ShapeResult(np.ascontiguousarray(g[f"vertices{i}"]).view(VERTEX_DTYPE).reshape(-1), g[f"indices{i}"],
                        g[f"extent{i}"]) for i in range(int(g["nshapes"]))]


def oracle_of(m):
    """The reference itself (oracle/_ref: its own sources compiled unmodified; built here, it travels to the GPU box)
    where it exists, else the plain-C restatement that tests/test_oracle.py pins to it."""
    from oracle import refcpu

    fn = refcpu.import_arrays if refcpu.available() else port.import_arrays
    return fn(m.pos, m.nrm, m.uv, m.idx, m.shape_offsets)


def gpu_of(m):
    return capi.import_shapes(m.pos, m.nrm, m.uv, m.idx, m.shape_offsets)


# ---------------------------------------------------------------- config 1: bundled assets (golden)
@pytest.mark.parametrize("name", ["falcon", "viking_room", "semantics"])
def test_import_matches_reference_golden(name):
    g = load_golden(name)
    got = capi.import_shapes(g["pos"], g["nrm"], g["uv"], g["idx"], g["shape_offsets"])
    assert_shapes_equal(got, golden_shapes(g))


def test_falcon_import_then_e3dm_is_the_reference_shipped_file(tmp_path):
    """End-to-end acceptance: tinyobj streams of falcon.obj -> CUDA import -> E3DM writer must
    give the byte-identical file the reference ships as assets/models/falcon.e3dm."""
    import hashlib

    g = load_golden("falcon")
    got = capi.import_shapes(g["pos"], g["nrm"], g["uv"], g["idx"], g["shape_offsets"])
    p = str(tmp_path / "falcon.e3dm")
    capi.e3dm_write(p, str(g["e3dm_name"]), got, [str(s) for s in g["shape_names"]], [str(s) for s in g["material_names"]])
    assert hashlib.sha256(open(p, "rb").read()).hexdigest() == str(g["e3dm_sha256"])


def test_cube_normals_then_tangents_golden():
    g = load_golden("cube")
    v = np.ascontiguousarray(g["vertices_in"]).view(capi.VERTEX_DTYPE).reshape(-1).copy()
    capi.generate_normals(v, g["indices"])
    capi.generate_tangents(v, g["indices"])
    ok, msg = vertices_equal(v, g["vertices_out"])
    assert ok, msg


# ---------------------------------------------------------------- synthetic meshes vs the oracle
@pytest.mark.parametrize("mesh", [
    lambda: synth.grid(1), lambda: synth.grid(2), lambda: synth.grid(37), lambda: synth.grid(200),
    lambda: synth.grid(96, permute_seed=11), lambda: synth.uv_sphere(16, 32), lambda: synth.uv_sphere(64, 1024),
    lambda: synth.adversarial(48), lambda: synth.adversarial(300, frac=5e-3, nan_frac=2e-3),
], ids=["grid1", "grid2", "grid37", "grid200", "grid96perm", "sphere16", "sphere_pole1024", "adv48", "adv300"])
def test_import_matches_oracle(mesh):
    m = mesh()
    assert_shapes_equal(gpu_of(m), oracle_of(m))


def test_config2_grid_1m_matches_oracle():
    m = synth.grid(708)  # 1 002 528 triangles
    got, want = gpu_of(m), oracle_of(m)
    assert len(want[0].vertices) == 709 * 709
    assert_shapes_equal(got, want)


def test_multi_shape_batch_matches_oracle():
    meshes = synth.batch(40, min_tris=20, max_tris=5000)
    m = synth.concat(meshes)
    assert_shapes_equal(gpu_of(m), oracle_of(m))


def test_shapes_sharing_attribute_arrays_dedup_per_shape():
    # one OBJ with two `o` groups that reference the same v/vt/vn records: indices restart per shape
    g = synth.grid(10)
    idx = np.concatenate([g.idx, g.idx[::-1].reshape(-1, 3, 3)[:, ::-1].reshape(-1, 3)])
    so = np.array([0, len(g.idx), 2 * len(g.idx)], np.uint64)
    got = capi.import_shapes(g.pos, g.nrm, g.uv, idx, so)
    want = port.import_arrays(g.pos, g.nrm, g.uv, idx, so)
    assert_shapes_equal(got, want)
    assert got[1].indices[0] == 0


def test_empty_and_ragged_inputs():
    g = synth.grid(3)
    # empty shape between two real ones, and an empty trailing shape
    so = np.array([0, 9, 9, len(g.idx), len(g.idx)], np.uint64)
    got = capi.import_shapes(g.pos, g.nrm, g.uv, g.idx, so)
    want = port.import_arrays(g.pos, g.nrm, g.uv, g.idx, so)
    assert_shapes_equal(got, want)
    assert got[1].vertices.size == 0 and got[1].indices.size == 0 and not got[1].extent.any()
    # nothing at all
    out = capi.import_shapes(g.pos, g.nrm, g.uv, np.empty((0, 3), np.int32), np.array([0, 0], np.uint64))
    assert out[0].vertices.size == 0 and out[0].indices.size == 0


def test_error_behaviour():
    g = synth.grid(3)
    bad = g.idx.copy()
    bad[5, 2] = -1  # tinyobj's "absent" marker: undefined behaviour in the reference, an error here
    with pytest.raises(capi.VkmError) as e:
        capi.import_shapes(g.pos, g.nrm, g.uv, bad, g.shape_offsets)
    assert e.value.code == 2
    bad = g.idx.copy()
    bad[7, 0] = len(g.pos)
    with pytest.raises(capi.VkmError) as e:
        capi.import_shapes(g.pos, g.nrm, g.uv, bad, g.shape_offsets)
    assert e.value.code == 2
    with pytest.raises(capi.VkmError) as e:  # not whole triangles
        capi.import_shapes(g.pos, g.nrm, g.uv, g.idx[:10], np.array([0, 10], np.uint64))
    assert e.value.code == 1
    # the library still works afterwards
    assert_shapes_equal(gpu_of(g), oracle_of(g))


def test_all_unique_corners_grows_tables():
    # no sharing at all: distinct triplets == corners, far above the optimistic table size
    rng = np.random.default_rng(5)
    n = 30000
    pos = rng.standard_normal((64, 3)).astype(np.float32)
    nrm = rng.standard_normal((64, 3)).astype(np.float32)
    uv = rng.random((n, 2)).astype(np.float32)
    idx = np.stack([rng.integers(0, 64, n), rng.integers(0, 64, n), np.arange(n)], 1).astype(np.int32)
    so = np.array([0, n], np.uint64)
    assert_shapes_equal(capi.import_shapes(pos, nrm, uv, idx, so), port.import_arrays(pos, nrm, uv, idx, so))


def test_many_triplets_per_attribute_record_falls_back_to_scrambled_table():
    # tiny attribute arrays, tens of thousands of distinct (v,n,t) combinations: the coherent (key-attribute)
    # addressing clusters, the library must switch to the scrambled table and grow it
    rng = np.random.default_rng(6)
    n = 30000
    pos = rng.standard_normal((64, 3)).astype(np.float32)
    nrm = rng.standard_normal((64, 3)).astype(np.float32)
    uv = rng.random((64, 2)).astype(np.float32)
    idx = rng.integers(0, 64, (n, 3)).astype(np.int32)
    so = np.array([0, n], np.uint64)
    want = port.import_arrays(pos, nrm, uv, idx, so)
    assert len(want[0].vertices) > 25000
    assert_shapes_equal(capi.import_shapes(pos, nrm, uv, idx, so), want)
    # and a coherent mesh right afterwards still works (the host entry retries coherent addressing per call)
    g = synth.grid(33)
    assert_shapes_equal(gpu_of(g), oracle_of(g))


def test_key_attribute_is_the_largest_array_whichever_it_is():
    """Mode-0 addressing keys the triplet table on the largest attribute array: v (grids), vt (uv sphere), and here
    vn (flat shading: one normal record per face) -- all three template instantiations of the insert kernel."""
    g = synth.grid(24)
    f = g.triangles
    rng = np.random.default_rng(12)
    nrm = rng.standard_normal((f, 3)).astype(np.float32)
    idx = g.idx.copy()
    idx[:, 1] = np.repeat(np.arange(f, dtype=np.int32), 3)
    assert len(nrm) > len(g.pos) and len(nrm) > len(g.uv)
    got = capi.import_shapes(g.pos, nrm, g.uv, idx, g.shape_offsets)
    assert_shapes_equal(got, port.import_arrays(g.pos, nrm, g.uv, idx, g.shape_offsets))
    # three variants of one key record (probe >= 2 -> displaced entries -> the exact fallback for triplet lookups)
    pos = np.tile(g.pos[:1], (4, 1)).astype(np.float32) + np.arange(4, dtype=np.float32)[:, None]
    uv = rng.random((40, 2)).astype(np.float32)
    nr = rng.standard_normal((3, 3)).astype(np.float32)
    idx2 = np.stack([rng.integers(0, 4, 300), rng.integers(0, 3, 300), np.arange(300) % 40], 1).astype(np.int32)
    so = np.array([0, 300], np.uint64)
    assert_shapes_equal(capi.import_shapes(pos, nr, uv, idx2, so), port.import_arrays(pos, nr, uv, idx2, so))


def test_value_duplicates_with_distinct_triplets_merge():
    # every attribute record exists twice: 8 different triplets per vertex value
    g = synth.grid(20)
    u = len(g.pos)
    pos, nrm, uv = np.tile(g.pos, (2, 1)), np.tile(g.nrm, (2, 1)), np.tile(g.uv, (2, 1))
    rng = np.random.default_rng(9)
    idx = g.idx + (rng.integers(0, 2, g.idx.shape) * u).astype(np.int32)
    got = capi.import_shapes(pos, nrm, uv, idx, g.shape_offsets)
    want = port.import_arrays(pos, nrm, uv, idx, g.shape_offsets)
    assert len(want[0].vertices) == u
    assert_shapes_equal(got, want)


# ---------------------------------------------------------------- stand-alone utilities
@pytest.mark.parametrize("mesh", [lambda: synth.grid(50), lambda: synth.uv_sphere(20, 40), lambda: synth.adversarial(40)])
def test_generate_tangents_and_normals_match_oracle(mesh):
    m = mesh()
    base = oracle_of(m)[0]
    rng = np.random.default_rng(3)
    for fn_gpu, fn_cpu in [(capi.generate_tangents, port.generate_tangents), (capi.generate_normals, port.generate_normals)]:
        a = base.vertices.copy()
        # poison the outputs so that untouched (unreferenced) vertices must stay untouched
        a["tangent"] = rng.standard_normal(a["tangent"].shape).astype(np.float32)
        b = a.copy()
        fn_gpu(a.view(capi.VERTEX_DTYPE), base.indices)
        fn_cpu(b, base.indices)
        ok, msg = vertices_equal(a, b)
        assert ok, msg


def test_generate_tangents_out_of_range_index_is_an_error():
    v = np.zeros(3, capi.VERTEX_DTYPE)
    with pytest.raises(capi.VkmError) as e:
        capi.generate_tangents(v, np.array([0, 1, 3], np.uint32))
    assert e.value.code == 2


def _dedup_case(rng, n, kind):
    v = np.zeros(n, capi.VERTEX_DTYPE)
    f = v.view(np.float32).reshape(n, 16)
    if kind == "exact":      # many exact duplicates, all 16 floats live
        base = rng.standard_normal((max(4, n // 5), 16)).astype(np.float32)
        f[:] = base[rng.integers(0, len(base), n)]
    elif kind == "epsilon":  # values on a 4e-8 lattice: chains A~B~C with A!~C (non-transitive epsilon equality)
        f[:, :3] = (rng.integers(0, 12, (n, 3)) * np.float32(4e-8)).astype(np.float32)
        f[:, 6] = rng.integers(0, 2, n)
    elif kind == "mixed":    # big coordinates (exact-only regime), +-0, NaN and Inf rows, epsilon-close small ones
        f[:, :3] = rng.integers(-3, 4, (n, 3)).astype(np.float32) * np.float32(1500.0)
        f[: n // 3, 0] = (rng.integers(0, 6, n // 3) * np.float32(5e-8)).astype(np.float32)
        f[rng.integers(0, n, n // 10), 1] = -0.0
        f[rng.integers(0, n, max(1, n // 20)), 2] = np.nan
        f[rng.integers(0, n, max(1, n // 20)), 0] = np.inf
        f[:, 12:16] = rng.integers(0, 2, (n, 4))
    idx = rng.integers(0, n, 3 * n).astype(np.uint32)
    return v, idx


@pytest.mark.parametrize("kind,n", [("exact", 300), ("epsilon", 400), ("mixed", 500), ("exact", 20000), ("epsilon", 5000)])
def test_deduplicate_vertices_literal_semantics_match_oracle(kind, n):
    rng = np.random.default_rng(n)
    v, idx = _dedup_case(rng, n, kind)
    want_v, want_i = port.deduplicate_vertices(v.view(VERTEX_DTYPE), idx)
    got_v, got_i = capi.deduplicate_vertices(v, idx)
    assert len(got_v) == len(want_v) and len(want_v) < n
    assert np.array_equal(got_i, want_i)
    ok, msg = vertices_equal(got_v, want_v)
    assert ok, msg


def test_deduplicate_vertices_edge_cases():
    v = np.zeros(0, capi.VERTEX_DTYPE)
    gv, gi = capi.deduplicate_vertices(v, np.zeros(0, np.uint32))
    assert len(gv) == 0 and len(gi) == 0
    v = np.zeros(3, capi.VERTEX_DTYPE)  # three identical vertices, an index beyond the array
    gv, gi = capi.deduplicate_vertices(v, np.array([0, 1, 2, 7], np.uint32))
    wv, wi = port.deduplicate_vertices(v.view(VERTEX_DTYPE), np.array([0, 1, 2, 7], np.uint32))
    assert len(gv) == len(wv) == 1 and np.array_equal(gi, wi)


@pytest.mark.parametrize("case", [(10.0, 5.0, 3, 2, 2.0, 1.0), (1.0, 1.0, 1, 1, 1.0, 1.0), (7.3, 0.9, 170, 53, 3.0, 0.5),
                                  (0.0, 0.0, 0, 0, 0.0, 0.0)])
def test_procedural_generators_match_oracle(case):
    w, h, xs, ys, tx, ty = case
    gv, gi = capi.generate_plane_config(w, h, xs, ys, tx, ty)
    pv, pi = port.generate_plane(w, h, xs, ys, tx, ty)
    assert gv.tobytes() == pv.tobytes() and np.array_equal(gi, pi)
    gv, gi = capi.generate_cube_config(w, h, 2.5 * w, tx, ty)
    pv, pi = port.generate_cube(w, h, 2.5 * w, tx, ty)
    ok, msg = vertices_equal(gv, pv)
    assert ok and np.array_equal(gi, pi), msg
    # the plane followed by DeduplicateVertices (the reference's TODO at geometry_system.cpp:227)
    if xs * ys <= 64:
        dv, di = capi.generate_plane_config(w, h, xs, ys, tx, ty, deduplicate=True)
        wv, wi = port.deduplicate_vertices(pv if False else port.generate_plane(w, h, xs, ys, tx, ty)[0],
                                           port.generate_plane(w, h, xs, ys, tx, ty)[1])
        assert dv.tobytes() == wv.tobytes() and np.array_equal(di, wi)


# ---------------------------------------------------------------- device-resident plan API
def test_plan_device_resident_matches_host_entry():
    m = synth.concat([synth.grid(64), synth.uv_sphere(12, 24), synth.adversarial(32)])
    dev = 0
    d = [capi.DeviceArray(dev, a) for a in (m.pos, m.nrm, m.uv, m.idx)]
    plan = capi.Plan(dev, m.corners, max_attrib=max(len(m.pos), len(m.nrm), len(m.uv)), max_shapes=3)
    plan.set_timing(True)
    for _ in range(2):  # plans are reusable
        plan.import_async(d[0].ptr, len(m.pos), d[1].ptr, len(m.nrm), d[2].ptr, len(m.uv), d[3].ptr, m.shape_offsets)
        r = plan.finish()
        got = plan.fetch(r, m.shape_offsets)
        assert_shapes_equal(got, oracle_of(m))
    assert sum(r.phase_ms) > 0
    with pytest.raises(capi.VkmError) as e:  # capacity is fixed at creation
        plan.import_async(d[0].ptr, len(m.pos), d[1].ptr, len(m.nrm), d[2].ptr, len(m.uv), d[3].ptr,
                          np.array([0, m.corners + 3], np.uint64))
    assert e.value.code == 6
    plan.destroy()


def test_sharded_batch_driver_matches_oracle():
    """The multi-GPU host logic (vulkanengine_b200/sharding.py) over the CUDA path, both shards on this GPU."""
    from vulkanengine_b200 import sharding

    ns, _ = sharding.batch_sizes(30, min_tris=50, max_tris=6000)
    want = sharding.run_sharded(ns, 0, 1, port.import_arrays)
    got = {}
    for rank in range(2):
        got.update(sharding.run_sharded(ns, rank, 2, capi.import_shapes, gather_fn=lambda o: [o]))
    assert got == want


def test_headless_renderer_handoff_pool():
    """f2: finished shapes land in first-fit ranges of one global vertex / index device buffer, byte-identical."""
    import ctypes as C

    L = capi.lib()
    m = synth.concat([synth.grid(20), synth.uv_sphere(8, 16), synth.grid(5)])
    d = [capi.DeviceArray(0, a) for a in (m.pos, m.nrm, m.uv, m.idx)]
    plan = capi.Plan(0, m.corners, max_attrib=max(len(m.pos), len(m.nrm), len(m.uv)), max_shapes=3)
    plan.import_async(d[0].ptr, len(m.pos), d[1].ptr, len(m.nrm), d[2].ptr, len(m.uv), d[3].ptr, m.shape_offsets)
    r = plan.finish()
    want = oracle_of(m)
    pool = C.c_void_p()
    capi._check_plain(L.vkm_pool_create(0, 1 << 20, 1 << 18, C.byref(pool)), "vkm_pool_create")
    so = np.ascontiguousarray(m.shape_offsets, np.uint64)
    geos = []
    for s in (2, 0, 1):  # any order: ranges are handed out first fit
        g = capi.PoolGeometry()
        capi._check_plain(L.vkm_pool_acquire(pool, C.byref(r), so.ctypes.data, s, C.byref(g), None), "vkm_pool_acquire")
        geos.append((s, g))
    capi._check(L.vkm_dev_sync(0))
    voff = 0
    for s, g in geos:
        assert g.vertex_offset == voff and g.vertex_bytes == 64 * len(want[s].vertices)
        voff += g.vertex_bytes
        v = capi.download(0, L.vkm_pool_vertex_buffer(pool) + g.vertex_offset, capi.VERTEX_DTYPE, g.vertex_count)
        i = capi.download(0, L.vkm_pool_index_buffer(pool) + g.index_offset, np.uint32, g.index_count)
        ok, msg = vertices_equal(v, want[s].vertices)
        assert ok, msg
        assert np.array_equal(i, want[s].indices)
    big = capi.PoolGeometry()
    assert L.vkm_pool_release(pool, C.byref(geos[0][1])) == 0
    L.vkm_pool_destroy(pool)
    plan.destroy()


def test_import_batch_pipelined_lanes_match_oracle():
    """vkm_import_batch: independent meshes over the selected GPUs, two overlapping lanes per GPU."""
    meshes = [synth.grid(40), synth.uv_sphere(12, 30), synth.adversarial(30), synth.grid(3),
              synth.concat([synth.grid(8), synth.grid(5)]), synth.grid(90, permute_seed=4), synth.adversarial(60)]
    got = capi.import_batch(meshes)
    assert len(got) == len(meshes)
    for g, m in zip(got, meshes):
        assert_shapes_equal(g, oracle_of(m))
    # a failing mesh is reported by position and does not poison the others
    bad = synth.grid(4)
    bad.idx = bad.idx.copy()
    bad.idx[2, 1] = -1
    with pytest.raises(capi.VkmError) as e:
        capi.import_batch([synth.grid(5), bad])
    assert e.value.code == 2 and "mesh 1" in str(e.value)


def test_import_batch_over_two_gpus_in_one_process():
    """The native multi-GPU scheduler: needs two visible devices (skipped on a one-GPU box)."""
    if capi.lib().vkm_device_count() < 2:
        pytest.skip("one GPU visible")
    meshes = [synth.grid(30 + 7 * i) for i in range(6)] + [synth.adversarial(25), synth.uv_sphere(10, 18)]
    capi.init([0, 1])
    try:
        got = capi.import_batch(meshes)
        for g, m in zip(got, meshes):
            assert_shapes_equal(g, oracle_of(m))
    finally:
        capi.init([0])


# ---------------------------------------------------------------- size-independent properties
# (full BASELINE sizes are compared with the reference itself in tests/test_gpu_fullsize.py)
def _check_import_properties(m, out, sample=2_000_000, seed=1):
    """Size-independent proof obligations that pin the result uniquely (for one shape):
       (a) every finite corner's vertex carries exactly the corner's expanded values (canonical -0 -> +0);
           every NaN/Inf corner wrote index 0 and created its own vertex with its own bits;
       (b) no two vertices referenced by finite corners are canonically equal (nothing left unmerged);
       (c) vertices are numbered in order of creation (first occurrence order)."""
    idx, vtx = out.indices, out.vertices
    C = m.corners
    U = len(vtx)
    pos, nrm, uv = m.pos[m.idx[:, 0]], None, None  # positions of all corners (cheap: 12 B each)
    finite = np.isfinite(pos).all(1)
    del pos
    finite &= np.isfinite(m.nrm[:, :][m.idx[:, 1]]).all(1)
    finite &= np.isfinite(m.uv[m.idx[:, 2]]).all(1)
    nonfinite_pos = np.nonzero(~finite)[0]
    assert np.all(idx[nonfinite_pos] == 0)

    def canon(a):
        b = a.view(np.uint32).copy()
        b[b == 0x80000000] = 0
        return b

    # (a) on a sample of finite corners
    rng = np.random.default_rng(seed)
    fin_pos = np.nonzero(finite)[0]
    c = fin_pos[rng.integers(0, len(fin_pos), min(sample, len(fin_pos)))]
    v = vtx[idx[c]]
    assert np.array_equal(canon(v["position"]), canon(m.pos[m.idx[c, 0]]))
    assert np.array_equal(canon(v["normal"]), canon(m.nrm[m.idx[c, 1]]))
    assert np.array_equal(canon(v["texcoord"]), canon(m.uv[m.idx[c, 2]]))
    assert not vtx["color"].any()
    # (c) creation order
    creation = np.full(U, C, dtype=np.int64)
    creation[idx[fin_pos][::-1]] = fin_pos[::-1]  # last write wins -> smallest corner per vertex id
    unref = np.nonzero(creation == C)[0]          # vertices no finite corner references = NaN/Inf-created
    assert len(unref) == len(nonfinite_pos), (len(unref), len(nonfinite_pos))
    creation[unref] = nonfinite_pos
    assert np.all(np.diff(creation) > 0)
    # NaN/Inf-created vertices carry their corner's own bits
    if len(unref):
        vb = vtx[unref]
        a = vb["position"].view(np.uint32)
        b = m.pos[m.idx[nonfinite_pos, 0]].view(np.uint32)
        assert np.all((a == b) | (np.isnan(vb["position"]) & np.isnan(m.pos[m.idx[nonfinite_pos, 0]])))
    # the stored vertex keeps the FIRST-SEEN bit pattern (e.g. which of +0 / -0 arrived first), not a canonical one
    ids = rng.integers(0, U, min(200_000, U))
    ids = ids[creation[ids] < C]
    cc = creation[ids]
    nanok = lambda a, b: np.all((a.view(np.uint32) == b.view(np.uint32)) | (np.isnan(a) & np.isnan(b)))
    assert nanok(vtx["position"][ids], m.pos[m.idx[cc, 0]])
    assert nanok(vtx["normal"][ids], m.nrm[m.idx[cc, 1]])
    assert nanok(vtx["texcoord"][ids], m.uv[m.idx[cc, 2]])
    # (b) uniqueness of canonical keys among finite-referenced vertices (64-bit mix of the 8 words)
    ref = np.ones(U, bool)
    ref[unref] = False
    k = canon(np.ascontiguousarray(vtx.view(np.uint32).reshape(-1, 16)[:, :8][ref])).astype(np.uint64)
    h = np.zeros(len(k), np.uint64)
    for j in range(8):
        h = (h ^ k[:, j]) * np.uint64(0x9E3779B97F4A7C15)
        h ^= h >> np.uint64(29)
    assert len(np.unique(h)) == len(h)


def test_properties_hold_on_medium_adversarial_and_match_oracle():
    m = synth.adversarial(700, frac=2e-3, nan_frac=1e-3)  # ~1M triangles
    out = gpu_of(m)[0]
    _check_import_properties(m, out, sample=500_000)
    assert_shapes_equal([out], oracle_of(m))


def test_host_plan_grows_when_vertices_outnumber_attribute_records():
    """The host entry points size the vertex buffer from the attribute arrays; a mesh that combines few records into
    many distinct corners (here 20 x 20 x 20 = 8000 vertices from 20-record arrays) makes the plan re-run with the
    reported count instead of failing with VKM_ERR_CAPACITY."""
    rng = np.random.default_rng(77)
    k = 20
    pos = rng.standard_normal((k, 3)).astype(np.float32)
    nrm = rng.standard_normal((k, 3)).astype(np.float32)
    uv = rng.random((k, 2)).astype(np.float32)
    combos = np.stack(np.meshgrid(np.arange(k), np.arange(k), np.arange(k), indexing="ij"), -1).reshape(-1, 3)
    combos = combos[rng.permutation(len(combos))]
    idx = np.concatenate([combos, combos[rng.integers(0, len(combos), 4000)]]).astype(np.int32)
    idx = idx[: len(idx) // 3 * 3]
    so = np.array([0, len(idx)], np.uint64)
    capi.shutdown()  # drop the cached host plan: the next call sizes a fresh one for this mesh
    for mode in (0, 1):
        capi.set_result_memory(mode)
        try:
            got = capi.import_shapes(pos, nrm, uv, idx, so)
        finally:
            capi.set_result_memory(0)
        assert got[0].vertices.shape[0] == k ** 3
        assert_shapes_equal(got, port.import_arrays(pos, nrm, uv, idx, so))
    # and the grown plan keeps serving smaller meshes
    g = synth.grid(12)
    assert_shapes_equal(gpu_of(g), oracle_of(g))
