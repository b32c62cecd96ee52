"""CPU tests: pin the plain-C oracle restatement (oracle/oracle_port.c) against
  (1) the reference-shipped known-answer file falcon.e3dm (via tests/golden/falcon.npz),
  (2) reference runs frozen as fixtures (viking_room, cube, semantics), and
  (3) when oracle/_ref/libvkref.so is present, the reference's own code on seeded synthetic meshes.
"""
import numpy as np
import pytest
from conftest import assert_shapes_equal, load_golden, vertices_equal

from oracle import port, refcpu
from oracle.refcpu import VERTEX_DTYPE, ShapeResult
from vulkanengine_b200 import synth


def golden_shapes(g):
    return [ShapeResult(np.ascontiguousarray(g[f"vertices{i}"]).view(VERTEX_DTYPE).reshape(-1), g[f"indices{i}"],
                        g[f"extent{i}"]) for i in range(int(g["nshapes"]))]


@pytest.mark.parametrize("name", ["falcon", "viking_room", "semantics"])
def test_port_matches_golden(name):
    g = load_golden(name)
    got = port.import_arrays(g["pos"], g["nrm"], g["uv"], g["idx"], g["shape_offsets"])
    assert_shapes_equal(got, golden_shapes(g))


def test_falcon_known_answer_shape():
    g = load_golden("falcon")
    assert g["vertices0"].shape == (5356, 64) and g["indices0"].shape == (25968,)
    idx = g["indices0"]
    # first-occurrence order: every new index is the previous maximum + 1
    running = np.maximum.accumulate(idx)
    assert idx[0] == 0 and np.all(np.diff(running) <= 1)


def test_cube_known_answer():
    g = load_golden("cube")
    v = np.ascontiguousarray(g["vertices_in"]).view(VERTEX_DTYPE).reshape(-1).copy()
    idx = g["indices"]
    port.generate_normals(v, idx)
    port.generate_tangents(v, idx)
    ok, msg = vertices_equal(v, g["vertices_out"])
    assert ok, msg
    want_n = [[0, 0, 1], [0, 0, -1], [-1, 0, 0], [1, 0, 0], [0, -1, 0], [0, 1, 0]]
    assert np.array_equal(v["normal"][::4], np.array(want_n, np.float32))


def test_hash_float_known_answers():
    # values measured on the reference build (libstdc++ 13 std::hash<float>)
    v = np.zeros(1, VERTEX_DTYPE)
    assert port.hash_vertex(v) == 0x1D02628F  # all-zero vertex
    neg = v.copy()
    neg["position"][0, 0] = -0.0
    assert port.hash_vertex(neg) == 0x1D02628F  # -0 hashes like +0


def test_semantics_rules():
    g = load_golden("semantics")
    s0, s1 = golden_shapes(g)
    # NaN position twice -> two appended vertices, index 0 written both times
    assert s0.indices[3:6].tolist() == [0, 0, 0] and len(s0.vertices) == 6
    # -0 twin merges with +0 and the stored vertex keeps the first-seen (+0) bit pattern
    assert s0.indices[6] == 0 and s0.vertices.view(np.uint32).reshape(-1, 16)[0, 0] == 0
    # 1-ulp twin does not merge; duplicate `v` record does
    assert s0.indices[10] == 5 and s0.indices[11] == 1
    # Inf behaves like NaN (|inf - inf| < eps is false)
    assert s1.indices[12] == 0
    # the box always contains the origin
    assert np.all(s0.extent[3:6] <= 0) and np.all(s0.extent[6:9] >= 0)


def test_degenerate_tangents():
    v = np.zeros(3, VERTEX_DTYPE)
    v["position"] = [[0, 0, 0], [1, 0, 0], [0, 1, 0]]
    v["texcoord"] = [[0, 0], [0, 0], [0, 0]]  # zero uv area -> fc = inf -> NaN xyz, w = +1
    idx = np.array([0, 1, 2], np.uint32)
    port.generate_tangents(v, idx)
    assert np.isnan(v["tangent"][:, :3]).all() and (v["tangent"][:, 3] == 1).all()
    v["texcoord"] = [[0, 0], [0, 1], [1, 0]]  # mirrored uv -> w = -1
    port.generate_tangents(v, idx)
    assert (v["tangent"][:, 3] == -1).all()


def test_last_triangle_wins():
    m = synth.grid(4)
    out = port.import_arrays(m.pos, m.nrm, m.uv, m.idx, m.shape_offsets)[0]
    v = out.vertices.copy()
    v["tangent"] = 0
    # recompute per vertex from the highest-numbered triangle only
    last = np.full(len(v), -1)
    for f in range(len(out.indices) // 3):
        last[out.indices[3 * f:3 * f + 3]] = f
    for vi in range(len(v)):
        tri = out.indices[3 * last[vi]:3 * last[vi] + 3].copy()
        tmp = out.vertices.copy()
        tmp["tangent"] = 0
        port.generate_tangents(tmp, tri)
        v["tangent"][vi] = tmp["tangent"][vi]
    ok, msg = vertices_equal(v, out.vertices)
    assert ok, msg


def test_dedup_utility_port_literal():
    v = np.zeros(6, VERTEX_DTYPE)
    v["position"][:, 0] = [0, 1, 2, 1, 0, 3]
    idx = np.array([0, 1, 2, 3, 4, 5], np.uint32)
    uv, ui = port.deduplicate_vertices(v, idx)
    assert len(uv) == 4 and uv["position"][:, 0].tolist() == [0, 1, 2, 3]
    assert ui.tolist() == [0, 1, 2, 1, 0, 3]


needs_ref = pytest.mark.skipif(not refcpu.available(), reason="oracle/_ref/libvkref.so not built")


@needs_ref
@pytest.mark.parametrize("mesh", [
    lambda: synth.grid(1), lambda: synth.grid(37), lambda: synth.grid(64, permute_seed=7),
    lambda: synth.uv_sphere(16, 32), lambda: synth.adversarial(48),
    lambda: synth.concat(synth.batch(6, min_tris=50, max_tris=3000)),
])
def test_port_matches_reference_code(mesh):
    m = mesh()
    want = refcpu.import_arrays(m.pos, m.nrm, m.uv, m.idx, m.shape_offsets)
    got = port.import_arrays(m.pos, m.nrm, m.uv, m.idx, m.shape_offsets)
    assert_shapes_equal(got, want)


@needs_ref
def test_port_hash_and_equality_match_reference_code():
    rng = np.random.default_rng(1)
    a = np.zeros(1, VERTEX_DTYPE)
    b = np.zeros(1, VERTEX_DTYPE)
    for i in range(3000):
        f = rng.standard_normal(16).astype(np.float32)
        if i % 3 == 0:
            f[rng.integers(0, 16)] = 0.0
        if i % 5 == 0:
            f[rng.integers(0, 16)] = -0.0
        if i % 7 == 0:
            f[rng.integers(0, 16)] = np.nan
        if i % 11 == 0:
            f[rng.integers(0, 16)] = np.inf
        a.view(np.float32)[:] = f
        g = f.copy()
        if i % 2 == 0:
            j = rng.integers(0, 16)
            g[j] = g[j] + np.float32(rng.choice([0, 5e-8, 1e-7, 2e-7, -1e-7]))
        b.view(np.float32)[:] = g
        assert refcpu.hash_vertex(a) == port.hash_vertex(a)
        assert refcpu.vertex_equal(a, b) == port.vertex_equal(a, b)


@needs_ref
def test_port_utilities_match_reference_code():
    rng = np.random.default_rng(2)
    m = synth.uv_sphere(12, 20)
    base = refcpu.import_arrays(m.pos, m.nrm, m.uv, m.idx, m.shape_offsets)[0]
    for fn_ref, fn_port in [(refcpu.generate_tangents, port.generate_tangents),
                            (refcpu.generate_normals, port.generate_normals)]:
        a, b = base.vertices.copy(), base.vertices.copy()
        fn_ref(a, base.indices)
        fn_port(b, base.indices)
        ok, msg = vertices_equal(a, b)
        assert ok, msg
    # DeduplicateVertices, literal semantics (quirks included)
    v = np.zeros(40, VERTEX_DTYPE)
    v["position"] = rng.integers(0, 4, size=(40, 3)).astype(np.float32)
    idx = rng.integers(0, 40, size=90).astype(np.uint32)
    ra, ri = refcpu.deduplicate_vertices(v, idx)
    pa, pi = port.deduplicate_vertices(v, idx)
    assert np.array_equal(ri, pi) and ra.tobytes() == pa.tobytes()


@needs_ref
def test_reference_obj_text_path_equals_injected_arrays(tmp_path):
    """The in-memory injection used for large inputs is the same code path as LoadOBJ on OBJ text."""
    m = synth.uv_sphere(10, 14)
    p = tmp_path / "sphere.obj"
    p.write_text(synth.to_obj_text(m))
    a = refcpu.load_obj(str(p), "sphere")
    b = refcpu.import_arrays(m.pos, m.nrm, m.uv, m.idx, m.shape_offsets)
    assert_shapes_equal(a, b)


GEN_CASES = [(10.0, 5.0, 3, 2, 2.0, 1.0), (1.0, 1.0, 1, 1, 1.0, 1.0), (7.3, 0.9, 17, 5, 3.0, 0.5), (0.0, 0.0, 0, 0, 0.0, 0.0)]


@needs_ref
@pytest.mark.parametrize("case", GEN_CASES)
def test_port_generators_match_reference_code(case):
    w, h, xs, ys, tx, ty = case
    rv, ri = refcpu.generate_plane(w, h, xs, ys, tx, ty)
    pv, pi = port.generate_plane(w, h, xs, ys, tx, ty)
    assert rv.tobytes() == pv.tobytes() and np.array_equal(ri, pi)
    rv, ri = refcpu.generate_cube(w, h, 2.5 * w, tx, ty)
    pv, pi = port.generate_cube(w, h, 2.5 * w, tx, ty)
    ok, msg = vertices_equal(rv, pv)
    assert ok and np.array_equal(ri, pi), msg


def test_port_cube_is_the_golden_cube():
    g = load_golden("cube")
    v, i = port.generate_cube(2.0, 2.0, 2.0, 1.0, 1.0)
    ok, msg = vertices_equal(v, g["vertices_out"])
    assert ok and np.array_equal(i, g["indices"]), msg
