"""vkm_set_result_memory(1): results of the host entry points in pooled page-locked blocks instead of malloc()ed
buffers.  Same bytes either way; blocks are recycled through vkm_free_shape; both kinds can be outstanding at once.
"""
import ctypes as C

import numpy as np
import pytest
from conftest import assert_shapes_equal

from oracle import port
from vulkanengine_b200 import capi, synth

pytestmark = pytest.mark.gpu


@pytest.fixture
def pooled():
    capi.set_result_memory(1)
    yield
    capi.set_result_memory(0)


def _meshes():
    rng_mesh = synth.adversarial(48, seed=11)  # value duplicates + NaN/Inf: the full schedule
    multi = synth.concat([synth.grid(17), synth.uv_sphere(9, 13), synth.grid(5, permute_seed=3)])
    return [synth.grid(64), rng_mesh, multi]


def test_pooled_results_equal_oracle_and_malloc_results(pooled):
    for m in _meshes():
        want = port.import_arrays(m.pos, m.nrm, m.uv, m.idx, m.shape_offsets)
        got = capi.import_shapes(m.pos, m.nrm, m.uv, m.idx, m.shape_offsets)
        assert_shapes_equal(got, want)
    capi.set_result_memory(0)
    for m in _meshes():
        want = port.import_arrays(m.pos, m.nrm, m.uv, m.idx, m.shape_offsets)
        assert_shapes_equal(capi.import_shapes(m.pos, m.nrm, m.uv, m.idx, m.shape_offsets), want)


def test_pooled_batch_equals_oracle(pooled):
    meshes = synth.batch(12, seed=5, min_tris=200, max_tris=5000)
    got = capi.import_batch(meshes)
    for m, g in zip(meshes, got):
        assert_shapes_equal(g, port.import_arrays(m.pos, m.nrm, m.uv, m.idx, m.shape_offsets))


def _raw_import(m):
    L = capi.lib()
    ns = len(m.shape_offsets) - 1
    out = (capi.ShapeOut * ns)()
    so = np.ascontiguousarray(m.shape_offsets, np.uint64)
    rc = L.vkm_import_shapes(m.pos.ctypes.data, m.pos.size // 3, m.nrm.ctypes.data, m.nrm.size // 3, m.uv.ctypes.data,
                             m.uv.size // 2, m.idx.ctypes.data, so.ctypes.data, ns, out)
    assert rc == 0, L.vkm_last_error().decode()
    return out, ns


def _snapshot(out, ns):
    return [(bytes(C.string_at(out[i].vertices, 64 * out[i].vertex_count)),
             bytes(C.string_at(out[i].indices, 4 * out[i].index_count))) for i in range(ns)]


def test_blocks_are_recycled_and_mixed_ownership_is_released_correctly(pooled):
    L = capi.lib()
    m = synth.concat([synth.grid(33), synth.grid(21)])
    a, ns = _raw_import(m)
    first_block = a[0].vertices
    # all buffers of one call live in one block: vertices of every shape, then indices of every shape
    assert a[1].vertices == a[0].vertices + 64 * a[0].vertex_count
    assert a[0].indices == a[0].vertices + 64 * (a[0].vertex_count + a[1].vertex_count)
    want = _snapshot(a, ns)
    # a second result while the first is outstanding must not alias it
    b, _ = _raw_import(m)
    assert b[0].vertices != first_block
    assert _snapshot(a, ns) == want and _snapshot(b, ns) == want
    # malloc()ed results while pooled ones are outstanding; vkm_free_shape releases each kind its own way
    capi.set_result_memory(0)
    c, _ = _raw_import(m)
    assert _snapshot(c, ns) == want
    capi.set_result_memory(1)
    # the block goes back to the pool only when its last buffer is released ...
    L.vkm_free_shape(C.byref(a[0]))
    assert a[0].vertices is None and a[0].indices is None
    d, _ = _raw_import(m)
    assert d[0].vertices != first_block
    assert (bytes(C.string_at(a[1].vertices, 64 * a[1].vertex_count)), bytes(C.string_at(a[1].indices, 4 * a[1].index_count))) == want[1]
    L.vkm_free_shape(C.byref(a[1]))
    # ... and is then the one the next call of that size gets
    e, _ = _raw_import(m)
    assert e[0].vertices == first_block
    assert _snapshot(e, ns) == want
    for r in (b, c, d, e):
        for i in range(ns):
            L.vkm_free_shape(C.byref(r[i]))


def test_empty_shapes_and_bad_mode(pooled):
    L = capi.lib()
    assert L.vkm_set_result_memory(7) != 0
    m = synth.grid(9)
    so = np.array([0, 0, len(m.idx), len(m.idx)], np.uint64)  # empty, whole mesh, empty
    got = capi.import_shapes(m.pos, m.nrm, m.uv, m.idx, so)
    assert_shapes_equal(got, port.import_arrays(m.pos, m.nrm, m.uv, m.idx, so))


def test_batch_group_with_one_bad_mesh_reports_per_mesh_status():
    """Small meshes of a batch are coalesced into one segmented import; a mesh with an out-of-range index must fail
    alone (its index may not alias a neighbour's records) while the others import normally."""
    L = capi.lib()
    meshes = synth.batch(9, seed=3, min_tris=50, max_tris=400)
    bad = 4
    meshes[bad].idx = meshes[bad].idx.copy()
    meshes[bad].idx[7, 1] = len(meshes[bad].nrm)  # one past the end of ITS normals, valid in the concatenation
    n = len(meshes)
    ins = (capi.MeshIn * n)()
    outs = (C.POINTER(capi.ShapeOut) * n)()
    keep = []
    for i, m in enumerate(meshes):
        so = np.ascontiguousarray(m.shape_offsets, np.uint64)
        arr = (capi.ShapeOut * 1)()
        keep += [so, arr]
        ins[i] = capi.MeshIn(m.pos.ctypes.data, len(m.pos), m.nrm.ctypes.data, len(m.nrm), m.uv.ctypes.data, len(m.uv),
                             m.idx.ctypes.data, so.ctypes.data, 1)
        outs[i] = C.cast(arr, C.POINTER(capi.ShapeOut))
    st = (C.c_int * n)()
    rc = L.vkm_import_batch(ins, n, outs, st)
    assert rc == 2  # VKM_ERR_BAD_INDEX
    assert f"mesh {bad}" in L.vkm_last_error().decode()
    for i, m in enumerate(meshes):
        if i == bad:
            assert st[i] != 0
            continue
        assert st[i] == 0
        o = outs[i][0]
        want = port.import_arrays(m.pos, m.nrm, m.uv, m.idx, m.shape_offsets)[0]
        got_idx = np.frombuffer(C.string_at(o.indices, 4 * o.index_count), np.uint32)
        assert np.array_equal(got_idx, want.indices)
        assert o.vertex_count == len(want.vertices)
        L.vkm_free_shape(C.byref(o))
