"""GPU test of vkm_generate_tangents_dev_noalloc, the tangent entry point for callers that cannot allocate
device memory (code running inside the renderer's frame loop, or engines that have given all device memory
to their own allocator).  Its scratch is a fixed array inside the library instead of a caller buffer.

Kept in its own file, named so that pytest collects it last: if it fails under `-x`, every other test has
already run.  It must pass both alone in a fresh process and after the rest of the suite.

What is checked, through the C ABI only:
  * the output equals that of the existing entry point vkm_generate_tangents_dev (caller scratch) bit for
    bit -- the tolerance the other parity tests use: 0 ULP, NaN equal to NaN -- on two calls in a row, the
    second on a different and smaller mesh, so that words left in the shared array by the first call
    would show up as wrong tangents;
  * the same output equals the CPU oracle's GenerateTangents (the reference's own function where oracle/_ref is
    present), so that the comparison does not rest on two CUDA paths agreeing with each other;
  * a mesh of exactly VKM_NOALLOC_MAX_VERTICES vertices, the largest the built-in array takes, comes out right;
  * the call really allocates nothing: the device's free memory (cudaMemGetInfo) is the same before and after a
    series of calls;
  * sizes beyond VKM_NOALLOC_MAX_VERTICES are refused with VKM_ERR_CAPACITY and a message that says so.
"""
import ctypes as C

import numpy as np
import pytest
from conftest import vertices_equal

from vulkanengine_b200 import capi, synth

pytestmark = pytest.mark.gpu

MAX_VERTICES = 1 << 20


def _mesh_buffers(m):
    out = capi.import_shapes(m.pos, m.nrm, m.uv, m.idx, m.shape_offsets)[0]
    v = out.vertices.copy()
    rng = np.random.default_rng(len(v))
    v["tangent"] = rng.standard_normal(v["tangent"].shape).astype(np.float32)  # must be overwritten / kept alike
    return v, out.indices


def _run_both(v, idx):
    L = capi.lib()
    dev = 0
    d_a = capi.DeviceArray(dev, v)
    d_b = capi.DeviceArray(dev, v)
    d_i = capi.DeviceArray(dev, idx)
    d_s = capi.DeviceArray(dev, nbytes=4 * len(v))
    capi._check(L.vkm_generate_tangents_dev(len(v), d_a.ptr, len(idx), d_i.ptr, d_s.ptr, None))
    capi._check(L.vkm_generate_tangents_dev_noalloc(len(v), d_b.ptr, len(idx), d_i.ptr, None))
    capi._check(L.vkm_dev_sync(dev))
    a = d_a.download(capi.VERTEX_DTYPE)
    b = d_b.download(capi.VERTEX_DTYPE)
    for d in (d_a, d_b, d_i, d_s):
        d.free()
    return a, b


def test_noalloc_tangents_match_existing_entry_point_twice_in_a_row():
    # first call: larger mesh, fills many scratch words; second call: smaller mesh with degenerate triangles,
    # vertex ids that overlap the first call's
    for mesh in (synth.grid(180), synth.adversarial(40)):
        v, idx = _mesh_buffers(mesh)
        assert len(v) <= MAX_VERTICES
        a, b = _run_both(v, idx)
        ok, msg = vertices_equal(b, a)
        assert ok, msg
        assert not np.array_equal(a["tangent"].view(np.uint32), v["tangent"].view(np.uint32))  # something was computed


def _oracle_tangents(v, idx):
    from oracle import port, refcpu

    w = v.copy()
    (refcpu.generate_tangents if refcpu.available() else port.generate_tangents)(w, idx.copy())
    return w


def _noalloc(v, idx, repeats=1):
    L = capi.lib()
    d_v = capi.DeviceArray(0, v)
    d_i = capi.DeviceArray(0, idx)
    for _ in range(repeats):
        capi._check(L.vkm_generate_tangents_dev_noalloc(len(v), d_v.ptr, len(idx), d_i.ptr, None))
    capi._check(L.vkm_dev_sync(0))
    out = d_v.download(capi.VERTEX_DTYPE)
    d_v.free()
    d_i.free()
    return out


def test_noalloc_tangents_equal_the_cpu_oracle():
    for mesh in (synth.grid(150), synth.adversarial(60), synth.uv_sphere(20, 40)):
        v, idx = _mesh_buffers(mesh)
        ok, msg = vertices_equal(_noalloc(v, idx), _oracle_tangents(v, idx))
        assert ok, msg


def test_noalloc_at_exactly_the_maximum_vertex_count():
    n = 1023  # (n + 1)^2 = 1024^2 = VKM_NOALLOC_MAX_VERTICES unique vertices
    m = synth.grid(n)
    v, idx = _mesh_buffers(m)
    assert len(v) == MAX_VERTICES
    ok, msg = vertices_equal(_noalloc(v, idx), _oracle_tangents(v, idx))
    assert ok, msg


def test_noalloc_calls_leave_the_free_device_memory_unchanged():
    L = capi.lib()
    v, idx = _mesh_buffers(synth.grid(200))
    d_v = capi.DeviceArray(0, v)
    d_i = capi.DeviceArray(0, idx)
    capi._check(L.vkm_generate_tangents_dev_noalloc(len(v), d_v.ptr, len(idx), d_i.ptr, None))  # first call: symbol look-up
    capi._check(L.vkm_dev_sync(0))
    free0, total = C.c_uint64(0), C.c_uint64(0)
    capi._check(L.vkm_dev_mem_info(0, C.byref(free0), C.byref(total)))
    for _ in range(25):
        capi._check(L.vkm_generate_tangents_dev_noalloc(len(v), d_v.ptr, len(idx), d_i.ptr, None))
    capi._check(L.vkm_dev_sync(0))
    free1 = C.c_uint64(0)
    capi._check(L.vkm_dev_mem_info(0, C.byref(free1), C.byref(total)))
    assert free1.value == free0.value, (free0.value, free1.value)
    d_v.free()
    d_i.free()


def test_noalloc_rejects_sizes_beyond_the_built_in_scratch():
    L = capi.lib()
    dummy = capi.DeviceArray(0, nbytes=256)
    rc = L.vkm_generate_tangents_dev_noalloc(MAX_VERTICES + 1, dummy.ptr, 3, dummy.ptr, None)
    assert rc == 6  # VKM_ERR_CAPACITY
    msg = L.vkm_last_error().decode()
    assert "exceed" in msg and str(MAX_VERTICES) in msg
    # the limit itself is accepted as a size (nothing to do for an empty index list)
    assert L.vkm_generate_tangents_dev_noalloc(MAX_VERTICES, dummy.ptr, 0, dummy.ptr, None) == 0
    dummy.free()
