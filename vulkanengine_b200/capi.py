"""ctypes bindings for the C ABI of libvkmesh_b200.so (include/vkmesh_b200.h).

These are what tests/ and bench.py call; the names mirror the reference's own interface for the path
(MeshLoader::LoadOBJ's per-shape loop, Geometry::GenerateTangents / GenerateNormals /
DeduplicateVertices).  There is no fallback: if the shared library is missing or there is no CUDA
device the calls raise.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass

import numpy as np

from . import build as _build

VERTEX_DTYPE = np.dtype(
    [("position", "<f4", 3), ("normal", "<f4", 3), ("texcoord", "<f4", 2), ("color", "<f4", 4), ("tangent", "<f4", 4)]
)

VKM_OK = 0
ERR_NAMES = {1: "INVALID_ARG", 2: "BAD_INDEX", 3: "OOM", 4: "CUDA", 5: "NO_DEVICE", 6: "CAPACITY", 7: "UNSUPPORTED"}

# every symbol include/vkmesh_b200.h declares (tests check that the library exports all of them)
SYMBOLS = [
    "vkm_device_count", "vkm_init", "vkm_shutdown", "vkm_last_error", "vkm_version",
    "vkm_generate_tangents", "vkm_generate_normals", "vkm_deduplicate_vertices",
    "vkm_import_shapes", "vkm_free_shape", "vkm_set_result_memory", "vkm_host_alloc", "vkm_host_free", "vkm_import_batch",
    "vkm_e3dm_write", "vkm_e3dm_read", "vkm_e3dm_free", "vkm_parse_obj", "vkm_free_obj_streams",
    "vkm_import_obj_text", "vkm_free_obj_import", "vkm_parse_obj_mtl", "vkm_import_obj_text_mtl", "vkm_generate_plane_config", "vkm_generate_cube_config",
    "vkm_plan_create", "vkm_plan_destroy", "vkm_plan_import", "vkm_plan_finish", "vkm_plan_set_timing",
    "vkm_plan_phase_name", "vkm_plan_get_stats", "vkm_last_import_stats", "vkm_generate_tangents_dev", "vkm_generate_normals_dev",
    "vkm_generate_tangents_dev_noalloc",
    "vkm_rangealloc_create", "vkm_rangealloc_destroy", "vkm_rangealloc_alloc", "vkm_rangealloc_free",
    "vkm_rangealloc_free_space", "vkm_pool_create", "vkm_pool_destroy", "vkm_pool_vertex_buffer",
    "vkm_pool_index_buffer", "vkm_pool_acquire", "vkm_pool_release",
    "vkm_dev_alloc", "vkm_dev_free", "vkm_dev_upload", "vkm_dev_download", "vkm_dev_sync", "vkm_dev_mem_info",
]


class VkmError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"vkmesh_b200 error {code} ({ERR_NAMES.get(code, '?')}): {msg}")
        self.code = code


class Extent(C.Structure):
    _fields_ = [("center", C.c_float * 3), ("min_extents", C.c_float * 3), ("max_extents", C.c_float * 3)]


class ShapeOut(C.Structure):
    _fields_ = [("vertex_size", C.c_uint32), ("vertex_count", C.c_uint32), ("vertices", C.c_void_p),
                ("index_size", C.c_uint32), ("index_count", C.c_uint32), ("indices", C.c_void_p),
                ("extent", Extent)]


class MeshIn(C.Structure):
    _fields_ = [("pos", C.c_void_p), ("np", C.c_uint64), ("nrm", C.c_void_p), ("nn", C.c_uint64), ("uv", C.c_void_p),
                ("nt", C.c_uint64), ("idx_vnt", C.c_void_p), ("shape_corner_offsets", C.c_void_p), ("nshapes", C.c_uint32)]


class E3dmConfig(C.Structure):
    _fields_ = [("shape", ShapeOut), ("name", C.c_char_p), ("material_name", C.c_char_p)]


class ObjStreams(C.Structure):
    _fields_ = [("pos", C.POINTER(C.c_float)), ("np", C.c_uint64), ("nrm", C.POINTER(C.c_float)), ("nn", C.c_uint64),
                ("uv", C.POINTER(C.c_float)), ("nt", C.c_uint64), ("idx_vnt", C.POINTER(C.c_int32)),
                ("shape_corner_offsets", C.POINTER(C.c_uint64)), ("nshapes", C.c_uint32),
                ("shape_names", C.POINTER(C.c_char_p)), ("usemtl_names", C.POINTER(C.c_char_p)),
                ("material_names", C.POINTER(C.c_char_p))]


class ObjImport(C.Structure):
    _fields_ = [("nshapes", C.c_uint32), ("shapes", C.POINTER(ShapeOut)), ("shape_names", C.POINTER(C.c_char_p)),
                ("usemtl_names", C.POINTER(C.c_char_p)), ("material_names", C.POINTER(C.c_char_p))]


class MtlSource(C.Structure):
    _fields_ = [("filename", C.c_char_p), ("text", C.c_char_p), ("len", C.c_uint64)]


class PoolGeometry(C.Structure):
    _fields_ = [("vertex_offset", C.c_uint64), ("vertex_bytes", C.c_uint64), ("index_offset", C.c_uint64),
                ("index_bytes", C.c_uint64), ("vertex_count", C.c_uint32), ("index_count", C.c_uint32)]


class PlanDesc(C.Structure):
    _fields_ = [("device", C.c_int), ("max_corners", C.c_uint64), ("max_vertices", C.c_uint64),
                ("max_attrib", C.c_uint64), ("max_shapes", C.c_uint32)]


class PlanStats(C.Structure):
    _fields_ = [("runs", C.c_uint32), ("schedule", C.c_uint32), ("mis_speculated", C.c_uint32),
                ("table_overflows", C.c_uint32), ("reseeds", C.c_uint32), ("hash_seed", C.c_uint32)]

    def as_dict(self):
        return {k: int(getattr(self, k)) for k, _ in self._fields_}


class PlanResult(C.Structure):
    _fields_ = [("total_vertices", C.c_uint64), ("total_indices", C.c_uint64), ("d_vertices", C.c_void_p),
                ("d_indices", C.c_void_p), ("shape_vertex_offsets", C.POINTER(C.c_uint32)),
                ("shape_extents", C.POINTER(Extent)), ("phase_ms", C.c_float * 8)]


_lib = None


def lib_path() -> str:
    return _build.LIB_PATH


def lib():
    """Loads the shared library (building it first if the sources are newer).  Raises if that fails."""
    global _lib
    if _lib is not None:
        return _lib
    path = os.environ.get("VKM_LIB_PATH") or _build.LIB_PATH  # (override: experimental builds of the same ABI)
    if not os.path.exists(path):
        path = _build.build()
    L = C.CDLL(path)
    L.vkm_last_error.restype = C.c_char_p
    L.vkm_version.restype = C.c_char_p
    L.vkm_device_count.restype = C.c_int
    L.vkm_init.argtypes = [C.POINTER(C.c_int), C.c_int]
    for fn in (L.vkm_generate_tangents, L.vkm_generate_normals):
        fn.restype = C.c_int
        fn.argtypes = [C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p]
    L.vkm_deduplicate_vertices.restype = C.c_uint32
    L.vkm_deduplicate_vertices.argtypes = [C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p, C.POINTER(C.c_void_p)]
    L.vkm_import_shapes.restype = C.c_int
    L.vkm_import_shapes.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p,
                                    C.c_void_p, C.c_uint32, C.POINTER(ShapeOut)]
    L.vkm_free_shape.argtypes = [C.POINTER(ShapeOut)]
    L.vkm_import_batch.restype = C.c_int
    L.vkm_import_batch.argtypes = [C.POINTER(MeshIn), C.c_uint32, C.POINTER(C.POINTER(ShapeOut)), C.POINTER(C.c_int)]
    L.vkm_e3dm_write.restype = C.c_int
    L.vkm_e3dm_write.argtypes = [C.c_char_p, C.c_char_p, C.POINTER(E3dmConfig), C.c_uint64]
    L.vkm_e3dm_read.restype = C.c_int
    L.vkm_e3dm_read.argtypes = [C.c_char_p, C.c_char_p, C.c_int, C.POINTER(C.POINTER(E3dmConfig)), C.POINTER(C.c_uint64)]
    L.vkm_e3dm_free.argtypes = [C.POINTER(E3dmConfig), C.c_uint64]
    L.vkm_parse_obj.restype = C.c_int
    L.vkm_parse_obj.argtypes = [C.c_int, C.c_char_p, C.c_uint64, C.POINTER(ObjStreams)]
    L.vkm_free_obj_streams.argtypes = [C.POINTER(ObjStreams)]
    L.vkm_import_obj_text.restype = C.c_int
    L.vkm_import_obj_text.argtypes = [C.c_int, C.c_char_p, C.c_uint64, C.POINTER(ObjImport)]
    L.vkm_free_obj_import.argtypes = [C.POINTER(ObjImport)]
    L.vkm_generate_plane_config.restype = C.c_int
    L.vkm_generate_plane_config.argtypes = [C.c_float, C.c_float, C.c_uint32, C.c_uint32, C.c_float, C.c_float, C.c_int, C.POINTER(ShapeOut)]
    L.vkm_generate_cube_config.restype = C.c_int
    L.vkm_generate_cube_config.argtypes = [C.c_float, C.c_float, C.c_float, C.c_float, C.c_float, C.POINTER(ShapeOut)]
    L.vkm_set_result_memory.restype = C.c_int
    L.vkm_set_result_memory.argtypes = [C.c_int]
    L.vkm_host_alloc.restype = C.c_void_p
    L.vkm_host_alloc.argtypes = [C.c_uint64]
    L.vkm_host_free.argtypes = [C.c_void_p]
    L.vkm_plan_create.restype = C.c_int
    L.vkm_plan_create.argtypes = [C.POINTER(PlanDesc), C.POINTER(C.c_void_p)]
    L.vkm_plan_destroy.argtypes = [C.c_void_p]
    L.vkm_plan_import.restype = C.c_int
    L.vkm_plan_import.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64,
                                  C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]
    L.vkm_plan_finish.restype = C.c_int
    L.vkm_plan_finish.argtypes = [C.c_void_p, C.POINTER(PlanResult)]
    L.vkm_plan_set_timing.argtypes = [C.c_void_p, C.c_int]
    L.vkm_plan_get_stats.argtypes = [C.c_void_p, C.POINTER(PlanStats)]
    L.vkm_last_import_stats.argtypes = [C.POINTER(PlanStats)]
    L.vkm_plan_phase_name.restype = C.c_char_p
    L.vkm_plan_phase_name.argtypes = [C.c_int]
    for fn in (L.vkm_generate_tangents_dev, L.vkm_generate_normals_dev):
        fn.restype = C.c_int
        fn.argtypes = [C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]
    L.vkm_generate_tangents_dev_noalloc.restype = C.c_int
    L.vkm_generate_tangents_dev_noalloc.argtypes = [C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p]
    L.vkm_rangealloc_create.restype = C.c_void_p
    L.vkm_rangealloc_create.argtypes = [C.c_uint64]
    L.vkm_rangealloc_destroy.argtypes = [C.c_void_p]
    L.vkm_rangealloc_alloc.restype = C.c_uint64
    L.vkm_rangealloc_alloc.argtypes = [C.c_void_p, C.c_uint64]
    L.vkm_rangealloc_free.restype = C.c_int
    L.vkm_rangealloc_free.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64]
    L.vkm_rangealloc_free_space.restype = C.c_uint64
    L.vkm_rangealloc_free_space.argtypes = [C.c_void_p]
    L.vkm_pool_create.restype = C.c_int
    L.vkm_pool_create.argtypes = [C.c_int, C.c_uint64, C.c_uint64, C.POINTER(C.c_void_p)]
    L.vkm_pool_destroy.argtypes = [C.c_void_p]
    L.vkm_pool_vertex_buffer.restype = C.c_void_p
    L.vkm_pool_vertex_buffer.argtypes = [C.c_void_p]
    L.vkm_pool_index_buffer.restype = C.c_void_p
    L.vkm_pool_index_buffer.argtypes = [C.c_void_p]
    L.vkm_pool_acquire.restype = C.c_int
    L.vkm_pool_acquire.argtypes = [C.c_void_p, C.POINTER(PlanResult), C.c_void_p, C.c_uint32, C.POINTER(PoolGeometry), C.c_void_p]
    L.vkm_pool_release.restype = C.c_int
    L.vkm_pool_release.argtypes = [C.c_void_p, C.POINTER(PoolGeometry)]
    L.vkm_dev_alloc.restype = C.c_int
    L.vkm_dev_alloc.argtypes = [C.c_int, C.c_uint64, C.POINTER(C.c_void_p)]
    L.vkm_dev_free.argtypes = [C.c_int, C.c_void_p]
    L.vkm_dev_upload.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_uint64]
    L.vkm_dev_download.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_uint64]
    L.vkm_dev_sync.argtypes = [C.c_int]
    L.vkm_dev_mem_info.argtypes = [C.c_int, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    _lib = L
    return L


def _check(rc: int) -> None:
    if rc != VKM_OK:
        raise VkmError(rc, lib().vkm_last_error().decode())


@dataclass
class ShapeResult:
    vertices: np.ndarray  # VERTEX_DTYPE [U]
    indices: np.ndarray  # uint32 [C]
    extent: np.ndarray  # float32 [9]: center, min_extents, max_extents


def _copy_out(ptr, nbytes, dtype):
    if not nbytes:
        return np.empty(0, dtype)
    return np.frombuffer((C.c_char * nbytes).from_address(ptr), dtype=dtype).copy()


def set_result_memory(mode: int) -> None:
    """0: malloc()ed results (engine-compatible default); 1: pooled page-locked results (release with vkm_free_shape)."""
    _check(lib().vkm_set_result_memory(int(mode)))


def shutdown() -> None:
    """vkm_shutdown: release the cached host-entry state (plans, staging, idle result blocks); the next call re-creates it."""
    lib().vkm_shutdown()


def init(devices=None) -> None:
    if devices is None:
        _check(lib().vkm_init(None, 0))
    else:
        arr = (C.c_int * len(devices))(*devices)
        _check(lib().vkm_init(arr, len(devices)))


def import_shapes(pos, nrm, uv, idx, shape_offsets) -> list:
    """vkm_import_shapes: host arrays in, list of ShapeResult out (the per-shape loop of LoadOBJ)."""
    L = lib()
    pos = np.ascontiguousarray(pos, np.float32)
    nrm = np.ascontiguousarray(nrm, np.float32)
    uv = np.ascontiguousarray(uv, np.float32)
    idx = np.ascontiguousarray(idx, np.int32)
    so = np.ascontiguousarray(shape_offsets, np.uint64)
    ns = len(so) - 1
    out = (ShapeOut * max(ns, 1))()
    _check(L.vkm_import_shapes(pos.ctypes.data, pos.size // 3, nrm.ctypes.data, nrm.size // 3, uv.ctypes.data,
                               uv.size // 2, idx.ctypes.data, so.ctypes.data, ns, out))
    res = []
    for i in range(ns):
        o = out[i]
        assert o.vertex_size == 64 and o.index_size == 4
        v = _copy_out(o.vertices, 64 * o.vertex_count, VERTEX_DTYPE)
        ix = _copy_out(o.indices, 4 * o.index_count, np.uint32)
        e = np.array(list(o.extent.center) + list(o.extent.min_extents) + list(o.extent.max_extents), np.float32)
        res.append(ShapeResult(v, ix, e))
        L.vkm_free_shape(C.byref(o))
    return res


def import_batch(meshes: list) -> list:
    """vkm_import_batch: a list of objects with pos/nrm/uv/idx/shape_offsets -> list of lists of ShapeResult."""
    L = lib()
    n = len(meshes)
    ins = (MeshIn * max(n, 1))()
    outs = (C.POINTER(ShapeOut) * max(n, 1))()
    keep, arrays = [], []
    for i, m in enumerate(meshes):
        pos = np.ascontiguousarray(m.pos, np.float32)
        nrm = np.ascontiguousarray(m.nrm, np.float32)
        uv = np.ascontiguousarray(m.uv, np.float32)
        idx = np.ascontiguousarray(m.idx, np.int32)
        so = np.ascontiguousarray(m.shape_offsets, np.uint64)
        keep += [pos, nrm, uv, idx, so]
        ns = len(so) - 1
        ins[i] = MeshIn(pos.ctypes.data, pos.size // 3, nrm.ctypes.data, nrm.size // 3, uv.ctypes.data, uv.size // 2,
                        idx.ctypes.data, so.ctypes.data, ns)
        arr = (ShapeOut * max(ns, 1))()
        arrays.append((arr, ns))
        outs[i] = C.cast(arr, C.POINTER(ShapeOut))
    st = (C.c_int * max(n, 1))()
    _check(L.vkm_import_batch(ins, n, outs, st))
    res = []
    for arr, ns in arrays:
        shapes = []
        for k in range(ns):
            o = arr[k]
            e = np.array(list(o.extent.center) + list(o.extent.min_extents) + list(o.extent.max_extents), np.float32)
            shapes.append(ShapeResult(_copy_out(o.vertices, 64 * o.vertex_count, VERTEX_DTYPE),
                                      _copy_out(o.indices, 4 * o.index_count, np.uint32), e))
            L.vkm_free_shape(C.byref(o))
        res.append(shapes)
    return res


def generate_tangents(vertices: np.ndarray, indices: np.ndarray) -> None:
    assert vertices.dtype == VERTEX_DTYPE and vertices.flags.c_contiguous
    idx = np.ascontiguousarray(indices, np.uint32)
    _check(lib().vkm_generate_tangents(len(vertices), vertices.ctypes.data, len(idx), idx.ctypes.data))


def generate_normals(vertices: np.ndarray, indices: np.ndarray) -> None:
    assert vertices.dtype == VERTEX_DTYPE and vertices.flags.c_contiguous
    idx = np.ascontiguousarray(indices, np.uint32)
    _check(lib().vkm_generate_normals(len(vertices), vertices.ctypes.data, len(idx), idx.ctypes.data))


def deduplicate_vertices(vertices: np.ndarray, indices: np.ndarray):
    assert vertices.dtype == VERTEX_DTYPE
    idx = np.ascontiguousarray(indices, np.uint32).copy()
    out = C.c_void_p()
    n = lib().vkm_deduplicate_vertices(len(vertices), vertices.ctypes.data, len(idx), idx.ctypes.data, C.byref(out))
    if n == 0xFFFFFFFF:
        raise VkmError(-1, lib().vkm_last_error().decode())
    v = _copy_out(out.value, 64 * n, VERTEX_DTYPE)
    libc = C.CDLL(None)
    libc.free.argtypes = [C.c_void_p]
    libc.free(out)  # malloc()ed by the library, as the reference's Platform::AllocMemory result is
    return v, idx


@dataclass
class ParsedObj:
    pos: np.ndarray
    nrm: np.ndarray
    uv: np.ndarray
    idx: np.ndarray
    shape_offsets: np.ndarray
    shape_names: list
    usemtl_names: list
    material_names: list = None  # GeometryConfig::material_name per shape ("" when the usemtl resolved to no material)


def _mtl_array(mtls):
    """{filename: text bytes} -> (MtlSource array, count, keep-alive list)."""
    items = list((mtls or {}).items())
    arr = (MtlSource * max(1, len(items)))()
    keep = []
    for i, (name, text) in enumerate(items):
        nb = name.encode() if isinstance(name, str) else name
        keep += [nb, text]
        arr[i] = MtlSource(nb, text, len(text))
    return arr, len(items), keep


def parse_obj(text: bytes, device: int = 0, mtls: dict | None = None) -> ParsedObj:
    """vkm_parse_obj[_mtl]: OBJ text (+ the .mtl texts it may name) -> tinyobj-shaped streams, parsed on the GPU."""
    L = lib()
    s = ObjStreams()
    arr, n, _keep = _mtl_array(mtls)
    _check(L.vkm_parse_obj_mtl(device, text, len(text), arr, n, C.byref(s)))
    try:
        ns = s.nshapes
        ncorners = int(s.shape_corner_offsets[ns]) if ns else 0
        f = lambda p, n: np.ctypeslib.as_array(p, shape=(n,)).copy() if n else np.empty(0, np.float32)
        pos = f(s.pos, 3 * s.np).reshape(-1, 3)
        nrm = f(s.nrm, 3 * s.nn).reshape(-1, 3)
        uv = f(s.uv, 2 * s.nt).reshape(-1, 2)
        idx = (np.ctypeslib.as_array(s.idx_vnt, shape=(3 * ncorners,)).copy() if ncorners else np.empty(0, np.int32)).reshape(-1, 3)
        so = np.array([s.shape_corner_offsets[i] for i in range(ns + 1)], np.uint64)
        names = [s.shape_names[i].decode() for i in range(ns)]
        mats = [s.usemtl_names[i].decode() for i in range(ns)]
        return ParsedObj(pos, nrm, uv, idx, so, names, mats, [s.material_names[i].decode() for i in range(ns)])
    finally:
        L.vkm_free_obj_streams(C.byref(s))


def import_obj_text(text: bytes, device: int = 0, mtls: dict | None = None):
    """vkm_import_obj_text[_mtl]: OBJ text -> finished shapes, all on the GPU.  Returns (shapes, names, material names):
    with `mtls` ({filename: bytes}) the material names are GeometryConfig::material_name as the engine derives it,
    without it the raw usemtl names (the entry point without .mtl sources resolves every material to "")."""
    L = lib()
    r = ObjImport()
    arr, n, _keep = _mtl_array(mtls)
    _check(L.vkm_import_obj_text_mtl(device, text, len(text), arr, n, C.byref(r)))
    try:
        shapes, names, mats = [], [], []
        for i in range(r.nshapes):
            o = r.shapes[i]
            e = np.array(list(o.extent.center) + list(o.extent.min_extents) + list(o.extent.max_extents), np.float32)
            shapes.append(ShapeResult(_copy_out(o.vertices, 64 * o.vertex_count, VERTEX_DTYPE),
                                      _copy_out(o.indices, 4 * o.index_count, np.uint32), e))
            names.append(r.shape_names[i].decode())
            mats.append((r.material_names[i] if mtls is not None else r.usemtl_names[i]).decode())
        return shapes, names, mats
    finally:
        L.vkm_free_obj_import(C.byref(r))


def _take_shape(o):
    v = _copy_out(o.vertices, 64 * o.vertex_count, VERTEX_DTYPE)
    ix = _copy_out(o.indices, 4 * o.index_count, np.uint32)
    lib().vkm_free_shape(C.byref(o))
    return v, ix


def generate_plane_config(width, height, xs, ys, tile_x, tile_y, deduplicate=False):
    o = ShapeOut()
    _check(lib().vkm_generate_plane_config(width, height, xs, ys, tile_x, tile_y, 1 if deduplicate else 0, C.byref(o)))
    return _take_shape(o)


def generate_cube_config(width, height, depth, tile_x, tile_y):
    o = ShapeOut()
    _check(lib().vkm_generate_cube_config(width, height, depth, tile_x, tile_y, C.byref(o)))
    return _take_shape(o)


def e3dm_write(path: str, mesh_name: str, shapes: list, names: list, material_names: list) -> None:
    """vkm_e3dm_write: the reference's WriteToE3DM file format (mesh_loader.cpp:63-126)."""
    n = len(shapes)
    cfgs = (E3dmConfig * max(n, 1))()
    keep = []
    for i, s in enumerate(shapes):
        v = np.ascontiguousarray(s.vertices)
        ix = np.ascontiguousarray(s.indices, np.uint32)
        keep += [v, ix]
        cfgs[i].shape.vertex_size, cfgs[i].shape.vertex_count, cfgs[i].shape.vertices = 64, len(v), v.ctypes.data
        cfgs[i].shape.index_size, cfgs[i].shape.index_count, cfgs[i].shape.indices = 4, len(ix), ix.ctypes.data
        cfgs[i].name = names[i].encode()
        cfgs[i].material_name = material_names[i].encode()
    _check_plain(lib().vkm_e3dm_write(path.encode(), mesh_name.encode(), cfgs, n), "vkm_e3dm_write")


def e3dm_read(path: str):
    """vkm_e3dm_read -> (mesh name, [ShapeResult], [config names], [material names])."""
    L = lib()
    cfgs = C.POINTER(E3dmConfig)()
    cnt = C.c_uint64()
    name = C.create_string_buffer(1024)
    _check_plain(L.vkm_e3dm_read(path.encode(), name, 1024, C.byref(cfgs), C.byref(cnt)), "vkm_e3dm_read")
    shapes, names, mats = [], [], []
    for i in range(cnt.value):
        s = cfgs[i].shape
        shapes.append(ShapeResult(_copy_out(s.vertices, s.vertex_count * s.vertex_size, VERTEX_DTYPE),
                                  _copy_out(s.indices, s.index_count * s.index_size, np.uint32), np.zeros(9, np.float32)))
        names.append(cfgs[i].name.decode())
        mats.append(cfgs[i].material_name.decode())
    L.vkm_e3dm_free(cfgs, cnt)
    return name.value.decode(), shapes, names, mats


def _check_plain(rc: int, what: str) -> None:
    if rc != VKM_OK:
        raise VkmError(rc, what + " failed")


class DeviceArray:
    """A device allocation made through vkm_dev_alloc (no second CUDA runtime needed)."""

    def __init__(self, device: int, host: np.ndarray | None = None, nbytes: int | None = None):
        self.device = device
        self.nbytes = int(host.nbytes if host is not None else nbytes)
        p = C.c_void_p()
        _check(lib().vkm_dev_alloc(device, self.nbytes, C.byref(p)))
        self.ptr = p.value
        if host is not None and host.nbytes:
            h = np.ascontiguousarray(host)
            _check(lib().vkm_dev_upload(device, self.ptr, h.ctypes.data, h.nbytes))

    def download(self, dtype, count: int | None = None, offset_bytes: int = 0) -> np.ndarray:
        dt = np.dtype(dtype)
        n = (self.nbytes - offset_bytes) // dt.itemsize if count is None else count
        out = np.empty(n, dt)
        if n:
            _check(lib().vkm_dev_download(self.device, out.ctypes.data, self.ptr + offset_bytes, n * dt.itemsize))
        return out

    def free(self):
        if self.ptr:
            lib().vkm_dev_free(self.device, self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def download(device: int, ptr: int, dtype, count: int) -> np.ndarray:
    dt = np.dtype(dtype)
    out = np.empty(count, dt)
    if count:
        _check(lib().vkm_dev_download(device, out.ctypes.data, ptr, count * dt.itemsize))
    return out


def last_import_stats() -> dict:
    """What the library had to do for this thread's last host-entry import (schedule, re-runs, reseeds)."""
    st = PlanStats()
    _check(lib().vkm_last_import_stats(C.byref(st)))
    return st.as_dict()


class Plan:
    """vkm_plan: device-resident import with preallocated scratch and outputs."""

    def __init__(self, device: int, max_corners: int, max_attrib: int = 0, max_vertices: int = 0, max_shapes: int = 1):
        d = PlanDesc(device, max_corners, max_vertices, max_attrib, max_shapes)
        h = C.c_void_p()
        _check(lib().vkm_plan_create(C.byref(d), C.byref(h)))
        self.h = h
        self.device = device
        self.nshapes = 0

    def set_timing(self, on: bool) -> None:
        lib().vkm_plan_set_timing(self.h, 1 if on else 0)

    def import_async(self, d_pos, np_, d_nrm, nn, d_uv, nt, d_idx, shape_offsets: np.ndarray, stream: int = 0) -> None:
        so = np.ascontiguousarray(shape_offsets, np.uint64)
        self.nshapes = len(so) - 1
        _check(lib().vkm_plan_import(self.h, d_pos, np_, d_nrm, nn, d_uv, nt, d_idx, so.ctypes.data, self.nshapes,
                                     C.c_void_p(stream)))

    def finish(self) -> PlanResult:
        r = PlanResult()
        _check(lib().vkm_plan_finish(self.h, C.byref(r)))
        return r

    def fetch(self, r: PlanResult, shape_offsets: np.ndarray) -> list:
        """Downloads the per-shape buffers of a finished import (test helper)."""
        so = np.asarray(shape_offsets, np.uint64)
        out = []
        for s in range(self.nshapes):
            v0, v1 = r.shape_vertex_offsets[s], r.shape_vertex_offsets[s + 1]
            c0, c1 = int(so[s]), int(so[s + 1])
            v = download(self.device, r.d_vertices + 64 * v0, VERTEX_DTYPE, v1 - v0)
            ix = download(self.device, r.d_indices + 4 * c0, np.uint32, c1 - c0)
            e = r.shape_extents[s]
            ext = np.array(list(e.center) + list(e.min_extents) + list(e.max_extents), np.float32)
            out.append(ShapeResult(v, ix, ext))
        return out

    def stats(self) -> dict:
        st = PlanStats()
        _check(lib().vkm_plan_get_stats(self.h, C.byref(st)))
        return st.as_dict()

    def phase_names(self):
        names, i = [], 0
        while True:
            n = lib().vkm_plan_phase_name(i)
            if not n:
                return names
            names.append(n.decode())
            i += 1

    def destroy(self):
        if self.h:
            lib().vkm_plan_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass
