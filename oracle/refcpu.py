"""TEST INFRASTRUCTURE ONLY (oracle) -- never imported by the product package.

ctypes bindings for ``oracle/_ref/libvkref.so``: the reference's own mesh-import
sources (mesh_loader.cpp, geometry.cpp, tinyobj, ...) compiled unmodified by
``oracle/Makefile``.  Used by tests, by ``tests/golden/make_golden.py`` and by
``bench.py``'s cpu_baseline / ``--impl reference`` legs.  The library is built
in the container that has ``/root/reference``; the prebuilt ``.so`` travels to
the GPU box, the reference tree itself does not.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libvkref.so")

VERTEX_DTYPE = np.dtype(
    [("position", "<f4", 3), ("normal", "<f4", 3), ("texcoord", "<f4", 2), ("color", "<f4", 4), ("tangent", "<f4", 4)]
)
assert VERTEX_DTYPE.itemsize == 64

_lib = None


def available() -> bool:
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(LIB_PATH)
        L.ref_parse_obj.restype = C.c_void_p
        L.ref_parse_obj.argtypes = [C.c_char_p, C.c_char_p]
        L.ref_parsed_sizes.argtypes = [C.c_void_p, C.POINTER(C.c_uint64)]
        L.ref_parsed_attribs.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_parsed_shape_info.restype = C.c_uint64
        L.ref_parsed_shape_info.argtypes = [C.c_void_p, C.c_uint32, C.c_char_p, C.c_int, C.POINTER(C.c_int), C.c_char_p, C.c_int]
        L.ref_parsed_shape_indices.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p]
        L.ref_parsed_free.argtypes = [C.c_void_p]
        L.ref_load_obj.restype = C.c_void_p
        L.ref_load_obj.argtypes = [C.c_char_p, C.c_char_p]
        L.ref_import_arrays.restype = C.c_void_p
        L.ref_import_arrays.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64,
                                        C.c_void_p, C.c_void_p, C.c_uint32]
        L.ref_mesh_config_count.restype = C.c_uint64
        L.ref_mesh_config_count.argtypes = [C.c_void_p]
        L.ref_mesh_config.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32),
                                      C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.c_void_p, C.c_char_p, C.c_int,
                                      C.c_char_p, C.c_int]
        L.ref_mesh_free.argtypes = [C.c_void_p]
        L.ref_write_e3dm.restype = C.c_int
        L.ref_write_e3dm.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p]
        L.ref_load_e3dm.restype = C.c_void_p
        L.ref_load_e3dm.argtypes = [C.c_char_p, C.c_char_p]
        for fn in (L.ref_generate_tangents, L.ref_generate_normals):
            fn.restype = None
            fn.argtypes = [C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p]
        L.ref_deduplicate_vertices.restype = C.c_uint32
        L.ref_deduplicate_vertices.argtypes = [C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p, C.POINTER(C.c_void_p)]
        L.ref_free.argtypes = [C.c_void_p]
        L.ref_hash_vertex.restype = C.c_uint64
        L.ref_hash_vertex.argtypes = [C.c_void_p]
        L.ref_vertex_equal.restype = C.c_int
        L.ref_vertex_equal.argtypes = [C.c_void_p, C.c_void_p]
        L.ref_sizeof_vertex.restype = C.c_uint32
        L.ref_generate_plane.restype = None
        L.ref_generate_plane.argtypes = [C.c_float, C.c_float, C.c_uint32, C.c_uint32, C.c_float, C.c_float,
                                         C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
        L.ref_generate_cube.restype = None
        L.ref_generate_cube.argtypes = [C.c_float, C.c_float, C.c_float, C.c_float, C.c_float,
                                        C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
        _lib = L
    return _lib


@dataclass
class ShapeResult:
    vertices: np.ndarray  # VERTEX_DTYPE [U]
    indices: np.ndarray  # uint32 [C]
    extent: np.ndarray  # float32 [9] = center, min, max
    name: str = ""
    material_name: str = ""


@dataclass
class ObjStreams:
    """tinyobj-shaped arrays: the input of the hot path."""

    pos: np.ndarray  # float32 [Np,3]
    nrm: np.ndarray  # float32 [Nn,3]
    uv: np.ndarray  # float32 [Nt,2]
    idx: np.ndarray  # int32 [C,3]  (v, n, t) -- tinyobj::index_t order
    shape_offsets: np.ndarray  # uint64 [nshapes+1], corner offsets
    shape_names: list
    material_names: list


def parse_obj(path: str) -> ObjStreams:
    L = lib()
    h = L.ref_parse_obj(path.encode(), (os.path.dirname(path) + "/").encode())
    if not h:
        raise RuntimeError(f"tinyobj failed on {path}")
    try:
        sz = (C.c_uint64 * 4)()
        L.ref_parsed_sizes(h, sz)
        np_, nn, nt, ns = (int(x) for x in sz)
        pos = np.empty((np_, 3), np.float32)
        nrm = np.empty((nn, 3), np.float32)
        uv = np.empty((nt, 2), np.float32)
        L.ref_parsed_attribs(h, pos.ctypes.data, nrm.ctypes.data, uv.ctypes.data)
        offs = [0]
        chunks, names, mats = [], [], []
        for i in range(ns):
            nb, mb, m0 = C.create_string_buffer(512), C.create_string_buffer(512), C.c_int(0)
            nc = int(L.ref_parsed_shape_info(h, i, nb, 512, C.byref(m0), mb, 512))
            a = np.empty((nc, 3), np.int32)
            L.ref_parsed_shape_indices(h, i, a.ctypes.data)
            chunks.append(a)
            offs.append(offs[-1] + nc)
            names.append(nb.value.decode())
            mats.append(mb.value.decode())
        idx = np.concatenate(chunks) if chunks else np.empty((0, 3), np.int32)
        return ObjStreams(pos, nrm, uv, idx, np.asarray(offs, np.uint64), names, mats)
    finally:
        L.ref_parsed_free(h)


def _collect(h) -> list:
    L = lib()
    out = []
    for i in range(int(L.ref_mesh_config_count(h))):
        vc, ic = C.c_uint32(), C.c_uint32()
        vp, ip = C.c_void_p(), C.c_void_p()
        ext = np.zeros(9, np.float32)
        nb, mb = C.create_string_buffer(512), C.create_string_buffer(512)
        L.ref_mesh_config(h, i, C.byref(vc), C.byref(ic), C.byref(vp), C.byref(ip), ext.ctypes.data, nb, 512, mb, 512)
        if vc.value:
            v = np.frombuffer((C.c_char * (64 * vc.value)).from_address(vp.value), dtype=VERTEX_DTYPE).copy()
        else:
            v = np.empty(0, VERTEX_DTYPE)
        if ic.value:
            ix = np.frombuffer((C.c_char * (4 * ic.value)).from_address(ip.value), dtype=np.uint32).copy()
        else:
            ix = np.empty(0, np.uint32)
        out.append(ShapeResult(v, ix, ext, nb.value.decode(), mb.value.decode()))
    return out


def load_obj(path: str, name: str = "mesh") -> list:
    """Unmodified MeshLoader::LoadOBJ (mesh_loader.cpp:200-328) on an OBJ file."""
    L = lib()
    h = L.ref_load_obj(path.encode(), name.encode())
    if not h:
        raise RuntimeError(f"reference LoadOBJ failed on {path}")
    try:
        return _collect(h)
    finally:
        L.ref_mesh_free(h)


def import_arrays(pos, nrm, uv, idx, shape_offsets, collect: bool = True):
    """Unmodified MeshLoader::LoadOBJ per-shape loop on in-memory tinyobj-shaped arrays."""
    L = lib()
    pos = np.ascontiguousarray(pos, np.float32)
    nrm = np.ascontiguousarray(nrm, np.float32)
    uv = np.ascontiguousarray(uv, np.float32)
    idx = np.ascontiguousarray(idx, np.int32)
    so = np.ascontiguousarray(shape_offsets, np.uint64)
    h = L.ref_import_arrays(pos.ctypes.data, pos.size // 3, nrm.ctypes.data, nrm.size // 3, uv.ctypes.data,
                            uv.size // 2, idx.ctypes.data, so.ctypes.data, len(so) - 1)
    if not h:
        raise RuntimeError("reference import failed")
    try:
        return _collect(h) if collect else None
    finally:
        L.ref_mesh_free(h)


def load_e3dm(path: str) -> list:
    L = lib()
    h = L.ref_load_e3dm(path.encode(), b"mesh")
    if not h:
        raise RuntimeError(f"reference LoadE3DM failed on {path}")
    try:
        return _collect(h)
    finally:
        L.ref_mesh_free(h)


def obj_to_e3dm(obj_path: str, out_path: str, name: str) -> None:
    """LoadOBJ followed by the reference's WriteToE3DM (mesh_loader.cpp:63-126)."""
    L = lib()
    h = L.ref_load_obj(obj_path.encode(), name.encode())
    if not h:
        raise RuntimeError("LoadOBJ failed")
    try:
        if not L.ref_write_e3dm(h, out_path.encode(), name.encode()):
            raise RuntimeError("WriteToE3DM failed")
    finally:
        L.ref_mesh_free(h)


def generate_tangents(vertices: np.ndarray, indices: np.ndarray) -> None:
    assert vertices.dtype == VERTEX_DTYPE and indices.dtype == np.uint32
    lib().ref_generate_tangents(len(vertices), vertices.ctypes.data, len(indices), indices.ctypes.data)


def generate_normals(vertices: np.ndarray, indices: np.ndarray) -> None:
    assert vertices.dtype == VERTEX_DTYPE and indices.dtype == np.uint32
    lib().ref_generate_normals(len(vertices), vertices.ctypes.data, len(indices), indices.ctypes.data)


def deduplicate_vertices(vertices: np.ndarray, indices: np.ndarray):
    """Returns (unique_vertices, rewritten_indices) per geometry.cpp:94-126 (indices rewritten in place there)."""
    assert vertices.dtype == VERTEX_DTYPE and indices.dtype == np.uint32
    idx = indices.copy()
    out = C.c_void_p()
    n = lib().ref_deduplicate_vertices(len(vertices), vertices.ctypes.data, len(idx), idx.ctypes.data, C.byref(out))
    v = np.frombuffer((C.c_char * (64 * n)).from_address(out.value), dtype=VERTEX_DTYPE).copy() if n else np.empty(0, VERTEX_DTYPE)
    lib().ref_free(out)
    return v, idx


def hash_vertex(v: np.ndarray) -> int:
    return int(lib().ref_hash_vertex(v.ctypes.data))


def vertex_equal(a: np.ndarray, b: np.ndarray) -> bool:
    return bool(lib().ref_vertex_equal(a.ctypes.data, b.ctypes.data))


def _take(vc, ic, vp, ip):
    v = np.frombuffer((C.c_char * (64 * vc.value)).from_address(vp.value), dtype=VERTEX_DTYPE).copy()
    ix = np.frombuffer((C.c_char * (4 * ic.value)).from_address(ip.value), dtype=np.uint32).copy()
    lib().ref_free(vp)
    lib().ref_free(ip)
    return v, ix


def generate_plane(width, height, xs, ys, tile_x, tile_y):
    """GeometrySystem::GeneratePlainConfig (geometry_system.cpp:182-294) -> (vertices, indices)."""
    vc, ic, vp, ip = C.c_uint32(), C.c_uint32(), C.c_void_p(), C.c_void_p()
    lib().ref_generate_plane(width, height, xs, ys, tile_x, tile_y, C.byref(vc), C.byref(ic), C.byref(vp), C.byref(ip))
    return _take(vc, ic, vp, ip)


def generate_cube(width, height, depth, tile_x, tile_y):
    """GeometrySystem::GenerateCubeConfig (geometry_system.cpp:296-439) -> (vertices, indices)."""
    vc, ic, vp, ip = C.c_uint32(), C.c_uint32(), C.c_void_p(), C.c_void_p()
    lib().ref_generate_cube(width, height, depth, tile_x, tile_y, C.byref(vc), C.byref(ic), C.byref(vp), C.byref(ip))
    return _take(vc, ic, vp, ip)


class Freelist:
    """The reference's own first-fit allocator (core/utils/freelist.cpp) behind ctypes.  NOTE: its FreeBlock merges
    neighbours with dangling pointers (freelist.cpp:36-72); only allocation-only and last-in-first-out release
    sequences are safe to run (random release orders crash inside FreeByOffset after a few dozen steps)."""

    def __init__(self, total: int):
        L = lib()
        L.ref_freelist_create.restype = C.c_void_p
        L.ref_freelist_create.argtypes = [C.c_uint64]
        L.ref_freelist_destroy.argtypes = [C.c_void_p]
        L.ref_freelist_alloc.restype = C.c_uint64
        L.ref_freelist_alloc.argtypes = [C.c_void_p, C.c_uint64]
        L.ref_freelist_free.argtypes = [C.c_void_p, C.c_uint64]
        L.ref_freelist_free_space.restype = C.c_uint64
        L.ref_freelist_free_space.argtypes = [C.c_void_p]
        self.L, self.h = L, L.ref_freelist_create(total)

    def alloc(self, size: int) -> int:
        return int(self.L.ref_freelist_alloc(self.h, size))

    def free(self, offset: int) -> bool:
        return bool(self.L.ref_freelist_free(self.h, offset))

    def free_space(self) -> int:
        return int(self.L.ref_freelist_free_space(self.h))

    def close(self):
        if self.h:
            self.L.ref_freelist_destroy(self.h)
            self.h = None
