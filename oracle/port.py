"""TEST INFRASTRUCTURE ONLY (oracle) -- never imported by the product package.

ctypes bindings for ``oracle/liboracle_port.so``, the plain-C restatement of the
reference's mesh-import algorithm (``oracle/oracle_port.c``).  Builds anywhere
(``make -C oracle port``), so it is the checker that is always available on the
GPU box; ``oracle/refcpu.py`` is the reference's own code and pins this one.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from .refcpu import VERTEX_DTYPE, ShapeResult

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liboracle_port.so")


class _Shape(C.Structure):
    _fields_ = [("vertex_count", C.c_uint32), ("index_count", C.c_uint32), ("vertices", C.c_void_p),
                ("indices", C.c_void_p), ("extent", C.c_float * 9)]


_lib = None


def build() -> None:
    subprocess.check_call(["make", "-C", _HERE, "port"], stdout=subprocess.DEVNULL)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = C.CDLL(LIB_PATH)
        L.orc_hash_vertex.restype = C.c_uint64
        L.orc_hash_vertex.argtypes = [C.c_void_p]
        L.orc_vertex_equal.restype = C.c_int
        L.orc_vertex_equal.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_import_shapes.restype = C.c_int
        L.orc_import_shapes.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64,
                                        C.c_void_p, C.c_void_p, C.c_uint32, C.POINTER(_Shape)]
        L.orc_free_shape.argtypes = [C.POINTER(_Shape)]
        for fn in (L.orc_generate_tangents, L.orc_generate_normals):
            fn.restype = None
            fn.argtypes = [C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p]
        L.orc_deduplicate_vertices.restype = C.c_uint32
        L.orc_deduplicate_vertices.argtypes = [C.c_uint32, C.c_void_p, C.c_uint32, C.c_void_p, C.POINTER(C.c_void_p)]
        L.orc_free.argtypes = [C.c_void_p]
        L.orc_generate_plane.restype = None
        L.orc_generate_plane.argtypes = [C.c_float, C.c_float, C.c_uint32, C.c_uint32, C.c_float, C.c_float, C.c_void_p, C.c_void_p]
        L.orc_generate_cube.restype = None
        L.orc_generate_cube.argtypes = [C.c_float, C.c_float, C.c_float, C.c_float, C.c_float, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


def import_arrays(pos, nrm, uv, idx, shape_offsets, collect: bool = True):
    L = lib()
    pos = np.ascontiguousarray(pos, np.float32)
    nrm = np.ascontiguousarray(nrm, np.float32)
    uv = np.ascontiguousarray(uv, np.float32)
    idx = np.ascontiguousarray(idx, np.int32)
    so = np.ascontiguousarray(shape_offsets, np.uint64)
    ns = len(so) - 1
    shapes = (_Shape * ns)()
    rc = L.orc_import_shapes(pos.ctypes.data, pos.size // 3, nrm.ctypes.data, nrm.size // 3, uv.ctypes.data,
                             uv.size // 2, idx.ctypes.data, so.ctypes.data, ns, shapes)
    if rc != 0:
        raise ValueError("oracle: index out of range")
    out = []
    for s in shapes:
        if collect:
            v = (np.frombuffer((C.c_char * (64 * s.vertex_count)).from_address(s.vertices), dtype=VERTEX_DTYPE).copy()
                 if s.vertex_count else np.empty(0, VERTEX_DTYPE))
            ix = (np.frombuffer((C.c_char * (4 * s.index_count)).from_address(s.indices), dtype=np.uint32).copy()
                  if s.index_count else np.empty(0, np.uint32))
            out.append(ShapeResult(v, ix, np.array(s.extent, np.float32)))
        L.orc_free_shape(C.byref(s))
    return out if collect else None


def generate_tangents(vertices: np.ndarray, indices: np.ndarray) -> None:
    assert vertices.dtype == VERTEX_DTYPE and indices.dtype == np.uint32
    lib().orc_generate_tangents(len(vertices), vertices.ctypes.data, len(indices), indices.ctypes.data)


def generate_normals(vertices: np.ndarray, indices: np.ndarray) -> None:
    assert vertices.dtype == VERTEX_DTYPE and indices.dtype == np.uint32
    lib().orc_generate_normals(len(vertices), vertices.ctypes.data, len(indices), indices.ctypes.data)


def deduplicate_vertices(vertices: np.ndarray, indices: np.ndarray):
    assert vertices.dtype == VERTEX_DTYPE and indices.dtype == np.uint32
    idx = indices.copy()
    out = C.c_void_p()
    n = lib().orc_deduplicate_vertices(len(vertices), vertices.ctypes.data, len(idx), idx.ctypes.data, C.byref(out))
    v = np.frombuffer((C.c_char * (64 * n)).from_address(out.value), dtype=VERTEX_DTYPE).copy() if n else np.empty(0, VERTEX_DTYPE)
    lib().orc_free(out)
    return v, idx


def hash_vertex(v: np.ndarray) -> int:
    return int(lib().orc_hash_vertex(v.ctypes.data))


def vertex_equal(a: np.ndarray, b: np.ndarray) -> bool:
    return bool(lib().orc_vertex_equal(a.ctypes.data, b.ctypes.data))


def generate_plane(width, height, xs, ys, tile_x, tile_y):
    xs, ys = max(xs, 1), max(ys, 1)
    v = np.zeros(4 * xs * ys, VERTEX_DTYPE)
    i = np.zeros(6 * xs * ys, np.uint32)
    lib().orc_generate_plane(width, height, xs, ys, tile_x, tile_y, v.ctypes.data, i.ctypes.data)
    return v, i


def generate_cube(width, height, depth, tile_x, tile_y):
    v = np.zeros(24, VERTEX_DTYPE)
    i = np.zeros(36, np.uint32)
    lib().orc_generate_cube(width, height, depth, tile_x, tile_y, v.ctypes.data, i.ctypes.data)
    return v, i
