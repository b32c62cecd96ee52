#!/usr/bin/env python
"""Regenerates the golden fixtures in this directory.  Runs only in the build container, where
/root/reference and oracle/_ref/libvkref.so (the reference's own sources, compiled unmodified by
oracle/Makefile) exist.  The fixtures travel to the GPU box; the reference tree does not.

  falcon.npz       inputs : tinyobj streams of assets/models/falcon.obj (parsed by the reference's tinyobj)
                   outputs: the buffers of the reference-SHIPPED known-answer file assets/models/falcon.e3dm
                            (read with the reference's LoadE3DM) + the extent from the reference's LoadOBJ
                            + sha256 / size of falcon.e3dm itself (for the E3DM writer byte-compare).
  viking_room.npz  inputs : tinyobj streams of assets/models/viking_room.obj
                   outputs: reference LoadOBJ run here (no golden is shipped for this asset).
  cube.npz         the 24-vertex / 36-index cube of GeometrySystem::GenerateCubeConfig
                   (geometry_system.cpp:296-439) after the reference's GenerateNormals + GenerateTangents.
  asset_text.npz   the bytes of falcon.obj / falcon.mtl / viking_room.obj themselves (reference assets; viking_room's
                   .mtl is not shipped), so that the GPU box can run OBJ TEXT -> E3DM without the reference tree.
  semantics.npz    tiny hand-built edge cases (NaN keys, +-0 twins, 1-ulp twins, degenerate triangles)
                   through the reference import loop.
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import refcpu  # noqa: E402

MODELS = "/root/reference/assets/models/"
HERE = os.path.dirname(os.path.abspath(__file__))


def asset(name, golden_e3dm=None):
    s = refcpu.parse_obj(MODELS + name + ".obj")
    ref = refcpu.load_obj(MODELS + name + ".obj", name)
    assert len(ref) == len(s.shape_names)
    out = dict(pos=s.pos, nrm=s.nrm, uv=s.uv, idx=s.idx, shape_offsets=s.shape_offsets,
               shape_names=np.array(s.shape_names), material_names=np.array(s.material_names), nshapes=len(ref))
    if golden_e3dm:
        g = refcpu.load_e3dm(golden_e3dm)
        raw = open(golden_e3dm, "rb").read()
        out["e3dm_sha256"] = np.array(hashlib.sha256(raw).hexdigest())
        out["e3dm_size"] = np.array(len(raw))
        out["e3dm_name"] = np.array(name)
        for i, (a, b) in enumerate(zip(ref, g)):  # the shipped golden must equal the reference run here
            assert a.vertices.tobytes() == b.vertices.tobytes() and np.array_equal(a.indices, b.indices), "golden drift"
            assert a.name == b.name and a.material_name == b.material_name
    for i, r in enumerate(ref):
        out[f"vertices{i}"] = r.vertices.view(np.uint8).reshape(-1, 64)
        out[f"indices{i}"] = r.indices
        out[f"extent{i}"] = r.extent
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, [(len(r.vertices), len(r.indices)) for r in ref])


def cube():
    # Restatement of the cube vertex table: the six faces of GenerateCubeConfig, width=height=depth=2, tile 1.
    hw = hh = hd = 1.0
    mn, mx = 0.0, 1.0
    faces = [  # (min corner, max corner pattern) as four positions per face in the reference's order
        [(-hw, -hh, hd), (hw, hh, hd), (-hw, hh, hd), (hw, -hh, hd)],      # front
        [(hw, -hh, -hd), (-hw, hh, -hd), (hw, hh, -hd), (-hw, -hh, -hd)],  # back
        [(-hw, -hh, -hd), (-hw, hh, hd), (-hw, hh, -hd), (-hw, -hh, hd)],  # left
        [(hw, -hh, hd), (hw, hh, -hd), (hw, hh, hd), (hw, -hh, -hd)],      # right
        [(hw, -hh, hd), (-hw, -hh, -hd), (hw, -hh, -hd), (-hw, -hh, hd)],  # bottom
        [(-hw, hh, hd), (hw, hh, -hd), (-hw, hh, -hd), (hw, hh, hd)],      # top
    ]
    uvs = [(mn, mn), (mx, mx), (mn, mx), (mx, mn)]
    v = np.zeros(24, refcpu.VERTEX_DTYPE)
    idx = np.zeros(36, np.uint32)
    for f, quad in enumerate(faces):
        for k in range(4):
            v["position"][4 * f + k] = quad[k]
            v["texcoord"][4 * f + k] = uvs[k]
        o = 4 * f
        idx[6 * f: 6 * f + 6] = [o + 0, o + 1, o + 2, o + 0, o + 3, o + 1]
    inp = v.copy()
    refcpu.generate_normals(v, idx)
    refcpu.generate_tangents(v, idx)
    np.savez_compressed(os.path.join(HERE, "cube.npz"), vertices_in=inp.view(np.uint8).reshape(-1, 64), indices=idx,
                        vertices_out=v.view(np.uint8).reshape(-1, 64))
    print("cube normals", v["normal"][::4].tolist())


def semantics():
    nan = np.float32(np.nan)
    one = np.float32(1.0)
    pos = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [nan, 0, 0], [-0.0, 0, 0], [np.nextafter(one, np.float32(2)), 0, 0],
                    [1, 0, 0], [0, 0, 1], [1e-9, 0, 0], [2e-9, 0, 0], [np.inf, 0, 0], [-1, -2, -3]], np.float32)
    nrm = np.array([[0, 0, 1], [0, 0, -1], [0, -0.0, 1]], np.float32)
    uv = np.array([[0, 0], [1, 0], [0, 1], [1, 1], [0, nan], [0.5, 0.5]], np.float32)
    tri = [
        [(0, 0, 0), (1, 0, 1), (2, 0, 2)],      # plain
        [(0, 0, 0), (3, 0, 1), (3, 0, 1)],      # NaN position twice
        [(4, 0, 0), (1, 0, 1), (2, 0, 2)],      # -0 twin of corner 0 -> merges
        [(0, 2, 0), (5, 0, 1), (6, 0, 1)],      # -0 normal twin merges; 1-ulp twin does not; duplicate v record merges
        [(0, 0, 0), (0, 0, 0), (1, 0, 1)],      # zero-area
        [(0, 0, 5), (1, 0, 5), (2, 0, 5)],      # zero-uv-area
        [(8, 0, 0), (9, 0, 0), (7, 0, 3)],      # denormal-ish neighbours: epsilon-close but different hash
        [(2, 1, 2), (1, 1, 1), (0, 1, 0)],      # mirrored winding / other normal
        [(0, 0, 4), (1, 0, 4), (2, 0, 3)],      # NaN uv
        [(10, 0, 0), (11, 1, 3), (0, 0, 0)],    # inf, negative extents
    ]
    idx = np.array([c for t in tri for c in t], np.int32)
    idx = idx[:, [0, 1, 2]]  # (v, n, t)
    so = np.array([0, 15, len(idx)], np.uint64)  # two shapes: dedup state resets between them
    ref = refcpu.import_arrays(pos, nrm, uv, idx, so)
    out = dict(pos=pos, nrm=nrm, uv=uv, idx=idx, shape_offsets=so, nshapes=len(ref))
    for i, r in enumerate(ref):
        out[f"vertices{i}"] = r.vertices.view(np.uint8).reshape(-1, 64)
        out[f"indices{i}"] = r.indices
        out[f"extent{i}"] = r.extent
        print("semantics shape", i, len(r.vertices), r.indices.tolist())
    np.savez_compressed(os.path.join(HERE, "semantics.npz"), **out)


def asset_text():
    out = {}
    for f in ("falcon.obj", "falcon.mtl", "viking_room.obj"):
        out[f.replace(".", "_")] = np.frombuffer(open(MODELS + f, "rb").read(), np.uint8)
    np.savez_compressed(os.path.join(HERE, "asset_text.npz"), **out)
    print("asset_text", {k: len(v) for k, v in out.items()})


if __name__ == "__main__":
    assert refcpu.available(), "build oracle/_ref first: make -C oracle ref"
    asset("falcon", MODELS + "falcon.e3dm")
    asset("viking_room")
    cube()
    semantics()
    asset_text()
