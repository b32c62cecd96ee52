"""Deterministic synthetic mesh generators, emitted directly as tinyobj-shaped arrays.

The shapes are the ones BASELINE.json / SURVEY.md section 8(d) name: the regular grid family
(1M and 16M triangles, valence-6 interior = "each vertex shared by ~6 faces"), a UV sphere
with a texture seam and poles, batches of independent grid patches, and an adversarial mesh
with injected degenerate / colliding / NaN cases.  Everything is integer-seeded numpy; no
file I/O.  Layout matches tinyobj (reference: engine/src/vendor/tinyobj/tinyobj.hpp:347-398):
``pos`` f32[Np,3], ``nrm`` f32[Nn,3], ``uv`` f32[Nt,2], ``idx`` i32[C,3] in (v, n, t) order,
``shape_offsets`` u64[nshapes+1] in corners.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np


@dataclass
class MeshStreams:
    pos: np.ndarray
    nrm: np.ndarray
    uv: np.ndarray
    idx: np.ndarray
    shape_offsets: np.ndarray
    meta: dict = field(default_factory=dict)

    @property
    def triangles(self) -> int:
        return int(self.idx.shape[0] // 3)

    @property
    def corners(self) -> int:
        return int(self.idx.shape[0])

    def input_bytes(self) -> int:
        return self.pos.nbytes + self.nrm.nbytes + self.uv.nbytes + self.idx.nbytes

    def algorithmic_bytes(self, unique_vertices: int) -> int:
        """36F + 12Np + 12Nn + 8Nt + 12F + 64U + 36 per shape."""
        f = self.triangles
        ns = len(self.shape_offsets) - 1
        return 36 * f + self.pos.nbytes + self.nrm.nbytes + self.uv.nbytes + 12 * f + 64 * unique_vertices + 36 * ns


def _grid_attributes(n: int, offset=(0.0, 0.0, 0.0), scale: float = 1.0):
    m = n + 1
    y, x = np.divmod(np.arange(m * m, dtype=np.int64), m)
    fx = (x / n).astype(np.float32)
    fy = (y / n).astype(np.float32)
    z = (np.float32(0.1) * (((7 * x + 13 * y) % 17).astype(np.float32) / np.float32(17))).astype(np.float32)
    pos = np.stack([fx, fy, z], axis=1).astype(np.float32)
    if scale != 1.0 or any(offset):
        pos = (pos * np.float32(scale) + np.asarray(offset, np.float32)).astype(np.float32)
    uv = np.stack([fx, fy], axis=1).astype(np.float32)
    hx = ((((3 * x + 5 * y) % 11) - 5).astype(np.float32) / np.float32(50)).astype(np.float32)
    hy = ((((5 * x + 2 * y) % 13) - 6).astype(np.float32) / np.float32(60)).astype(np.float32)
    nz = np.ones_like(hx)
    ln = np.sqrt(hx * hx + hy * hy + nz * nz).astype(np.float32)
    nrm = np.stack([hx / ln, hy / ln, nz / ln], axis=1).astype(np.float32)
    return pos, nrm, uv


def _grid_faces(n: int) -> np.ndarray:
    """Vertex ids [F,3] of the 2*n*n triangles, row-major quads, each quad -> (a,b,d), (a,d,c)."""
    m = n + 1
    qy, qx = np.divmod(np.arange(n * n, dtype=np.int64), n)
    a = qy * m + qx
    b = a + 1
    c = a + m
    d = c + 1
    tri = np.empty((n * n, 2, 3), dtype=np.int32)
    tri[:, 0, 0] = a
    tri[:, 0, 1] = b
    tri[:, 0, 2] = d
    tri[:, 1, 0] = a
    tri[:, 1, 1] = d
    tri[:, 1, 2] = c
    return tri.reshape(-1, 3)


def grid(n: int, permute_seed: int | None = None, offset=(0.0, 0.0, 0.0), scale: float = 1.0) -> MeshStreams:
    """n x n quads: F = 2 n^2 triangles, U = (n+1)^2 vertices, Np = Nn = Nt = U.

    n = 708 -> 1 002 528 triangles (config 2);  n = 2896 -> 16 773 632 (config 3);
    n = 5793 -> 67 117 698 (config 5 base).
    """
    pos, nrm, uv = _grid_attributes(n, offset, scale)
    faces = _grid_faces(n)
    if permute_seed is not None:
        rng = np.random.Generator(np.random.PCG64(permute_seed))
        faces = faces[rng.permutation(faces.shape[0])]
    flat = faces.reshape(-1)
    idx = np.empty((flat.shape[0], 3), dtype=np.int32)
    idx[:, 0] = flat
    idx[:, 1] = flat
    idx[:, 2] = flat
    so = np.array([0, idx.shape[0]], dtype=np.uint64)
    return MeshStreams(pos, nrm, uv, idx, so, {"kind": "grid", "n": n, "unique": (n + 1) ** 2,
                                                "permute_seed": permute_seed})


def uv_sphere(rings: int, segments: int) -> MeshStreams:
    """Lat/long sphere.  Positions are shared across the texture seam and at the poles while uv
    records are not (Nt > Np), so many position indices carry several (n, t) combinations; pole
    triangles have two corners with identical position (zero-area in uv or in space)."""
    rr, ss = rings, segments
    # positions: (rings+1) latitudes x segments longitudes, poles collapsed to one record each
    lat = np.arange(1, rr, dtype=np.float64) / rr * np.pi
    lon = np.arange(ss, dtype=np.float64) / ss * 2 * np.pi
    body = np.stack([np.outer(np.sin(lat), np.cos(lon)), np.outer(np.cos(lat), np.ones(ss)),
                     np.outer(np.sin(lat), np.sin(lon))], axis=2).reshape(-1, 3)
    pos = np.concatenate([[[0.0, 1.0, 0.0]], body, [[0.0, -1.0, 0.0]]]).astype(np.float32)
    nrm = pos.copy()  # unit sphere: normal == position, one per position record
    # uv: (rings+1) x (segments+1) grid (seam column duplicated)
    vv, uu = np.divmod(np.arange((rr + 1) * (ss + 1)), ss + 1)
    uv = np.stack([uu / ss, 1.0 - vv / rr], axis=1).astype(np.float32)

    def pid(r, s):  # position record of lattice point (ring r, segment s mod ss)
        r = np.asarray(r)
        s = np.asarray(s) % ss
        return np.where(r == 0, 0, np.where(r == rr, 1 + (rr - 1) * ss, 1 + (r - 1) * ss + s))

    def tid(r, s):
        return np.asarray(r) * (ss + 1) + np.asarray(s)

    r, s = np.divmod(np.arange(rr * ss), ss)
    quads_p = np.stack([pid(r, s), pid(r, s + 1), pid(r + 1, s), pid(r + 1, s + 1)], axis=1)
    quads_t = np.stack([tid(r, s), tid(r, s + 1), tid(r + 1, s), tid(r + 1, s + 1)], axis=1)
    order = np.array([[0, 1, 3], [0, 3, 2]])
    fp = quads_p[:, order].reshape(-1, 3)
    ft = quads_t[:, order].reshape(-1, 3)
    idx = np.empty((fp.size, 3), dtype=np.int32)
    idx[:, 0] = fp.reshape(-1)
    idx[:, 1] = fp.reshape(-1)
    idx[:, 2] = ft.reshape(-1)
    so = np.array([0, idx.shape[0]], dtype=np.uint64)
    return MeshStreams(pos, nrm, uv, idx, so, {"kind": "uv_sphere", "rings": rr, "segments": ss})


def concat(meshes: list) -> MeshStreams:
    """Concatenate meshes into one multi-shape stream (one shape per input shape).  Attribute indices
    are rebased so that every mesh keeps its own attribute records."""
    pos, nrm, uv, idx, offs = [], [], [], [], [0]
    bp = bn = bt = 0
    corner_base = 0
    for m in meshes:
        pos.append(m.pos)
        nrm.append(m.nrm)
        uv.append(m.uv)
        idx.append((m.idx.astype(np.int64) + np.array([bp, bn, bt])).astype(np.int32))
        offs.extend(corner_base + int(o) for o in m.shape_offsets[1:])
        corner_base += m.idx.shape[0]
        bp += m.pos.shape[0]
        bn += m.nrm.shape[0]
        bt += m.uv.shape[0]
    return MeshStreams(np.concatenate(pos), np.concatenate(nrm), np.concatenate(uv), np.concatenate(idx),
                       np.asarray(offs, np.uint64), {"kind": "concat", "meshes": len(meshes)})


def batch(count: int, seed: int = 0x5EED0004, min_tris: int = 1000, max_tris: int = 100_000) -> list:
    """``count`` independent grid patches with log-uniform triangle counts in [min_tris, max_tris]
    (config 4).  Returned as a list of single-shape meshes; ``concat`` packs a sub-batch."""
    rng = np.random.Generator(np.random.PCG64(seed))
    logs = rng.uniform(np.log(min_tris), np.log(max_tris), size=count)
    out = []
    for lg in logs:
        n = max(2, int(round(np.sqrt(np.exp(lg) / 2.0))))
        off = rng.uniform(-10.0, 10.0, size=3)
        sc = float(rng.uniform(0.5, 4.0))
        out.append(grid(n, offset=tuple(float(v) for v in off), scale=sc))
    return out


def adversarial(n: int, seed: int = 0x5EED0005, frac: float = 1e-3, nan_frac: float = 1e-4) -> MeshStreams:
    """Grid mesh with seeded injections (config 5):
      * zero-area triangles (a corner repeated) and zero-uv-area triangles,
      * duplicated v / vn / vt records (different index triplets, identical values -> must merge),
      * +0 / -0 twins (must merge; the stored vertex keeps the first-seen bit pattern),
      * 1-ulp-apart twins (must NOT merge),
      * NaN attributes (every occurrence appends a vertex and writes index 0).
    """
    rng = np.random.Generator(np.random.PCG64(seed))
    m = grid(n)
    pos, nrm, uv, idx = m.pos.copy(), m.nrm.copy(), m.uv.copy(), m.idx.copy()
    u = pos.shape[0]
    f = idx.shape[0] // 3
    k = max(4, int(f * frac))

    # duplicated attribute records appended at the end; some corners are redirected to the copies
    def dup(arr, col):
        src = rng.integers(0, u, size=k)
        base = arr.shape[0]
        arr2 = np.concatenate([arr, arr[src]])
        corners = rng.integers(0, idx.shape[0], size=4 * k)
        hit = np.isin(idx[corners, col], src)
        cc = corners[hit]
        lookup = {int(s): base + j for j, s in enumerate(src)}
        idx[cc, col] = np.array([lookup[int(v)] for v in idx[cc, col]], dtype=np.int32) if cc.size else idx[cc, col]
        return arr2

    pos = dup(pos, 0)
    nrm = dup(nrm, 1)
    uv = dup(uv, 2)

    # +0 / -0 twins: copy of a position record whose zero components are negated (x == 0 column of the grid)
    zsrc = np.arange(0, (n + 1) * (n + 1), n + 1)[: max(2, k // 4)]  # x == 0 -> position.x == +0
    ztw = pos[zsrc].copy()
    ztw[:, 0] = np.float32(-0.0)
    zbase = pos.shape[0]
    pos = np.concatenate([pos, ztw])
    zmap = {int(s): zbase + j for j, s in enumerate(zsrc)}
    cand = np.nonzero(np.isin(idx[:, 0], zsrc))[0]
    pick = cand[rng.random(cand.size) < 0.5]
    idx[pick, 0] = np.array([zmap[int(v)] for v in idx[pick, 0]], dtype=np.int32) if pick.size else idx[pick, 0]

    # 1-ulp twins: copies with uv.x bumped by one ulp
    usrc = rng.integers(0, u, size=k)
    utw = uv[usrc].copy()
    utw[:, 0] = np.nextafter(utw[:, 0], np.float32(2.0)).astype(np.float32)
    ubase = uv.shape[0]
    uv = np.concatenate([uv, utw])
    umap = {int(s): ubase + j for j, s in enumerate(usrc)}
    cand = np.nonzero(np.isin(idx[:, 2], usrc))[0]
    pick = cand[rng.random(cand.size) < 0.5]
    idx[pick, 2] = np.array([umap[int(v)] for v in idx[pick, 2]], dtype=np.int32) if pick.size else idx[pick, 2]

    # zero-area triangles: corner 1 := corner 0 ; zero-uv-area: all three uv indices equal
    tri = rng.integers(0, f, size=k)
    idx[3 * tri + 1] = idx[3 * tri]
    tri = rng.integers(0, f, size=k)
    idx[3 * tri + 1, 2] = idx[3 * tri, 2]
    idx[3 * tri + 2, 2] = idx[3 * tri, 2]

    # NaN attributes
    kn = max(2, int(u * nan_frac))
    pos[rng.integers(0, u, size=kn), rng.integers(0, 3, size=kn)] = np.float32(np.nan)
    uv[rng.integers(0, u, size=kn), 1] = np.float32(np.nan)
    nrm[rng.integers(0, u, size=max(1, kn // 2)), 2] = np.float32(np.nan)

    so = np.array([0, idx.shape[0]], dtype=np.uint64)
    return MeshStreams(pos.astype(np.float32), nrm.astype(np.float32), uv.astype(np.float32), idx, so,
                       {"kind": "adversarial", "n": n, "seed": seed})


def with_duplicate_record(m: MeshStreams) -> MeshStreams:
    """The same mesh made "not clean" by ONE duplicated position record: the last corner is redirected to a copy of
    its own position, so that two distinct (v, n, t) triplets carry one value and must merge.  Unique vertices, indices
    and sizes stay what they were; only the schedule the importer needs changes (bench.py: full_schedule, mis_speculated)."""
    pos = np.concatenate([m.pos, m.pos[m.idx[-1, 0]][None, :]])
    idx = m.idx.copy()
    idx[-1, 0] = len(pos) - 1
    return MeshStreams(pos, m.nrm, m.uv, idx, m.shape_offsets, dict(m.meta, duplicate_record=True))


def to_obj_text(m: MeshStreams) -> str:
    """OBJ text of a mesh (one `o` per shape), for driving the unmodified reference LoadOBJ in tests.
    Floats are printed with 9 significant digits so that text -> float32 round-trips exactly."""
    lines = []
    for p in m.pos:
        lines.append("v %.9g %.9g %.9g" % (p[0], p[1], p[2]))
    for t in m.uv:
        lines.append("vt %.9g %.9g" % (t[0], t[1]))
    for n in m.nrm:
        lines.append("vn %.9g %.9g %.9g" % (n[0], n[1], n[2]))
    for s in range(len(m.shape_offsets) - 1):
        lines.append("o shape%d" % s)
        c0, c1 = int(m.shape_offsets[s]), int(m.shape_offsets[s + 1])
        tri = m.idx[c0:c1].reshape(-1, 3, 3) + 1
        for t in tri:
            lines.append("f %d/%d/%d %d/%d/%d %d/%d/%d" % (t[0][0], t[0][2], t[0][1], t[1][0], t[1][2], t[1][1],
                                                        t[2][0], t[2][2], t[2][1]))
    return "\n".join(lines) + "\n"
