"""Key sets engineered to collide in the BUILD'S OWN table hashes.

Test-input generation only: a numpy restatement of the device hash functions of csrc/vkm_device.cuh (`mix64`,
`hash_key8`, `hash_triplet`, hash generation 0) and searches for

  * value clusters   -- many distinct 8-float keys whose low 32 hash bits fall in a narrow window, so that for ANY table
                        size they land in a handful of neighbouring slots / buckets: linear-probe runs beyond
                        `kMaxProbe`, bucket chains beyond the fingerprint table's limit -> overflow -> re-run in
                        the next hash generation (vkm_plan_stats.reseeds);
  * triplet clusters -- many distinct (v, n, t) triplets that share one key record (so all but one go to the side
                        table / probe on in the general table) AND cluster in `hash_triplet`;
  * fingerprint twins -- two DIFFERENT value keys with equal high 32 hash bits in one bucket of the record schedule's
                        fingerprint table: a false "maybe duplicate" that must cost a re-run, never a merge.

`inject` appends the records and a few triangles that use them to a mesh.  The expected output is whatever the
reference does with the same arrays (the values are ordinary finite floats): collisions must not change results.
"""
from __future__ import annotations

import numpy as np

from .synth import MeshStreams

M64 = np.uint64(0xFFFFFFFFFFFFFFFF)
K_MAX_PROBE = 2048  # csrc/vkm_device.cuh kMaxProbe


def mix64(x: np.ndarray) -> np.ndarray:
    x = x.astype(np.uint64, copy=True)
    with np.errstate(over="ignore"):
        x ^= x >> np.uint64(33)
        x *= np.uint64(0xFF51AFD7ED558CCD)
        x ^= x >> np.uint64(33)
        x *= np.uint64(0xC4CEB9FE1A85EC53)
        x ^= x >> np.uint64(33)
    return x


def canon_bits(a: np.ndarray) -> np.ndarray:
    b = np.ascontiguousarray(a, np.float32).view(np.uint32).copy()
    b[b == 0x80000000] = 0
    return b


def hash_key8(k: np.ndarray, shape: int = 0, seed: int = 0) -> np.ndarray:
    """k: uint32[N, 8] canonical bits (pos xyz, nrm xyz, uv xy)."""
    k = k.astype(np.uint64)
    with np.errstate(over="ignore"):
        h = mix64((k[:, 0] | (k[:, 1] << np.uint64(32))) + np.uint64(0x9E3779B97F4A7C15) * np.uint64(shape + 1)
                  + np.uint64(0xD6E8FEB86659FD93) * np.uint64(seed))
        h = mix64(h ^ (k[:, 2] | (k[:, 3] << np.uint64(32))))
        h = mix64(h ^ (k[:, 4] | (k[:, 5] << np.uint64(32))))
        h = mix64(h ^ (k[:, 6] | (k[:, 7] << np.uint64(32))))
    return h


def hash_triplet(v, n, t, shape: int = 0, seed: int = 0) -> np.ndarray:
    v, n, t = (np.asarray(x).astype(np.uint64) for x in (v, n, t))
    with np.errstate(over="ignore"):
        a = v | (n << np.uint64(32))
        b = t | (np.uint64(shape) << np.uint64(32))
        return mix64(a ^ mix64(b + np.uint64(0x9E3779B97F4A7C15) + np.uint64(0xD6E8FEB86659FD93) * np.uint64(seed)))


def reduce32(x: np.ndarray, m: int) -> np.ndarray:
    return ((x.astype(np.uint64) & np.uint64(0xFFFFFFFF)) * np.uint64(m)) >> np.uint64(32)


def _uv_candidates(rng, count):
    """Distinct, finite, non-zero uv pairs in [1, 2) x [1, 2): no +-0 / NaN side effects, exact in float32."""
    bits = rng.integers(0x3F800000, 0x40000000, size=(count, 2), dtype=np.uint32)  # [1.0, 2.0)
    return bits


def value_cluster(pos3: np.ndarray, nrm3: np.ndarray, count: int, window_slots: int, table_slots: int, rng,
                  shape: int = 0, batch: int = 1 << 22, max_batches: int = 400) -> np.ndarray:
    """`count` distinct uv records (float32[count, 2]) such that the value keys (pos3, nrm3, uv) have the low 32
    bits of hash_key8 below 2^32 * window_slots / table_slots: in a table of `table_slots` (or fewer) slots they all
    start probing within the first `window_slots` slots."""
    base = np.concatenate([canon_bits(pos3), canon_bits(nrm3)]).astype(np.uint32)
    limit = np.uint64((1 << 32) * window_slots // max(table_slots, 1))
    found, seen = [], set()
    for _ in range(max_batches):
        uvb = _uv_candidates(rng, batch)
        k = np.empty((batch, 8), np.uint32)
        k[:, :6] = base
        k[:, 6:] = uvb
        h = hash_key8(k, shape)
        hit = np.nonzero((h & np.uint64(0xFFFFFFFF)) < limit)[0]
        for i in hit:
            key = (int(uvb[i, 0]), int(uvb[i, 1]))
            if key not in seen:
                seen.add(key)
                found.append(uvb[i])
        if len(found) >= count:
            break
    if len(found) < count:
        raise RuntimeError(f"value_cluster: found {len(found)} of {count} keys; widen the window")
    return np.asarray(found[:count], np.uint32).view(np.float32)


def fingerprint_twins(pos3, nrm3, buckets: int, pairs: int, rng, shape: int = 0, batch: int = 1 << 22,
                      max_batches: int = 64):
    """`pairs` pairs of DIFFERENT uv records whose value keys share the high 32 hash bits (the fingerprint) and the
    bucket reduce32(low 32 bits, buckets).  Returns float32[pairs, 2, 2]."""
    base = np.concatenate([canon_bits(pos3), canon_bits(nrm3)]).astype(np.uint32)
    pool_key = np.empty(0, np.uint64)
    pool_uv = np.empty((0, 2), np.uint32)
    out = []
    for _ in range(max_batches):
        uvb = _uv_candidates(rng, batch)
        k = np.empty((batch, 8), np.uint32)
        k[:, :6] = base
        k[:, 6:] = uvb
        h = hash_key8(k, shape)
        ident = ((h >> np.uint64(32)) << np.uint64(32)) | reduce32(h, buckets)  # (fingerprint, bucket)
        pool_key = np.concatenate([pool_key, ident])
        pool_uv = np.concatenate([pool_uv, uvb])
        order = np.argsort(pool_key, kind="stable")
        sk = pool_key[order]
        eq = np.nonzero(sk[1:] == sk[:-1])[0]
        out = []
        for j in eq:
            a, b = pool_uv[order[j]], pool_uv[order[j + 1]]
            if not np.array_equal(a, b):
                out.append(np.stack([a, b]))
        if len(out) >= pairs:
            break
    if len(out) < pairs:
        raise RuntimeError(f"fingerprint_twins: found {len(out)} of {pairs} pairs for {buckets} buckets")
    return np.asarray(out[:pairs], np.uint32).view(np.float32)


def triplet_cluster(v: int, nn: int, nt: int, count: int, window_slots: int, table_slots: int, shape: int = 0):
    """`count` distinct (n, t) pairs for the fixed position index `v` whose hash_triplet low 32 bits fall below
    2^32 * window_slots / table_slots.  Returns int32[count, 2]."""
    limit = np.uint64((1 << 32) * window_slots // max(table_slots, 1))
    found = []
    tt = np.arange(nt, dtype=np.uint64)
    for n in range(nn):
        h = hash_triplet(np.full(nt, v, np.uint64), np.full(nt, n, np.uint64), tt, shape)
        hit = np.nonzero((h & np.uint64(0xFFFFFFFF)) < limit)[0]
        found.extend((n, int(t)) for t in hit)
        if len(found) >= count:
            break
    if len(found) < count:
        raise RuntimeError(f"triplet_cluster: found {len(found)} of {count}; widen the window or enlarge nn x nt")
    return np.asarray(found[:count], np.int32)


def inject(m: MeshStreams, value_keys: int = 0, triplet_keys: int = 0, twin_pairs: int = 0, window_slots: int = 48,
           table_slots: int | None = None, fingerprint_buckets: int | None = None, seed: int = 0x5EED00C0) -> MeshStreams:
    """Appends engineered records to a single-shape mesh and triangles that reference them.

    `table_slots`: the largest table the clusters must stay clustered in (default: 2.5 x the largest attribute
    array + 2 K, the plan's own first sizing).  `fingerprint_buckets`: bucket count of the record schedule's
    fingerprint table for the RESULTING mesh; default derives it from the resulting array sizes the way
    vkmesh.cu does for a fresh plan ((2 * na + 1024) & ~7) / 8."""
    assert len(m.shape_offsets) == 2, "inject works on single-shape meshes"
    rng = np.random.Generator(np.random.PCG64(seed))
    pos, nrm, uv, idx = m.pos, m.nrm, m.uv, m.idx
    na_after = max(len(pos), len(nrm), len(uv) + value_keys + 2 * twin_pairs)
    if table_slots is None:
        table_slots = 2 * (na_after + na_after // 4 + 1024) + 64
    if fingerprint_buckets is None:
        fingerprint_buckets = ((2 * na_after + 1024) & ~7) // 8
    new_uv, corners = [], []  # corners: (v, n, t) of appended corners, t relative to len(uv) when >= 0 marker
    v0 = int(rng.integers(0, len(pos)))
    n0 = int(rng.integers(0, len(nrm)))
    t_base = len(uv)
    if value_keys:
        vc = value_cluster(pos[v0], nrm[n0], value_keys, window_slots, table_slots, rng)
        for j in range(value_keys):
            corners.append((v0, n0, t_base + len(new_uv) + j))
        new_uv.append(vc)
    if twin_pairs:
        done = sum(len(x) for x in new_uv)
        tw = fingerprint_twins(pos[v0], nrm[n0], fingerprint_buckets, twin_pairs, rng).reshape(-1, 2)
        for j in range(len(tw)):
            corners.append((v0, n0, t_base + done + j))
        new_uv.append(tw)
    if new_uv:
        uv = np.concatenate([uv] + [np.ascontiguousarray(x, np.float32) for x in new_uv])
    if triplet_keys:
        v1 = int(rng.integers(0, len(pos)))
        tc = triplet_cluster(v1, len(nrm), len(uv), triplet_keys, window_slots, table_slots)
        for n, t in tc:
            corners.append((v1, int(n), int(t)))
    # whole triangles: pad with repeats of the last corner (a degenerate triangle is fine)
    while len(corners) % 3:
        corners.append(corners[-1])
    extra = np.asarray(corners, np.int32).reshape(-1, 3)
    idx = np.concatenate([idx, extra])
    so = np.array([0, idx.shape[0]], np.uint64)
    meta = dict(m.meta, collisions={"value_keys": value_keys, "triplet_keys": triplet_keys, "twin_pairs": twin_pairs,
                                    "table_slots": table_slots, "fingerprint_buckets": fingerprint_buckets})
    return MeshStreams(pos, nrm, np.ascontiguousarray(uv, np.float32), idx, so, meta)
