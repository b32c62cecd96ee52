"""Host-side helpers that need no GPU: the soak's mesh generators (valid streams, the oracle accepts them), the hash
restatement used to engineer colliding keys, and bench.py's launch bookkeeping."""
import importlib.util
import os

import numpy as np

from conftest import ROOT
from oracle import port
from vulkanengine_b200 import hash_collide, synth


def _load(path, name):
    spec = importlib.util.spec_from_file_location(name, os.path.join(ROOT, path))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_soak_generators_make_valid_streams_the_oracle_accepts():
    soak = _load("scripts/soak_fuzz.py", "soak_fuzz")
    rng = np.random.default_rng(3)
    kinds, dirty = set(), 0
    for _ in range(12):
        # (the soup and sphere generators at reduced size: the CPU suite stays short)
        m = soak.soup(rng) if rng.random() < 0.5 else synth.uv_sphere(int(rng.integers(8, 40)), int(rng.integers(8, 40)))
        idx = soak.windowed_permutation(rng, m.idx[: 3 * 2000], int(rng.choice([1, 64, 1 << 30])))
        so = soak.cut_shapes(rng, len(idx) // 3, 5)
        m = synth.MeshStreams(m.pos, m.nrm, m.uv, np.ascontiguousarray(idx), so, dict(m.meta))
        if rng.random() < 0.5:
            m = soak.make_dirty(rng, m)
            dirty += 1
        assert m.idx.min() >= 0 and m.idx[:, 0].max() < len(m.pos) and m.idx[:, 1].max() < len(m.nrm) and m.idx[:, 2].max() < len(m.uv)
        assert m.shape_offsets[0] == 0 and m.shape_offsets[-1] == len(m.idx) and (np.diff(m.shape_offsets.astype(np.int64)) % 3 == 0).all()
        out = port.import_arrays(m.pos, m.nrm, m.uv, m.idx, m.shape_offsets)
        assert len(out) == len(m.shape_offsets) - 1
        assert sum(len(s.indices) for s in out) == len(m.idx)
        kinds.add(len(m.shape_offsets) > 2)
    assert dirty >= 2 and kinds == {False, True}


def test_value_cluster_lands_in_the_window_of_the_device_hash():
    rng = np.random.Generator(np.random.PCG64(7))
    pos3 = np.array([0.25, -1.5, 3.0], np.float32)
    nrm3 = np.array([0.0, 1.0, 0.0], np.float32)
    table, window = 1 << 16, 8
    uv = hash_collide.value_cluster(pos3, nrm3, 40, window, table, rng, batch=1 << 20)
    k = np.empty((len(uv), 8), np.uint32)
    k[:, :3], k[:, 3:6], k[:, 6:] = hash_collide.canon_bits(pos3), hash_collide.canon_bits(nrm3), uv.view(np.uint32)
    home = hash_collide.reduce32(hash_collide.hash_key8(k), table)
    assert len(np.unique(uv.view(np.uint64))) == 40 and home.max() < window
    # a smaller table keeps them clustered too (the plan may start smaller than the estimate)
    assert hash_collide.reduce32(hash_collide.hash_key8(k), table // 4).max() < window


def test_bench_counts_the_lattice_launches_of_the_general_schedule():
    bench = _load("bench.py", "bench_mod")
    small, big = synth.grid(64), synth.grid(8)
    big = synth.MeshStreams(np.zeros((4_000_000, 3), np.float32), big.nrm, big.uv, big.idx, big.shape_offsets)
    assert bench.general_launches(small, {"hash_seed": 0}) == 13      # table below 32 MB: hashed homes
    assert bench.general_launches(big, {"hash_seed": 0}) == 15        # + k_sample_box, k_cell_map
    assert bench.general_launches(big, {"hash_seed": 1}) == 13        # a later hash generation scrambles
    two = synth.MeshStreams(big.pos, big.nrm, big.uv, big.idx, np.array([0, 3, len(big.idx)], np.uint64))
    assert bench.general_launches(two, {"hash_seed": 0}) == 13        # several shapes: hashed homes
