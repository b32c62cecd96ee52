"""CPU tests of the multi-GPU host logic: LPT partition of a mesh batch and the
world_size-2 run over gloo.  The compute stand-in on CPU is the oracle (the product has no CPU path);
the same `run_sharded` drives the CUDA path in tests/test_gpu_parity.py and bench.py."""
import os
import socket

import numpy as np
import pytest

from vulkanengine_b200 import sharding


def test_lpt_partition_covers_every_mesh_once_and_balances():
    ns, tris = sharding.batch_sizes(2000)
    for w in (1, 2, 4, 8):
        parts = sharding.lpt_partition(tris, w)
        flat = sorted(i for p in parts for i in p)
        assert flat == list(range(2000))
        assert sharding.imbalance(tris, parts) < 1.01
    # deterministic
    assert sharding.lpt_partition(tris, 8) == sharding.lpt_partition(tris, 8)
    # degenerate inputs
    assert sharding.lpt_partition([], 4) == [[], [], [], []]
    assert sharding.lpt_partition([5], 3) == [[0], [], []]


def test_batch_sizes_match_config4_shape():
    ns, tris = sharding.batch_sizes(10000)
    assert tris.min() >= 900 and tris.max() <= 101_000
    assert 1.9e8 < tris.sum() < 2.4e8  # ~2.15e8 triangles, log-uniform 1k..100k


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, ns, q):
    import torch.distributed as dist

    from oracle import port as oracle_port

    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)

    def gather(obj):
        out = [None] * world
        dist.all_gather_object(out, obj)
        return out

    merged = sharding.run_sharded(ns, rank, world, oracle_port.import_arrays, gather)
    dist.barrier()
    if rank == 0:
        q.put(merged)
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_world_size_2_gloo_matches_single_process():
    import torch.multiprocessing as mp

    ns, _ = sharding.batch_sizes(24, min_tris=50, max_tris=4000)
    from oracle import port as oracle_port

    single = sharding.run_sharded(ns, 0, 1, oracle_port.import_arrays)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, ns, q)) for r in range(2)]
    [p.start() for p in procs]
    merged = q.get(timeout=240)
    [p.join(timeout=60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    assert sorted(merged) == list(range(24))
    assert merged == single
