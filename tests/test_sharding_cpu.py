"""Host-side multi-rank logic on CPU: shard arithmetic, and a world-size-2 gloo run in which each rank
answers its shard of a query batch and rank 0 reassembles the batch (no GPU: the per-rank work is done by
the CPU oracle, the thing under test is the sharding / gathering plumbing of mola_b200.sharding)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_range_partitions():
    from mola_b200.sharding import shard_range, shard_scans
    for n in (0, 1, 7, 64, 1_000_003):
        for world in (1, 2, 3, 4, 8):
            spans = [shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    assert sum(len(shard_scans(64, r, 8)) for r in range(8)) == 64
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from mola_b200 import sharding, synth
        from oracle import mmo
        xs, ys, zs, pose = synth.lidar_scan("tiny")
        m = mmo.OracleMap(0, 0.5)           # every rank holds a replica (deterministic build)
        m.insert(xs, ys, zs, pose)
        T = synth.perturbed_pose(pose)
        x, y, z = xs.astype(np.float64), ys.astype(np.float64), zs.astype(np.float64)
        g = [(((T[3 * r] * x + T[3 * r + 1] * y) + T[3 * r + 2] * z) + T[9 + r]).astype(np.float32) for r in range(3)]
        n = g[0].size
        lx, ly, lz = sharding.shard_points(*g, rank, world)
        local = m.nn_single(lx, ly, lz)
        full = sharding.gather_to_root(local, n, dist)
        if rank == 0:
            ref = m.nn_single(*g)
            ok = all(np.array_equal(a, b) for a, b in zip(full, ref)) and full[3].dtype == np.uint64
            q.put(bool(ok))
        else:
            assert full is None
    finally:
        dist.destroy_process_group()


def test_world2_gloo_sharded_queries_gather():
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True
