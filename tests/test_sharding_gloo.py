"""N>1 host logic on CPU: world_size-2 gloo processes shard a unit range and all-gather their summaries
(the one collective of the path)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from openbeam_b200 import sharding


def test_shard_range_partitions():
    for total in (0, 1, 7, 64, 1_000_001):
        for world in (1, 2, 3, 8):
            r = [sharding.shard_range(total, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == total
            assert all(r[k][1] == r[k + 1][0] for k in range(world - 1))
            sizes = [b - a for a, b in r]
            assert max(sizes) - min(sizes) <= 1
    assert sharding.weak_range(10, 3) == (30, 40)


def _worker(rank, world, port, total, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = sharding.shard_range(total, rank, world)
    # stand-in for obgpu_summaries_device output: row i = f(global unit index)
    idx = torch.arange(lo, hi, dtype=torch.float64)
    local = torch.stack([idx, 2 * idx, -idx, torch.zeros_like(idx)], 1)
    g = sharding.gather_summaries(local)
    # equal-shard fast path
    eq = torch.full((5, 4), float(rank), dtype=torch.float64)
    out = torch.empty((world * 5, 4), dtype=torch.float64)
    g2 = sharding.gather_summaries(eq, out)
    if rank == 0:
        q.put((g.numpy(), g2.numpy()))
    dist.barrier()
    dist.destroy_process_group()


def test_gather_summaries_world2():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    total, world = 11, 2   # ragged: 6 + 5
    procs = [ctx.Process(target=_worker, args=(r, world, port, total, q)) for r in range(world)]
    for p in procs:
        p.start()
    g, g2 = q.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    idx = np.arange(total, dtype=np.float64)
    assert np.array_equal(g, np.stack([idx, 2 * idx, -idx, 0 * idx], 1))
    assert np.array_equal(g2[:5], np.zeros((5, 4))) and np.array_equal(g2[5:], np.ones((5, 4)))
