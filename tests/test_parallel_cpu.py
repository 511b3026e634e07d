"""world_size-2 gloo tests (CPU) of the host-side sharding logic used at N > 1."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from pathnet_gym_b200 import parallel


def test_shard_range_partitions_population():
    for P in (1, 7, 64, 512):
        for ws in (1, 2, 3, 8):
            got, prev = [], 0
            for r in range(ws):
                lo, hi = parallel.shard_range(P, r, ws)
                assert lo == prev and hi >= lo
                prev = hi
                got.append(hi - lo)
            assert prev == P and max(got) - min(got) <= 1
    assert parallel.shard_range(512, 3, 8) == (192, 256)


def _worker(rank, ws, port, q):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    try:
        # gradient slab: rank r contributes r+1 everywhere -> sum = ws(ws+1)/2
        g = torch.full((1000,), float(rank + 1))
        parallel.allreduce_sum_(g)
        ok = bool(torch.all(g == ws * (ws + 1) / 2))
        # scores of uneven slices gathered in rank order
        P = 5
        lo, hi = parallel.shard_range(P, rank, ws)
        local = torch.arange(lo, hi, dtype=torch.float32)
        sizes = [parallel.shard_range(P, r, ws) for r in range(ws)]
        allv = parallel.allgather_concat(local, [b - a for a, b in sizes])
        ok = ok and torch.equal(allv, torch.arange(P, dtype=torch.float32))
        ok = ok and parallel.world() == (rank, ws)
        q.put((rank, ok))
    finally:
        dist.destroy_process_group()


def test_gloo_world_size_2():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(res) == [(0, True), (1, True)]
