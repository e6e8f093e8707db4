"""CPU, world_size 2 over gloo: the image-sharding host logic bench.py uses at N > 1 (static
round-robin of image indices, max-over-ranks timing, final gather of per-image metrics) gives the same
per-image results whatever the shard count.  The per-image work is the oracle here (no GPU)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import helpers
from oracle import zstacks as oz


def shard(n_items, rank, world):
    """bench.py / compute_branches.py:585 partitioning: image i belongs to rank i % world."""
    return list(range(rank, n_items, world))


def _worker(rank, world, port, n_items, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = shard(n_items, rank, world)
    vals = torch.full((n_items,), -1.0, dtype=torch.float64)
    for i in mine:
        vals[i] = float(oz.proj_max(helpers.synth_stack(3, 24, 20, 100 + i)).mean())
    t = torch.tensor([0.1 * (rank + 1)], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    gathered = [torch.zeros_like(vals) for _ in range(world)]
    dist.all_gather(gathered, vals)
    merged = torch.stack(gathered).max(0).values
    if rank == 0:
        out.put((merged.numpy(), float(t.item())))
    dist.barrier()
    dist.destroy_process_group()


def test_image_sharding_is_shard_count_independent():
    n_items = 5
    assert sorted(shard(n_items, 0, 2) + shard(n_items, 1, 2)) == list(range(n_items))
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_items, q)) for r in range(2)]
    for p in procs:
        p.start()
    merged, tmax = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    single = np.array([float(oz.proj_max(helpers.synth_stack(3, 24, 20, 100 + i)).mean()) for i in range(n_items)])
    assert np.array_equal(merged, single)
    assert abs(tmax - 0.2) < 1e-12
