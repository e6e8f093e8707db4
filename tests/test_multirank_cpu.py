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


# ---- row-strip sharding of ONE large image (host geometry; the kernels are covered by tests/test_gpu_stp.py) ----
def _ref_patches_for_rows(h, w, ws, sd, row0, row1):
    """Independent restatement with the oracle's D4 helpers: a patch is needed iff its window in the padded
    ORIGINAL frame (oracle.stp.tile_origin) intersects padded rows [row0+aug, row1+aug)."""
    from oracle import stp
    aug, step = stp.aug_of(ws, sd), ws // sd
    PM, PN = h + 2 * aug, w + 2 * aug
    ids, per = [], None
    for g in range(8):
        Lm, Ln = (PM, PN) if g % 2 == 0 else (PN, PM)
        rows, cols = range(0, Lm - ws + 1, step), range(0, Ln - ws + 1, step)
        per = len(rows) * len(cols)
        for a, i in enumerate(rows):
            for b, j in enumerate(cols):
                r0, _ = stp.tile_origin(g, i, j, PM, PN, ws)
                if r0 < row1 + aug and r0 + ws > row0 + aug:
                    ids.append(g * per + a * len(cols) + b)
    return np.array(ids, np.int32), 8 * per


def _strip_worker(rank, world, port, shape, ws, sd, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from tmat_b200.smooth_tiled_predictions import strip_rows, patches_for_rows
    h, w = shape
    row0, row1 = strip_rows(h, rank, world)
    ids = patches_for_rows(h, w, ws, sd, row0, row1)
    ref, n_total = _ref_patches_for_rows(h, w, ws, sd, row0, row1)
    ok = torch.tensor([int(np.array_equal(ids, ref)), row0, row1, len(ids), n_total], dtype=torch.int64)
    gathered = [torch.zeros_like(ok) for _ in range(world)]
    dist.all_gather(gathered, ok)
    if rank == 0:
        out.put(torch.stack(gathered).numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_row_strip_sharding_geometry_world2():
    shape, ws, sd = (1000, 700), 64, 2
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_strip_worker, args=(r, 2, port, shape, ws, sd, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res[:, 0].all()                                   # C geometry == oracle restatement on every rank
    assert res[0, 1] == 0 and res[0, 2] == res[1, 1] and res[1, 2] == shape[0]   # strips tile the rows
    n_total = res[0, 4]
    assert res[:, 3].max() < n_total                          # each rank evaluates a strict subset ...
    assert n_total <= res[:, 3].sum() <= 1.35 * n_total       # ... with a bounded halo of redundant patches


def test_strip_rows_cover_every_split():
    from tmat_b200.smooth_tiled_predictions import strip_rows
    for h in (1, 7, 64, 8192):
        for world in (1, 2, 3, 8):
            if world > h:
                continue
            edges = [strip_rows(h, r, world) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == h
            assert all(edges[i][1] == edges[i + 1][0] for i in range(world - 1))
            assert max(b - a for a, b in edges) - min(b - a for a, b in edges) <= 1
