#!/usr/bin/env python
"""BASELINE config C5: ONE large stitched image (default 8192x8192, 256-px windows, subdivisions 2), sharded
over the ranks by owner-computes row strips (no data-path reduction; one all_gather of the result strips).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        scripts/bench_c5_rowshard.py [--size 8192] [--steps 2] [--check]

Prints one JSON line on rank 0: MP/s of the whole image (max-over-ranks device time), patches per rank, and --
with --check, at sizes where a single GPU can hold the full prediction -- whether the gathered image is
bit-identical to the single-GPU result."""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=8192)
    ap.add_argument("--ws", type=int, default=256)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--check", action="store_true")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    from bench import synth_image, FILTERS
    from tmat_b200 import _ffi, models
    from tmat_b200.smooth_tiled_predictions import strip_rows, patches_for_rows, predict_img_with_smooth_windowing

    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = _ffi.Context(local)
    _ffi._tls.ctx = ctx
    L = _ffi.lib()
    net = models.CudaUNet(models.random_init_weights(FILTERS, seed=0), args.ws, FILTERS, impl="tc", ctx=ctx)
    size = args.size
    base = synth_image(min(size, 2048), 5000)
    reps = -(-size // base.shape[0])
    img = np.tile(base, (reps, reps))[:size, :size].copy()      # mosaic of C1-style fields
    d_img = torch.from_numpy(img).cuda()
    row0, row1 = strip_rows(size, rank, world)
    d_out = torch.empty((row1 - row0, size), dtype=torch.float64, device="cuda")
    rows_max = -(-size // world)
    gather_in = torch.zeros((rows_max, size), dtype=torch.float64, device="cuda")
    gathered = [torch.empty_like(gather_in) for _ in range(world)] if world > 1 else None
    stream = torch.cuda.current_stream()
    sptr = _ffi.C.c_void_p(stream.cuda_stream)

    def step():
        _ffi.check(L.tmat_stp_predict_rows_dev(ctx.handle, net._tmat_handle, _ffi.ptr(d_img), size, size, 2, row0, row1,
                                               _ffi.ptr(d_out), None, sptr))
        if world > 1:
            gather_in[: row1 - row0].copy_(d_out)
            dist.all_gather(gathered, gather_in)

    step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / args.steps], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    n_mine = len(patches_for_rows(size, size, args.ws, 2, row0, row1))
    counts = torch.tensor([n_mine], dtype=torch.int64, device="cuda")
    all_counts = [torch.zeros_like(counts) for _ in range(world)]
    if world > 1:
        dist.all_gather(all_counts, counts)
    else:
        all_counts = [counts]
    identical = None
    if args.check:
        if world > 1:
            full_sharded = np.concatenate([gathered[r][: strip_rows(size, r, world)[1] - strip_rows(size, r, world)[0]]
                                           .cpu().numpy() for r in range(world)])
        else:
            full_sharded = d_out.cpu().numpy()
        if rank == 0:
            ref = predict_img_with_smooth_windowing(img, args.ws, 2, net.predict)
            identical = bool(np.array_equal(ref, full_sharded))
    if rank == 0:
        n_total = len(patches_for_rows(size, size, args.ws, 2, 0, size))
        ms = float(t.item())
        print(json.dumps({"workload": f"C5: one {size}x{size} image, ws {args.ws}, subdiv 2, row strips over {world} GPU(s)",
                          "n_gpus": world, "ms_per_image": ms, "value": size * size / 1e6 / (ms / 1e3), "unit": "MP/s",
                          "patches_total": n_total, "patches_per_rank": [int(c.item()) for c in all_counts],
                          "redundant_patch_fraction": sum(int(c.item()) for c in all_counts) / n_total - 1.0,
                          "bit_identical_to_single_gpu": identical}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
