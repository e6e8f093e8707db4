#!/usr/bin/env python
"""bench.py -- megapixels/s segmented (tiled UNet + 8x D4 + spline blend) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--size 2048]
                    [--no-cpu-baseline] [--no-e2e] [--no-extras]

Headline (the one JSON line the driver reads).  One "step" = one synthetic 2048x2048 float32 Z-projection
(BASELINE.json config C4's per-image unit: 13x13 windows x 8 orientations = 1352 patches of 320x320) through
``predict_img_with_smooth_windowing`` with the random-init binary-segmentation UNet.  With N ranks (torchrun)
every rank segments its own image per step (images are independent, scripts/compute_branches.py:585): weak
scaling, no collective on the data path; NCCL carries only the timing reduction and the final gather of
per-image metrics.

Keys beyond the base contract:
  roofline      tensor bound.  `achieved`/`frac` are EXECUTED tensor-pipe FLOPs of the tcgen05 GEMM kernels (what the
                pipe really did: the up path is evaluated per output parity, 4 taps instead of 9) over their summed
                CUDA-event time, against the measured sustained bf16 peak.  `effective_tflops` is the ALGORITHMIC count
                (31.026 GFLOP per 320^2 patch as written in models.py:110-171) over the same time -- a speed-up
                figure, not a utilisation: it is not divided by the hardware peak.
  cpu_baseline  N=1 only: the reference's own smooth_tiled_predictions module (baseline/_ref, imported unchanged)
                driving the oracle UNet (torch-CPU fp32; TensorFlow is not installable) on a bounded crop.
  e2e           the public Python call with host buffers, H2D/D2H inside the timed region.
  stages        after the timed region: C2 (five Z projections of a 32x2048x2048 uint16 stack: ms, algorithmic GB/s,
                fraction of the measured HBM copy peak) and C3 (auto_threshold_well + filter_branch_seg_mask over a
                batch of 2048^2 images per rank: images/s kernels-only and end-to-end) and C4 (the whole mask path, stack
                in -> cleaned mask out, one image per rank: ms, MP/s).
  c5            one 8192x8192 image, 256-px windows, owner-computes row strips over the N ranks (+ one all_gather):
                ms, MP/s, redundant patches -- so the driver's 1/2/4/8 run records the row-strip split too.

`--impl reference`: the reference's CPU path on the host cores for the same config -- the unmodified
smooth_tiled_predictions.predict_img_with_smooth_windowing from baseline/_ref with the oracle UNet as pred_func, whole
images per step, steps bounded so the run ends within a few minutes.  Rank 0 only; other ranks exit before any
process group exists.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WS, SUBDIV = 320, 2
FILTERS = (64, 128, 256, 512)
METRIC = "megapixels/sec segmented (tiled UNet+D4+blend)"
REF_DIR = os.path.join(ROOT, "baseline", "_ref")


def synth_image(size, seed):
    """C1/C4-style projection: noisy background + bright vessel-like polylines, in [0, 1]."""
    rng = np.random.default_rng(seed)
    img = rng.normal(0.15, 0.03, (size, size)).astype(np.float32)
    yy, xx = np.mgrid[0:size, 0:size].astype(np.float32)
    for _ in range(24):
        x0, y0, x1, y1 = rng.uniform(0, size, 4)
        width, inten = rng.uniform(1.5, 4.5), rng.uniform(0.5, 1.0)
        dx, dy = x1 - x0, y1 - y0
        t = np.clip(((xx - x0) * dx + (yy - y0) * dy) / (dx * dx + dy * dy + 1e-9), 0, 1)
        d = np.hypot(xx - (x0 + t * dx), yy - (y0 + t * dy))
        np.maximum(img, inten * np.clip(width - d + 0.5, 0, 1), out=img)
    return np.clip(img, 0, 1).astype(np.float32)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"bf16_burst": d.get("bf16_tflops"), "bf16_sustained": d.get("bf16_tflops_sustained"),
                "hbm": d.get("hbm_gbs"), "source": "MEASURED_PEAKS.json"}
    return {"bf16_burst": 1590.0, "bf16_sustained": 1400.0, "hbm": 6650.0, "source": "fallback (B200_PROFILING.md)"}


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons with NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.samples, self.reasons, self.max_mhz = index, False, [], set(), None

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40,
                     "sw_thermal_slowdown": 0x20, "hw_power_brake": 0x80}
            while not self.stop_flag:
                self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
                time.sleep(0.05)
        except Exception as e:   # NVML missing: report that rather than inventing clocks
            self.reasons.add(f"nvml_unavailable:{type(e).__name__}")

    def result(self):
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None,
                "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


# ---- the reference's CPU path ---------------------------------------------------------------------------------
def reference_stp():
    """The reference's own tiling/blending module, unmodified, from baseline/_ref (pip-installed there by
    __graft_entry__.build()).  Falls back to the oracle restatement when the install is absent."""
    if os.path.isdir(os.path.join(REF_DIR, "fl_tissue_model_tools")):
        if REF_DIR not in sys.path:
            sys.path.insert(0, REF_DIR)
        try:
            from fl_tissue_model_tools import smooth_tiled_predictions as r
            return r, "fl_tissue_model_tools.smooth_tiled_predictions from baseline/_ref (unmodified reference module)"
        except Exception as e:   # noqa: BLE001 -- report and fall back to the restatement
            why = f"{type(e).__name__}: {e}"
    else:
        why = "baseline/_ref not installed"
    from oracle import stp as r
    return r, f"oracle/stp.py restatement ({why})"


def cpu_pred_func():
    import torch
    from oracle import unet as ou
    cores = len(os.sched_getaffinity(0))
    torch.set_num_threads(cores)
    w = ou.make_weights(FILTERS, seed=0)
    return ou.make_pred_func(w), cores


def cpu_segment(img, pred, stp_mod):
    """One whole image through the reference's predict_img_with_smooth_windowing (batches of 16 tiles,
    smooth_tiled_predictions.py:174-182).  Returns (seconds, result)."""
    import contextlib
    import io
    t0 = time.perf_counter()
    with contextlib.redirect_stdout(io.StringIO()), contextlib.redirect_stderr(io.StringIO()):   # its tqdm bars
        out = stp_mod.predict_img_with_smooth_windowing(img, window_size=WS, subdivisions=SUBDIV, pred_func=pred)
    return time.perf_counter() - t0, out


def patches_of(size):
    n_r = (size + 2 * (WS // 2) - WS) // (WS // SUBDIV) + 1
    return 8 * n_r * n_r


def reference_arm(args):
    """`--impl reference`: whole images of the SAME config through the reference module on the host cores."""
    pred, cores = cpu_pred_func()
    stp_mod, how = reference_stp()
    size = args.size
    img = synth_image(size, 4000)
    # warm-up: two 16-tile batches through pred_func (thread pool, allocator), which also estimate the step time
    batch16 = np.ascontiguousarray(img[:WS, :WS][None].repeat(16, 0))
    pred(batch16, verbose=0)
    t0 = time.perf_counter()
    pred(batch16, verbose=0)
    est = (time.perf_counter() - t0) * patches_of(size) / 16
    budget = 130.0
    steps = int(max(1, min(args.steps, budget // max(est, 1e-3))))
    times = []
    for _ in range(steps):
        dt, _ = cpu_segment(img, pred, stp_mod)
        times.append(dt)
    dt = float(np.mean(times))
    value = size * size / 1e6 / dt
    sample = (f"{steps} whole {size}x{size} image(s) = {patches_of(size)} patches of 320x320 each, through {how} with "
              f"pred_func = oracle UNet (torch-CPU fp32, {cores} threads, batches of 16); no extrapolation")
    line = {"metric": METRIC, "value": value, "unit": "MP/s", "n_gpus": args.gpus, "steps": steps,
            "warmup": 2, "ms_per_step": dt * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "impl": "reference",
            "config": {"workload": f"C4 per-image unit: one {size}x{size} float32 projection per step -> "
                                   f"{patches_of(size)} patches of 320x320 (ws 320, subdiv 2, 8x D4) through the "
                                   "random-init binary_segmentation UNet (64/128/256/512) + spline blend",
                       "steps_requested": args.steps, "steps_bound": f"{budget:.0f} s of CPU work",
                       "warmup_unit": "batches of 16 tiles through pred_func"},
            "cpu_baseline": {"value": value, "unit": "MP/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "MP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "note": "kind=port: the tiling/D4/blend code is the reference's own module; the UNet forward is the torch-CPU "
                    "restatement of models.py:110-171 because TensorFlow/Keras is installable neither here nor on the "
                    "GPU box (probed)"}
    print(json.dumps(line), flush=True)


def cpu_baseline_leg(size):
    """N=1 leg of the GPU arm: the same reference module on a bounded crop (about 10-30 s of CPU work)."""
    pred, cores = cpu_pred_func()
    stp_mod, how = reference_stp()
    crop = min(size, 512)
    img = synth_image(size, 4000)[:crop, :crop].copy()
    pred(np.ascontiguousarray(img[:WS, :WS][None].repeat(16, 0)), verbose=0)
    dt, _ = cpu_segment(img, pred, stp_mod)
    n = patches_of(crop)
    return {"value": crop * crop / 1e6 / dt, "unit": "MP/s", "cores": cores, "kind": "port",
            "sample": f"one {crop}x{crop} crop of the step's image = {n} patches of 320x320 ({dt:.1f} s), through {how} "
                      f"with pred_func = oracle UNet (torch-CPU fp32, batches of 16); the crop pays "
                      f"{n / (crop * crop / 1e6):.0f} patches per MP against {patches_of(size) / (size * size / 1e6):.0f} "
                      f"for the full image (padding), so this figure is a lower bound of the full-image CPU rate"}


# ---- extras: the other named configs, measured after the headline's timed region ------------------------------
def stages_leg(ctx, world, rank, pk):
    """C2 and C3 on this rank's GPU (scripts/bench_stages.py holds the per-kernel version with CPU timings)."""
    import torch
    import torch.distributed as dist
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import helpers
    from tmat_b200 import _ffi, well_mask_generation as W, transforms as T
    L, C = _ffi.lib(), _ffi.C
    stream = torch.cuda.current_stream()
    sptr = C.c_void_p(stream.cuda_stream)
    S, Z = 2048, 32
    out = {}

    def timed(fn, iters=8, warm=3):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
        for e0, e1 in evs:
            e0.record(stream)
            fn()
            e1.record(stream)
        torch.cuda.synchronize()
        ts = sorted(e0.elapsed_time(e1) for e0, e1 in evs)
        return ts[len(ts) // 2]

    # C2: the stack is generated on the device (content does not change the traffic): textured base, per-slice gain
    g = torch.Generator(device="cuda").manual_seed(2000 + rank)
    base = torch.randint(200, 4000, (S, S), generator=g, device="cuda", dtype=torch.int32)
    stack = torch.empty((Z, S, S), dtype=torch.int16, device="cuda")
    for z in range(Z):
        noise = torch.randint(0, 600, (S, S), generator=g, device="cuda", dtype=torch.int32)
        stack[z] = ((base * (20 + abs(z - Z // 2))) // 24 + noise).clamp_(0, 32767).to(torch.int16)
    c2 = {}
    for name in ("max", "min", "avg", "med", "fs"):
        wide = name in ("avg", "med")
        d_out = torch.empty((S, S), dtype=torch.float64 if wide else torch.int16, device="cuda")

        def fn():
            _ffi.check(L.tmat_zproj_dev(ctx.handle, _ffi.ptr(stack), _ffi.U16, Z, S, S, _ffi.ZPROJ[name], 5,
                                        _ffi.ptr(d_out), sptr))
        ms = timed(fn)
        nbytes = Z * S * S * 2 + S * S * (8 if wide else 2)
        gbs = nbytes / (ms * 1e-3) / 1e9
        c2[name] = {"ms": round(ms, 4), "GB/s": round(gbs, 1), "frac": round(gbs / pk["hbm"], 3)}
    out["C2"] = {"workload": f"five Z projections of a ({Z},{S},{S}) uint16 stack, device-resident (268 MB > L2)",
                 "bytes": "Z*2 B read + 2 B (max/min/fs) or 8 B (avg/med) written per pixel", "hbm_peak_GB/s": pk["hbm"],
                 "proj": c2}
    del stack

    # C3: B images per rank; per image auto_threshold_well (Gaussian + Otsu + disk(5) erosion) and
    # filter_branch_seg_mask (median, labelling, perimeter, thinning, fork test) -- device-resident, then end to end
    B = 32
    well = torch.from_numpy(helpers.synth_well(S, S, 3000 + rank).view(np.int16)).cuda()
    wells = torch.empty((B, S, S), dtype=torch.int16, device="cuda")
    for b in range(B):
        wells[b] = torch.roll(well, shifts=(7 * b, 13 * b), dims=(0, 1))
    m512 = helpers.synth_mask(512, 512, 3001)
    seg = torch.from_numpy(np.tile(m512, (S // 512, S // 512)).astype(np.uint8)).cuda()
    segs = torch.empty((B, S, S), dtype=torch.uint8, device="cuda")
    for b in range(B):
        segs[b] = torch.roll(seg, shifts=(11 * b, 5 * b), dims=(0, 1))
    d_m = torch.empty((S, S), dtype=torch.uint8, device="cuda")
    d_o = torch.empty((S, S), dtype=torch.uint8, device="cuda")
    def batch():
        for b in range(B):
            _ffi.check(L.tmat_auto_threshold_well_dev(ctx.handle, _ffi.ptr(wells[b]), _ffi.U16, S, S, _ffi.ptr(d_m), None,
                                                      sptr))
            _ffi.check(L.tmat_filter_branch_seg_mask_dev(ctx.handle, _ffi.ptr(segs[b]), S, S, 1, 1, _ffi.ptr(d_o), sptr))
    ms = timed(batch, iters=3, warm=1)
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    well_h = well.cpu().numpy().view(np.uint16)
    seg_h = seg.cpu().numpy().astype(bool)
    W.auto_threshold_well(well_h)
    T.filter_branch_seg_mask(seg_h)
    n_e2e = 8
    t0 = time.perf_counter()
    for _ in range(n_e2e):
        W.auto_threshold_well(well_h)
        T.filter_branch_seg_mask(seg_h)
    dt = torch.tensor([(time.perf_counter() - t0) / n_e2e], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    floor_bytes = S * S * 3 + S * S * 2
    out["C3"] = {"workload": f"{B} images of {S}x{S} per rank: auto_threshold_well (uint16 in) + filter_branch_seg_mask "
                             "(uint8 mask in)", "images_per_rank": B, "n_gpus": world,
                 "ms_per_image_kernels": round(ms / B, 4), "images_per_s_kernels": round(world * B / (ms / 1e3), 1),
                 "images_per_s_e2e": round(world / float(dt.item()), 1),
                 "e2e": "numpy in -> H2D -> kernels -> D2H -> numpy out through the Python mirrors, one image at a time",
                 "floor_bytes_per_image": floor_bytes,
                 "frac_of_hbm_floor": round(floor_bytes * B / (ms * 1e-3) / 1e9 / pk["hbm"], 4)}
    del wells, segs

    # C4: the WHOLE branching-mask path for one image per rank per step (compute_branches.py:228, 316-337):
    # (32, 2048, 2048) uint16 stack -> max projection -> [0, 1] rescale -> tiled UNet x8 D4 + blend -> > 0.5 -> cleanup
    from tmat_b200 import models
    net = models.CudaUNet(models.random_init_weights(FILTERS, seed=0), WS, FILTERS, impl="tc", ctx=ctx)
    g = torch.Generator(device="cuda").manual_seed(4000 + rank)
    st4 = torch.randint(200, 30000, (Z, S, S), generator=g, device="cuda", dtype=torch.int32).to(torch.int16)
    d_pred = torch.empty((S, S), dtype=torch.float64, device="cuda")
    d_mask = torch.empty((S, S), dtype=torch.uint8, device="cuda")

    def full_path():
        _ffi.check(L.tmat_branch_mask_dev(ctx.handle, net._tmat_handle, _ffi.ptr(st4), Z, S, S, SUBDIV, _ffi.ptr(d_pred),
                                          _ffi.ptr(d_mask), sptr))
    ms = timed(full_path, iters=3, warm=2)
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    out["C4"] = {"workload": f"whole mask path per image: ({Z},{S},{S}) uint16 stack -> max projection -> rescale -> tiled UNet "
                             "(ws 320, 8x D4) + blend -> > 0.5 -> filter_branch_seg_mask; one image per rank per step",
                 "n_gpus": world, "ms_per_image": round(ms, 3), "MP/s": round(world * S * S / 1e6 / (ms / 1e3), 2)}
    net.close()
    return out


def c5_leg(ctx, world, rank, steps=2, size=8192, ws=256):
    """One large image over the N ranks by owner-computes row strips + one all_gather of the strips."""
    import torch
    import torch.distributed as dist
    from tmat_b200 import _ffi, models
    from tmat_b200.smooth_tiled_predictions import strip_rows, patches_for_rows
    L = _ffi.lib()
    net = models.CudaUNet(models.random_init_weights(FILTERS, seed=0), ws, FILTERS, impl="tc", ctx=ctx)
    base = synth_image(2048, 5000)
    reps = -(-size // base.shape[0])
    d_img = torch.from_numpy(np.tile(base, (reps, reps))[:size, :size].copy()).cuda()
    row0, row1 = strip_rows(size, rank, world)
    d_out = torch.empty((row1 - row0, size), dtype=torch.float64, device="cuda")
    rows_max = -(-size // world)
    gather_in = torch.zeros((rows_max, size), dtype=torch.float64, device="cuda")
    gathered = [torch.empty_like(gather_in) for _ in range(world)] if world > 1 else None
    stream = torch.cuda.current_stream()
    sptr = _ffi.C.c_void_p(stream.cuda_stream)

    def step():
        _ffi.check(L.tmat_stp_predict_rows_dev(ctx.handle, net._tmat_handle, _ffi.ptr(d_img), size, size, SUBDIV, row0, row1,
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
    for _ in range(steps):
        step()
    e1.record(stream)
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / steps], dtype=torch.float64, device="cuda")
    n_mine = torch.tensor([len(patches_for_rows(size, size, ws, SUBDIV, row0, row1))], dtype=torch.int64, device="cuda")
    counts = [torch.zeros_like(n_mine) for _ in range(world)]
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_gather(counts, n_mine)
    else:
        counts = [n_mine]
    n_total = len(patches_for_rows(size, size, ws, SUBDIV, 0, size))
    ms = float(t.item())
    net.close()
    return {"workload": f"one {size}x{size} image, ws {ws}, subdiv 2 ({n_total} patches), owner-computes row strips over "
                        f"{world} GPU(s) + one all_gather of the float64 strips", "n_gpus": world, "steps": steps,
            "scaling": "strong", "ms_per_image": round(ms, 2), "MP/s": round(size * size / 1e6 / (ms / 1e3), 2),
            "patches_per_rank": [int(c.item()) for c in counts],
            "redundant_patch_fraction": round(sum(int(c.item()) for c in counts) / n_total - 1.0, 4)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--size", type=int, default=2048)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="skip the end-to-end leg (profiling runs)")
    ap.add_argument("--no-extras", action="store_true", help="skip the C2/C3 stage table and the C5 leg (profiling runs)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        if rank == 0:          # the other ranks leave before any process group exists: no GPU spins in a barrier
            reference_arm(args)
        return

    import torch
    import torch.distributed as dist
    from tmat_b200 import _ffi, models
    from tmat_b200.smooth_tiled_predictions import predict_img_with_smooth_windowing

    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    warmup = max(args.warmup, 3)
    ctx = _ffi.Context(local)
    _ffi._tls.ctx = ctx
    L = _ffi.lib()
    net = models.CudaUNet(models.random_init_weights(FILTERS, seed=0), WS, FILTERS, impl="tc", ctx=ctx)
    flops_patch = net.flops_per_patch()
    size = args.size
    img_host = torch.from_numpy(synth_image(size, 4000 + rank)).pin_memory()
    img_dev = img_host.cuda(non_blocking=True)
    out_dev = torch.empty((size, size), dtype=torch.float64, device="cuda")
    stream = torch.cuda.current_stream()
    sptr = _ffi.C.c_void_p(stream.cuda_stream)

    def step():
        _ffi.check(L.tmat_stp_predict_dev(ctx.handle, net._tmat_handle, _ffi.ptr(img_dev), size, size, SUBDIV,
                                          _ffi.ptr(out_dev), sptr))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(warmup):
        step()
    barrier()
    ctx.profile(True)
    ctx.profile_collect()
    launches0 = ctx.launch_count()
    sampler = ClockSampler(local)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        step()
    e1.record(stream)
    barrier()
    sampler.stop_flag = True
    ms_total = e0.elapsed_time(e1)
    tc_ms, tc_launches, tc_flops = ctx.profile_collect()
    ctx.profile(False)
    launches = ctx.launch_count() - launches0
    sampler.join(timeout=2)
    t = torch.tensor([ms_total], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t.item()) / args.steps
    mp = size * size / 1e6
    value = world * mp / (ms_step / 1e3)

    # e2e: the public Python call with pinned host buffers (H2D of the image, D2H of the result inside)
    e2e_steps = 1 if args.no_e2e else max(3, args.steps // 2)
    img_np = img_host.numpy()
    predict_img_with_smooth_windowing(img_np, WS, SUBDIV, net.predict)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        res = predict_img_with_smooth_windowing(img_np, WS, SUBDIV, net.predict)
    torch.cuda.synchronize()
    dt = torch.tensor([(time.perf_counter() - t0) / e2e_steps], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    e2e_value = world * mp / float(dt.item())

    # per-image metric gathered at the end (the only collective of the path)
    frac = torch.tensor([float((res > 0.5).mean())], dtype=torch.float64, device="cuda")
    fracs = [torch.zeros_like(frac) for _ in range(world)]
    if world > 1:
        dist.all_gather(fracs, frac)
    else:
        fracs = [frac]

    pk = peaks()
    extras = {}
    if not args.no_extras and size == 2048:
        net.close()                      # the headline's network is done: its weights go before the extras allocate
        del img_dev, out_dev
        torch.cuda.empty_cache()
        for key, leg in (("stages", lambda: stages_leg(ctx, world, rank, pk)), ("c5", lambda: c5_leg(ctx, world, rank))):
            ok = torch.ones(1, device="cuda")
            try:
                extras[key] = leg()
            except Exception as e:   # noqa: BLE001 -- the headline must survive a failing extra
                extras[key] = {"error": f"{type(e).__name__}: {e}"}
                ok.zero_()
            torch.cuda.synchronize()

    if rank == 0:
        patches = patches_of(size)
        algo_flops_step = flops_patch * patches
        tc_ms_step = tc_ms / args.steps
        exec_flops_step = tc_flops / args.steps
        # DRAM bytes of the GEMM kernels over one step, from the committed ncu pass of this same command
        # (a profiler number, so it is read, not re-measured here)
        traffic, traffic_note = None, "no committed ncu DRAM capture for this size"
        try:
            if args.size == 2048:
                name = "r2_dram_traffic.json" if os.path.exists(os.path.join(ROOT, "profiles", "r2_dram_traffic.json")) \
                    else "r1_dram_traffic.json"
                with open(os.path.join(ROOT, "profiles", name)) as f:
                    tr = json.load(f)
                traffic = tr["gemm_dram_bytes_per_step"]
                traffic_note = ("dram__bytes_read.sum + dram__bytes_write.sum summed over the GEMM launches of one step "
                                "(ncu, cache-control none; profiles/%s, captured with %d GEMM launches = 4 chunks of 338 "
                                "patches per step; the same work runs as 2 chunks of 676 now); all launches of the step: "
                                "%.1f GB" % (name, tr["gemm_launches"], tr["all_dram_bytes_per_step"] / 1e9))
        except (OSError, KeyError, ValueError):
            pass

        achieved = exec_flops_step / (tc_ms_step / 1e3) / 1e12
        effective = algo_flops_step / (tc_ms_step / 1e3) / 1e12
        line = {
            "metric": METRIC, "value": value, "unit": "MP/s", "n_gpus": world, "steps": args.steps, "warmup": warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": f"C4 per-image unit: one {size}x{size} float32 projection per rank per step -> "
                                   f"{patches} patches of 320x320 (ws 320, subdiv 2, 8x D4) through the random-init "
                                   "binary_segmentation UNet (64/128/256/512) + spline blend",
                       "l2": "working set per step (554 MB of patch probabilities + ~27 GB of activations per "
                             "676-patch chunk) exceeds the 126 MB L2; no explicit flush",
                       "parallelism": f"image-sharded x{world}"},
            "roofline": {"bound": "tensor", "achieved": achieved, "peak": pk["bf16_sustained"], "unit": "TFLOP/s",
                         "frac": achieved / pk["bf16_sustained"], "traffic": traffic, "traffic_note": traffic_note,
                         "basis": "EXECUTED tensor-pipe FLOPs (parity evaluation of the up path: 4 taps instead of 9) over "
                                  "the summed CUDA-event time of the GEMM launches",
                         "kernel": "the tcgen05 GEMM kernels of the UNet: conv_tc_kernel (implicit-GEMM convolution), "
                                   "conv_dwpw_kernel (fused SeparableConv2D), conv_halo_kernel (3x3, haloed tile), head_taps_kernel (head, taps in N)",
                         "executed_flops_per_step": exec_flops_step, "algorithmic_flops_per_step": algo_flops_step,
                         "effective_tflops": effective,
                         "effective_note": "algorithmic FLOPs (31.026 GFLOP per 320^2 patch as written in models.py) over "
                                           "the same kernel time: a speed-up figure, not divided by the hardware peak",
                         "frac_of_whole_step": exec_flops_step / (ms_step / 1e3) / 1e12 / pk["bf16_sustained"],
                         "kernel_ms_per_step": tc_ms_step, "launches_per_step": tc_launches / args.steps,
                         "kernel_share_of_step": tc_ms_step / ms_step, "peak_source": pk["source"] + " (sustained bf16)"},
            "e2e": {"value": e2e_value, "unit": "MP/s", "h2d_bytes_per_step": size * size * 4,
                    "d2h_bytes_per_step": size * size * 8, "steps": e2e_steps},
            "gpu_launches": int(launches), "clocks": sampler.result(),
            "image_metrics": {"foreground_fraction": [float(f.item()) for f in fracs]},
        }
        line.update(extras)
        if not args.no_cpu_baseline and world == 1:
            line["cpu_baseline"] = cpu_baseline_leg(size)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
