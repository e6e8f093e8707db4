#!/usr/bin/env python
"""bench.py -- megapixels/s segmented (tiled UNet + 8x D4 + spline blend) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--size 2048]

One "step" = one synthetic 2048x2048 float32 Z-projection (BASELINE.json config C4's per-image
unit: 13x13 windows x 8 orientations = 1352 patches of 320x320) through
``predict_img_with_smooth_windowing`` with the random-init binary-segmentation UNet.  With N
ranks (torchrun) every rank segments its own image per step (images are independent,
scripts/compute_branches.py:585): weak scaling, no collective on the data path; NCCL carries only
the timing reduction and the final gather of per-image metrics.

The JSON line carries `roofline` (tensor-pipe: the tcgen05 convolution kernel, timed per launch
with CUDA events inside the timed region), `cpu_baseline` (oracle port on the host cores, bounded
sample), `e2e` (same call with pinned host buffers, H2D/D2H inside), `clocks`, `gpu_launches`.
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


def cpu_arm(steps, warmup, size, patches_per_step=16):
    """The reference's CPU path for the same workload, bounded: the oracle UNet (torch-CPU fp32,
    all host threads, batches of 16 as smooth_tiled_predictions.py:174-182) + window multiply and
    overlap-add for the sampled tiles.  Returns (MP/s, seconds per step, description)."""
    import torch
    from oracle import unet as ou, stp as ostp
    cores = len(os.sched_getaffinity(0))
    torch.set_num_threads(cores)
    w = ou.make_weights(FILTERS, seed=0)
    pred = ou.make_pred_func(w)
    rng = np.random.default_rng(0)
    tiles = rng.random((patches_per_step, WS, WS), dtype=np.float32)
    w2 = ostp.window_2d(WS)
    canvas = np.zeros((WS + (patches_per_step - 1) * (WS // SUBDIV), WS))

    def step():
        p = pred(tiles, verbose=0)
        p = np.array([q * w2 for q in p])
        for i in range(patches_per_step):
            canvas[i * 160:i * 160 + WS, :] += p[i, :, :, 0]

    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / steps
    n_r = (size + 2 * 160 - WS) // 160 + 1
    patches_per_image = 8 * n_r * n_r
    mp_per_patch = size * size / 1e6 / patches_per_image
    value = patches_per_step * mp_per_patch / dt
    sample = (f"{patches_per_step} of the {patches_per_image} 320x320 patches of one {size}x{size} image per step "
              f"(oracle UNet torch-CPU fp32 + window/overlap-add), extrapolated linearly")
    return value, dt, cores, sample


def reference_arm(args, rank):
    if rank != 0:
        return
    value, dt, cores, sample = cpu_arm(args.steps, min(args.warmup, 1), args.size)
    line = {"metric": METRIC, "value": value, "unit": "MP/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": min(args.warmup, 1), "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "impl": "reference",
            "config": {"workload": f"tiled UNet+D4+blend, {args.size}x{args.size} image, ws 320, subdiv 2 "
                                   "(CPU sample, see cpu_baseline.sample)"},
            "cpu_baseline": {"value": value, "unit": "MP/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "MP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "note": "TensorFlow/Keras (the reference's runtime) is not installable offline; the reference arm is the "
                    "oracle port of smooth_tiled_predictions.py + models.py on the host cores"}
    print(json.dumps(line), flush=True)


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
        reference_arm(args, rank)
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

    if rank == 0:
        pk = peaks()
        flops_patch = net.flops_per_patch()
        n_r = (size + 2 * 160 - WS) // 160 + 1
        patches = 8 * n_r * n_r
        algo_flops_step = flops_patch * patches
        tc_ms_step = tc_ms / args.steps
        # DRAM bytes of the GEMM kernels over one step, from the committed ncu pass of this same command
        # (profiles/r1_dram_traffic.json; a profiler number, so it is read, not re-measured here)
        traffic, traffic_note = None, "no committed ncu DRAM capture for this size"
        try:
            if args.size == 2048:
                with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r1_dram_traffic.json")) as f:
                    tr = json.load(f)
                traffic = tr["gemm_dram_bytes_per_step"]
                traffic_note = ("dram__bytes_read.sum + dram__bytes_write.sum summed over the %d GEMM launches of one step "
                                "(ncu, cache-control none); all %d launches of the step: %.1f GB" %
                                (tr["gemm_launches"], tr["launches_in_step"], tr["all_dram_bytes_per_step"] / 1e9))
        except (OSError, KeyError, ValueError):
            pass

        achieved = algo_flops_step / (tc_ms_step / 1e3) / 1e12
        line = {
            "metric": METRIC, "value": value, "unit": "MP/s", "n_gpus": world, "steps": args.steps, "warmup": warmup,
            "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": f"C4 per-image unit: one {size}x{size} float32 projection per rank per step -> "
                                   f"{patches} patches of 320x320 (ws 320, subdiv 2, 8x D4) through the random-init "
                                   "binary_segmentation UNet (64/128/256/512) + spline blend",
                       "l2": "working set per step (554 MB of patch probabilities + ~13 GB of activations per "
                             "338-patch chunk) exceeds the 126 MB L2; no explicit flush",
                       "parallelism": f"image-sharded x{world}"},
            "roofline": {"bound": "tensor", "achieved": achieved, "peak": pk["bf16_sustained"], "unit": "TFLOP/s",
                         "frac": achieved / pk["bf16_sustained"], "traffic": traffic, "traffic_note": traffic_note,
                         "kernel": "the tcgen05 GEMM kernels of the UNet: conv_tc_kernel (implicit-GEMM convolution), "
                                   "conv_dwpw_kernel (fused SeparableConv2D), conv_halo_kernel (3x3, haloed tile)",
                         "algorithmic_flops_per_step": algo_flops_step, "executed_flops_per_step": tc_flops / args.steps,
                         "kernel_ms_per_step": tc_ms_step, "launches_per_step": tc_launches / args.steps,
                         "kernel_share_of_step": tc_ms_step / ms_step, "peak_source": pk["source"] + " (sustained bf16)"},
            "e2e": {"value": e2e_value, "unit": "MP/s", "h2d_bytes_per_step": size * size * 4,
                    "d2h_bytes_per_step": size * size * 8, "steps": e2e_steps},
            "gpu_launches": int(launches), "clocks": sampler.result(),
            "image_metrics": {"foreground_fraction": [float(f.item()) for f in fracs]},
        }
        if not args.no_cpu_baseline:
            v, sdt, cores, sample = cpu_arm(1, 1, size)
            line["cpu_baseline"] = {"value": v, "unit": "MP/s", "cores": cores, "kind": "port", "sample": sample}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
