#!/usr/bin/env python
"""BASELINE configs C2, C3 and C4 -- the bandwidth-bound stages either side of the tiled UNet, on one B200.

    python scripts/bench_stages.py [--config C2|C3|C4|all] [--size 2048] [--z 32] [--iters 10] [--no-cpu]

  C2  the five Z projections of a synthetic (32, 2048, 2048) uint16 stack (`zstacks.py:153-249`)
  C3  stages 3/4 on a synthetic 2048x2048 image: skimage Gaussian, auto_threshold_well (Gaussian + uint8 rescale +
      Otsu + disk(5) erosion, `well_mask_generation.py:236-277`), disk(2) median, 8-connected labelling,
      skeletonisation and the whole filter_branch_seg_mask (`transforms.py:306-361`)
  C4  tmat_branch_mask_dev: stack -> max projection -> [0,1] rescale -> tiled UNet + D4 + blend -> > 0.5 -> cleanup
      (`compute_branches.py:228, 316-337`) for one 2048x2048 field

Every stage is called through the C ABI with device pointers, timed with CUDA events on the launching stream over
`--iters` calls after 3 warm-up calls, one event pair per call, median reported (C2's 268 MB stack exceeds the
126 MB L2; the 2048^2 images of C3 are 4-32 MB and would stay L2-resident between calls, so a 256 MB memset flushes
L2 before each timed call there, outside the event pair).  One JSON line per stage: ms, MP/s, algorithmic GB/s (bytes = the stage's compulsory reads + writes, stated in `bytes`) and the fraction of the
measured HBM copy bandwidth (MEASURED_PEAKS.json); `cpu_ms` is the oracle (numpy / scipy / cv2 restatement) on the host.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def hbm_peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "MEASURED_PEAKS.json"
    except Exception:
        return 6550.0, "fallback (B200_PROFILING.md)"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="all")
    ap.add_argument("--size", type=int, default=2048)
    ap.add_argument("--z", type=int, default=32)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    import torch
    import helpers
    from tmat_b200 import _ffi, models
    torch.cuda.set_device(0)
    ctx = _ffi.Context(0)
    _ffi._tls.ctx = ctx
    L = _ffi.lib()
    C = _ffi.C
    stream = torch.cuda.current_stream()
    sptr = C.c_void_p(stream.cuda_stream)
    peak, peak_src = hbm_peak()
    S, Z = args.size, args.z
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def timed(fn, iters, flush):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
        for e0, e1 in evs:
            if flush:
                flush_buf.zero_()
            e0.record(stream)
            fn()
            e1.record(stream)
        torch.cuda.synchronize()
        ts = sorted(e0.elapsed_time(e1) for e0, e1 in evs)
        return ts[len(ts) // 2], ts[0]

    def cpu_time(fn):
        if args.no_cpu:
            return None
        t0 = time.perf_counter()
        fn()
        return (time.perf_counter() - t0) * 1e3

    def emit(config, stage, ms, best, nbytes, bytes_desc, cpu_ms, extra=None):
        gbs = nbytes / (ms * 1e-3) / 1e9
        line = {"config": config, "stage": stage, "ms": round(ms, 4), "ms_best": round(best, 4),
                "MP/s": round(S * S / 1e6 / (ms * 1e-3), 1), "bytes": bytes_desc,
                "algorithmic_GB/s": round(gbs, 1), "hbm_peak_GB/s": peak, "frac_of_hbm_peak": round(gbs / peak, 3),
                "peak_source": peak_src, "cpu_ms": None if cpu_ms is None else round(cpu_ms, 1),
                "cpu_cores": len(os.sched_getaffinity(0)),
                "speedup_vs_cpu": None if cpu_ms is None else round(cpu_ms / ms, 1)}
        if extra:
            line.update(extra)
        print(json.dumps(line), flush=True)

    want = args.config.upper()
    if want in ("C2", "ALL"):
        from oracle import zstacks as oz
        stack = helpers.synth_stack(Z, S, S, 2000)
        d_stack = torch.from_numpy(stack.view(np.int16)).cuda()
        for name in ("max", "min", "avg", "med", "fs"):
            wide = name in ("avg", "med")
            d_out = torch.empty((S, S), dtype=torch.float64 if wide else torch.int16, device="cuda")

            def fn():
                _ffi.check(L.tmat_zproj_dev(ctx.handle, _ffi.ptr(d_stack), _ffi.DTYPES[np.dtype(np.uint16)], Z, S, S,
                                            _ffi.ZPROJ[name], 5, _ffi.ptr(d_out), sptr))
            ms, best = timed(fn, args.iters, flush=False)
            got = d_out.cpu().numpy()
            got = got if wide else got.view(np.uint16)
            t0 = time.perf_counter()
            ref = oz.METHODS[name](stack)
            cpu_ms = None if args.no_cpu else (time.perf_counter() - t0) * 1e3
            nbytes = Z * S * S * 2 + S * S * (8 if wide else 2)
            emit("C2", "proj_" + name, ms, best, nbytes, f"{Z}*2 B read + {8 if wide else 2} B written per pixel",
                 cpu_ms, {"bit_exact_vs_oracle": bool(np.array_equal(got, ref)), "stack": [Z, S, S], "dtype": "u16"})
        del d_stack

    if want in ("C3", "ALL"):
        from oracle import wellmask as ow, morphology as om
        well = helpers.synth_well(S, S, 3000)
        d_well = torch.from_numpy(well.view(np.int16)).cuda()
        d_g = torch.empty((S, S), dtype=torch.float64, device="cuda")
        d_m = torch.empty((S, S), dtype=torch.uint8, device="cuda")

        def f_gauss():
            _ffi.check(L.tmat_gaussian_dev(ctx.handle, _ffi.ptr(d_well), 1, S, S, C.c_double(1.0), _ffi.ptr(d_g), sptr))
        ms, best = timed(f_gauss, args.iters, True)
        ref = ow.gaussian(well, 1.0) if not args.no_cpu else None
        cpu_ms = cpu_time(lambda: ow.gaussian(well, 1.0))
        emit("C3", "gaussian_sigma1_u16_to_f64", ms, best, S * S * 10, "2 B read + 8 B written per pixel", cpu_ms,
             {"bit_exact_vs_oracle": None if ref is None else bool(np.array_equal(d_g.cpu().numpy(), ref))})

        info = (C.c_int * 4)()

        def f_atw():
            _ffi.check(L.tmat_auto_threshold_well_dev(ctx.handle, _ffi.ptr(d_well), 1, S, S, _ffi.ptr(d_m), info, sptr))
        ms, best = timed(f_atw, args.iters, True)
        ref = ow.auto_threshold_well(well) if not args.no_cpu else None
        cpu_ms = cpu_time(lambda: ow.auto_threshold_well(well))
        emit("C3", "auto_threshold_well", ms, best, S * S * 3, "2 B read + 1 B written per pixel (floor; intermediates "
             "f64 image + u8 image are written and re-read)", cpu_ms,
             {"bit_exact_vs_oracle": None if ref is None else bool(np.array_equal(d_m.cpu().numpy() != 0, ref != 0)),
              "host_syncs": "0 (Otsu and the corner medians are decided on the device)"})

        base = helpers.synth_mask(min(S, 512), min(S, 512), 3001)      # the generator is O(minutes) at 2048^2: mosaic
        reps = -(-S // base.shape[0])
        mask = np.tile(base, (reps, reps))[:S, :S].copy()
        d_mask = torch.from_numpy(mask.astype(np.uint8)).cuda()
        d_o = torch.empty((S, S), dtype=torch.uint8, device="cuda")
        d_lab = torch.empty((S, S), dtype=torch.int32, device="cuda")
        nlab = C.c_int(0)
        iters_out = C.c_int(0)

        def f_med():
            _ffi.check(L.tmat_median_disk2_dev(ctx.handle, _ffi.ptr(d_mask), S, S, _ffi.ptr(d_o), sptr))

        def f_lab():
            _ffi.check(L.tmat_label_dev(ctx.handle, _ffi.ptr(d_mask), S, S, 2, _ffi.ptr(d_lab), C.byref(nlab), sptr))

        def f_skel():
            _ffi.check(L.tmat_skeletonize_dev(ctx.handle, _ffi.ptr(d_mask), S, S, _ffi.ptr(d_o), C.byref(iters_out), sptr))

        def f_filter():
            _ffi.check(L.tmat_filter_branch_seg_mask_dev(ctx.handle, _ffi.ptr(d_mask), S, S, 1, 1, _ffi.ptr(d_o), sptr))

        mb = mask.astype(bool)
        for stage, fn, nbytes, desc, cpu_fn, cmp in (
            ("median_disk2", f_med, S * S * 2, "1 B read + 1 B written per pixel",
             lambda: om.median_bool(mb, om.disk(2)), lambda r: np.array_equal(d_o.cpu().numpy() != 0, r != 0)),
            ("label_8conn", f_lab, S * S * 5, "1 B read + 4 B written per pixel (floor)",
             lambda: om.label(mb, 2), lambda r: np.array_equal(d_lab.cpu().numpy(), r[0] if isinstance(r, tuple) else r)),
            ("skeletonize", f_skel, S * S * 2, "1 B read + 1 B written per pixel (floor; one pass per thinning round)",
             lambda: om.skeletonize(mb), lambda r: np.array_equal(d_o.cpu().numpy() != 0, r != 0)),
            ("filter_branch_seg_mask", f_filter, S * S * 2, "1 B read + 1 B written per pixel (floor)",
             lambda: om.filter_branch_seg_mask(mb), lambda r: np.array_equal(d_o.cpu().numpy() != 0, r != 0)),
        ):
            ms, best = timed(fn, args.iters, True)
            if args.no_cpu:
                cpu_ms, ok = None, None
            else:
                t0 = time.perf_counter()
                r = cpu_fn()
                cpu_ms = (time.perf_counter() - t0) * 1e3
                fn()
                torch.cuda.synchronize()
                ok = bool(cmp(r))
            extra = {"bit_exact_vs_oracle": ok}
            if stage == "label_8conn":
                extra["components"] = int(nlab.value)
            if stage == "skeletonize":
                extra["thinning_rounds"] = int(iters_out.value)
            emit("C3", stage, ms, best, nbytes, desc, cpu_ms, extra)

    if want in ("C4", "ALL"):
        from bench import FILTERS
        ws = 320
        net = models.CudaUNet(models.random_init_weights(FILTERS, seed=0), ws, FILTERS, impl="tc", ctx=ctx)
        stack = helpers.synth_stack(Z, S, S, 4000)
        d_stack = torch.from_numpy(stack.view(np.int16)).cuda()
        d_pred = torch.empty((S, S), dtype=torch.float64, device="cuda")
        d_mask = torch.empty((S, S), dtype=torch.uint8, device="cuda")

        def f_path():
            _ffi.check(L.tmat_branch_mask_dev(ctx.handle, net._tmat_handle, _ffi.ptr(d_stack), Z, S, S, 2,
                                              _ffi.ptr(d_pred), _ffi.ptr(d_mask), sptr))
        ms, best = timed(f_path, max(3, args.iters // 2), False)
        fg = float(d_mask.float().mean().item())
        emit("C4", "branch_mask (zproj max -> rescale -> tiled UNet x8 D4 -> blend -> >0.5 -> cleanup)", ms, best,
             Z * S * S * 2 + S * S, f"{Z}*2 B read + 1 B written per pixel (stage floor; the UNet is tensor-bound)", None,
             {"foreground_fraction": round(fg, 4), "note": "UNet share and its roofline: bench.py"})


if __name__ == "__main__":
    main()
