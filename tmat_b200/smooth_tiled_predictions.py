"""Sliding-window prediction with D4 augmentation and spline blending on the B200: drop-in for
``fl_tissue_model_tools/smooth_tiled_predictions.py`` (``predict_img_with_smooth_windowing:220``).

Two routes behind the same signature:

* ``pred_func`` is the ``predict`` of a :class:`tmat_b200.models.CudaUNet` (what
  ``UNetXceptionPatchSegmentor.model.predict`` is in this package): the fused CUDA pipeline
  (``tmat_stp_predict_host``) -- D4 folded into the stem's tile read, tcgen05 convolutions,
  one blend pass.
* any other callable ``(batch (n<=16, ws, ws) float32, verbose=0) -> (n, ws, ws, 1)``: tiles are
  gathered on the GPU per orientation, ``pred_func`` is called on the host in batches of 16 in
  the reference's order (:174-182), and the blend runs on the GPU in float64 in the reference's
  summation order, so the result is bit-identical to the reference for the same ``pred_func``.
"""
import numpy as np

from . import _ffi

INFERENCE_BATCH_SIZE = 16   # smooth_tiled_predictions.py:23

_WINDOW_CACHE = {}


def _spline_window(window_size, power=2):
    """Squared-spline window of smooth_tiled_predictions.py:26-41, normalised to mean 1."""
    n = np.arange(1, (window_size + 1) // 2 + 1)
    if window_size % 2 == 0:
        half = (2 * n - 1.0) / window_size
        tri = np.r_[half, half[::-1]]
    else:
        half = 2 * n / (window_size + 1.0)
        tri = np.r_[half, half[-2::-1]]
    q = int(window_size / 4)
    outer = (abs(2 * tri) ** power) / 2
    outer[q:-q] = 0
    inner = 1 - (abs(2 * (tri - 1)) ** power) / 2
    inner[:q] = 0
    inner[-q:] = 0
    wind = inner + outer
    return wind / np.average(wind)


def _window_1d(window_size):
    w = _WINDOW_CACHE.get(window_size)
    if w is None:
        w = np.ascontiguousarray(_spline_window(window_size), dtype=np.float64)
        _WINDOW_CACHE[window_size] = w
    return w


def geometry(h, w, window_size, subdivisions):
    g = (_ffi.C.c_int * 8)()
    _ffi.check(_ffi.lib().tmat_stp_geometry(h, w, window_size, subdivisions, g))
    return {"aug": g[0], "PM": g[1], "PN": g[2], "step": g[3],
            "tiles_even": (g[4], g[5]), "tiles_odd": (g[6], g[7])}


def predict_img_with_smooth_windowing(input_img, window_size, subdivisions, pred_func):
    """Same contract as smooth_tiled_predictions.py:220-267: 2-D image in, float64 (H, W) out."""
    img = np.asarray(input_img)
    if img.ndim != 2:
        raise ValueError(f"input_img must be 2-D, got shape {img.shape}")
    img = _ffi.as_c(img, np.float32)
    h, w = img.shape
    ctx = _ffi.default_context()
    L = _ffi.lib()
    net = getattr(getattr(pred_func, "__self__", None), "_tmat_handle", None)
    # the result of the all-device path lives in page-locked memory (D2H at PCIe speed)
    out = _ffi.pinned_empty((h, w), np.float64) if net is not None else np.empty((h, w), dtype=np.float64)
    if net is not None:
        owner = pred_func.__self__
        if owner.patch != window_size:
            raise ValueError(f"window_size {window_size} != UNet patch size {owner.patch}")
        _ffi.check(L.tmat_stp_predict_host(ctx.handle, net, _ffi.ptr(img), h, w, subdivisions,
                                           _ffi.ptr(out)))
        return out

    geo = geometry(h, w, window_size, subdivisions)
    n_tiles = geo["tiles_even"][0] * geo["tiles_even"][1]
    tiles = np.empty((n_tiles, window_size, window_size), dtype=np.float32)
    preds = np.empty((8, n_tiles, window_size, window_size), dtype=np.float32)
    for g in range(8):
        _ffi.check(L.tmat_stp_gather_host(ctx.handle, _ffi.ptr(img), h, w, window_size, subdivisions,
                                          g, _ffi.ptr(tiles)))
        if n_tiles > INFERENCE_BATCH_SIZE:
            for s in range(0, n_tiles, INFERENCE_BATCH_SIZE):
                p = np.asarray(pred_func(tiles[s:s + INFERENCE_BATCH_SIZE], verbose=0))
                preds[g, s:s + INFERENCE_BATCH_SIZE] = p.reshape(-1, window_size, window_size)
        else:
            preds[g] = np.asarray(pred_func(tiles, verbose=0)).reshape(n_tiles, window_size, window_size)
    w1d = _window_1d(window_size)
    _ffi.check(L.tmat_stp_blend_host(ctx.handle, _ffi.ptr(preds), h, w, window_size, subdivisions,
                                     _ffi.ptr(w1d), _ffi.ptr(out)))
    return out


# ---- one large image sharded over ranks by row strips (BASELINE "large-well tiles"; no reduction) ----------
def strip_rows(h, rank, world):
    """Output rows [row0, row1) owned by ``rank``: contiguous strips, sizes differ by at most one row."""
    base, extra = divmod(h, world)
    row0 = rank * base + min(rank, extra)
    return row0, row0 + base + (1 if rank < extra else 0)


def patches_for_rows(h, w, window_size, subdivisions, row0, row1):
    """Ids of the (orientation, tile) patches whose window touches output rows [row0, row1) (host geometry)."""
    n = _ffi.C.c_int()
    _ffi.check(_ffi.lib().tmat_stp_rows_patches(h, w, window_size, subdivisions, row0, row1, None, 0, _ffi.C.byref(n)))
    ids = np.empty(n.value, np.int32)
    _ffi.check(_ffi.lib().tmat_stp_rows_patches(h, w, window_size, subdivisions, row0, row1, _ffi.ptr(ids), n.value,
                                                _ffi.C.byref(n)))
    return ids


def predict_rows_with_smooth_windowing(input_img, window_size, subdivisions, pred_func, row0, row1):
    """Rows [row0, row1) of ``predict_img_with_smooth_windowing`` (bit-identical to those rows of the full
    result), evaluating only the patches that touch them.  ``pred_func`` must be ``CudaUNet.predict``."""
    img = _ffi.as_c(np.asarray(input_img), np.float32)
    if img.ndim != 2:
        raise ValueError(f"input_img must be 2-D, got shape {img.shape}")
    owner = getattr(pred_func, "__self__", None)
    net = getattr(owner, "_tmat_handle", None)
    if net is None:
        raise TypeError("predict_rows_with_smooth_windowing needs the CUDA UNet's predict method")
    if owner.patch != window_size:
        raise ValueError(f"window_size {window_size} != UNet patch size {owner.patch}")
    h, w = img.shape
    out = np.empty((row1 - row0, w), dtype=np.float64)
    _ffi.check(_ffi.lib().tmat_stp_predict_rows_host(_ffi.default_context().handle, net, _ffi.ptr(img), h, w,
                                                     subdivisions, row0, row1, _ffi.ptr(out), None))
    return out


def predict_img_row_sharded(input_img, window_size, subdivisions, pred_func, rank=None, world=None):
    """``predict_img_with_smooth_windowing`` of ONE image over ``world`` ranks: rank r computes its row strip
    (owner computes; ~(ws/2) rows of overlap work per strip boundary, no data-path reduction) and the strips
    are all-gathered over the default process group (NCCL on GPUs) so every rank returns the full image."""
    import torch
    import torch.distributed as dist
    if world is None:
        world = dist.get_world_size() if dist.is_initialized() else 1
    if rank is None:
        rank = dist.get_rank() if dist.is_initialized() else 0
    h, w = np.asarray(input_img).shape
    row0, row1 = strip_rows(h, rank, world)
    mine = predict_rows_with_smooth_windowing(input_img, window_size, subdivisions, pred_func, row0, row1)
    if world == 1:
        return mine
    rows_max = -(-h // world)
    dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
    buf = torch.zeros((rows_max, w), dtype=torch.float64, device=dev)
    buf[: row1 - row0] = torch.from_numpy(mine).to(dev)
    parts = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(parts, buf)
    out = np.empty((h, w), np.float64)
    for r in range(world):
        a, b = strip_rows(h, r, world)
        out[a:b] = parts[r][: b - a].cpu().numpy()
    return out
