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
    out = np.empty((h, w), dtype=np.float64)

    net = getattr(getattr(pred_func, "__self__", None), "_tmat_handle", None)
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
