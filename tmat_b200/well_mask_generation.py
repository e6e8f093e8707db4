"""Well-mask generation on the B200: drop-in for the reference's ``fl_tissue_model_tools/well_mask_generation.py``.

* ``auto_threshold_well`` (:236-277) -- the heavy, full-resolution part -- runs on the GPU end to end;
* ``get_superellipse_hull`` (:16-91), ``gen_superellipse_mask`` (:94-118), ``create_convex_hull_mask`` (:121-140):
  the random parameter sets are drawn here with numpy exactly as the reference draws them (so ``np.random.seed`` decides
  the same draw), the 25000 x P evaluation / the per-pixel tests run on the GPU;
* ``generate_well_mask`` (:143-233) composes them as the reference does.  Its glue on the <= 200-pixel thumbnail
  (``rescale``, ``canny``, ``resize``, ``draw.disk``) is scikit-image's and ``ConvexHull`` is scipy's, as in the
  reference: they are imported when the function is called.
"""
import ctypes as C
import math

import numpy as np

from . import _ffi


def gaussian(image, sigma=1.0):
    """skimage.filters.gaussian(image, sigma) as called at well_mask_generation.py:245."""
    img = np.asarray(image)
    if img.dtype == np.float16:
        img = img.astype(np.float32)
    if img.dtype not in _ffi.DTYPES:
        raise TypeError(f"unsupported dtype {img.dtype}")
    img = np.ascontiguousarray(img)
    out = np.empty(img.shape, dtype=np.float32 if img.dtype == np.float32 else np.float64)
    _ffi.check(_ffi.lib().tmat_gaussian_host(_ffi.default_context().handle, _ffi.ptr(img), _ffi.DTYPES[img.dtype],
                                             *img.shape, float(sigma), _ffi.ptr(out)))
    return out


def auto_threshold_well(image, return_info=False):
    """well_mask_generation.py:236-277 -> boolean mask."""
    img = np.asarray(image)
    if img.ndim != 2:
        raise ValueError(f"expected a 2-D image, got shape {img.shape}")
    if img.dtype == np.float16:
        img = img.astype(np.float32)
    if img.dtype not in _ffi.DTYPES:
        raise TypeError(f"unsupported dtype {img.dtype}")
    img = np.ascontiguousarray(img)
    mask = np.empty(img.shape, dtype=np.uint8)
    info = (C.c_int * 4)() if return_info else None          # asking for it costs a device round trip
    _ffi.check(_ffi.lib().tmat_auto_threshold_well_host(_ffi.default_context().handle, _ffi.ptr(img),
                                                        _ffi.DTYPES[img.dtype], *img.shape, _ffi.ptr(mask), info))
    mask = mask.view(np.bool_)                                # the kernel writes 0 / 1: no copy
    if return_info:
        return mask, {"invert": bool(info[0]), "otsu": info[1], "u8_min": info[2], "u8_max": info[3]}
    return mask


# ---- superellipse fit / hull mask (well_mask_generation.py:16-140) ------------------------------------------------
_BOUNDS = ((-np.pi / 20, np.pi / 20), (0.67, 1.33), (0.9, 1.1), (0.9, 1.1), (-0.3, 0.3), (-0.3, 0.3))


def _gamma(v):
    try:
        from scipy.special import gamma              # what the reference calls (:11)
        return float(gamma(v))
    except ImportError:                               # pragma: no cover -- scipy is a dependency of the reference
        return math.gamma(v)


def get_superellipse_hull(x, y, n, num_iters=25000):
    """well_mask_generation.py:16-91 -> (t, d, s_a, s_b, c_x, c_y) of the smallest random superellipse that encloses
    the points.  Raises ValueError when no parameter set does, as the reference's ``np.argmin`` of an empty array."""
    x = np.ascontiguousarray(x, dtype=np.float64).ravel()
    y = np.ascontiguousarray(y, dtype=np.float64).ravel()
    if x.size != y.size or x.size == 0:
        raise ValueError("x and y must be non-empty and of equal length")
    linear_weights = np.random.rand(num_iters, 6)                                    # :35
    bounds = np.array(_BOUNDS)
    param_values = np.ascontiguousarray((bounds[:, 1] - bounds[:, 0]) * linear_weights + bounds[:, 0])   # :47
    cos_t = sin_t = None
    if n != 2:
        cos_t = np.ascontiguousarray(np.cos(param_values[:, 0]))
        sin_t = np.ascontiguousarray(np.sin(param_values[:, 0]))
    best = C.c_int(-1)
    _ffi.check(_ffi.lib().tmat_superellipse_hull_host(
        _ffi.default_context().handle, _ffi.ptr(x), _ffi.ptr(y), int(x.size), int(n), _ffi.ptr(param_values),
        _ffi.ptr(cos_t), _ffi.ptr(sin_t), int(num_iters), _gamma(1 + 1 / n) ** 2, _gamma(1 + 2 / n), C.byref(best)))
    if best.value < 0:
        raise ValueError("attempt to get argmin of an empty sequence")
    return tuple(param_values[best.value])


def gen_superellipse_mask(t, d, s_a, s_b, c_x, c_y, n, shape):
    """well_mask_generation.py:94-118 -> boolean mask of ``shape``."""
    xs = np.ascontiguousarray(np.linspace(-1, 1, shape[0]))
    ys = np.ascontiguousarray(np.linspace(-1, 1, shape[1]))
    out = np.empty((int(shape[0]), int(shape[1])), dtype=np.uint8)
    _ffi.check(_ffi.lib().tmat_superellipse_mask_host(
        _ffi.default_context().handle, _ffi.ptr(xs), int(shape[0]), _ffi.ptr(ys), int(shape[1]), float(c_x), float(c_y),
        float(np.cos(t)), float(np.sin(t)), float(d * s_a), float(d * s_b), int(n), _ffi.ptr(out)))
    return out.view(np.bool_)


def _hull_order(points):
    """Vertices of the convex hull of integer points in polygon order (Andrew's monotone chain, exact)."""
    pts = sorted(set((int(r), int(c)) for r, c in points))

    def half(seq):
        h = []
        for p in seq:
            while len(h) >= 2 and ((h[-1][0] - h[-2][0]) * (p[1] - h[-2][1]) - (h[-1][1] - h[-2][1]) * (p[0] - h[-2][0])) <= 0:
                h.pop()
            h.append(p)
        return h

    lower, upper = half(pts), half(pts[::-1])
    return lower[:-1] + upper[:-1]


def create_convex_hull_mask(array_shape, hull_vertices):
    """well_mask_generation.py:121-140 -> boolean mask: pixels inside or on the convex hull of ``hull_vertices``
    ((row, col) pixel coordinates)."""
    hv = np.asarray(hull_vertices)
    if hv.ndim != 2 or hv.shape[1] != 2 or not np.issubdtype(hv.dtype, np.integer):
        raise TypeError("hull_vertices must be an (N, 2) integer array of pixel coordinates")
    poly = _hull_order(hv)
    if len(poly) < 3:
        try:
            from scipy.spatial import QhullError
        except ImportError:                           # pragma: no cover
            QhullError = RuntimeError
        raise QhullError("the hull vertices are collinear: no 2-D triangulation (scipy's Delaunay raises here too)")
    verts = np.ascontiguousarray(poly, dtype=np.int32)
    H, W = int(array_shape[0]), int(array_shape[1])
    out = np.empty((H, W), dtype=np.uint8)
    _ffi.check(_ffi.lib().tmat_convex_hull_mask_host(_ffi.default_context().handle, H, W, _ffi.ptr(verts), len(poly),
                                                     _ffi.ptr(out)))
    return out.view(np.bool_)


def generate_well_mask(image, mask_val=1, return_superellipse_params=False):
    """well_mask_generation.py:143-233, same steps in the same order; the threshold, the hull mask and the
    superellipse search / mask run on the GPU."""
    import traceback
    try:
        from skimage.feature import canny
        from skimage.transform import rescale, resize
        from skimage.draw import disk as circle
        from scipy.spatial import ConvexHull
    except ImportError as e:
        raise ImportError("generate_well_mask: the thumbnail glue (rescale, canny, resize, draw.disk, ConvexHull) is "
                          f"scikit-image's / scipy's, as in the reference ({e})") from e
    im_thresh = auto_threshold_well(image)
    downsamp_ratio = min(1, 200 / np.max(im_thresh.shape))
    im_thresh = rescale(im_thresh, downsamp_ratio, order=0, preserve_range=True)

    def with_edges(border, src):
        border[0, :] += src[0, :]
        border[-1, :] += src[-1, :]
        border[:, 0] += src[:, 0]
        border[:, -1] += src[:, -1]
        return border

    im_thresh_border = with_edges(canny(im_thresh), im_thresh)

    def get_circ_mask():
        circ_mask = np.zeros(image.shape, dtype=np.uint8)
        center = image.shape[0] // 2, image.shape[1] // 2
        radius = int(image.shape[0] * 0.5 * (1 - 0.95))
        rr, cc = circle((center[0], center[1]), radius, shape=image.shape)
        circ_mask[rr, cc] = mask_val
        return circ_mask

    border_points = np.argwhere(im_thresh_border)
    try:
        hull = ConvexHull(border_points)
    except ValueError:
        return get_circ_mask()
    hull_vertices = border_points[hull.vertices]
    well_mask = create_convex_hull_mask(im_thresh.shape, hull_vertices)
    well_mask_border = with_edges(canny(well_mask), well_mask)
    area = np.sum(well_mask)
    perimeter = np.sum(well_mask_border)
    n = 8 if perimeter / area > .027 else 2
    x = hull_vertices[:, 0] / im_thresh.shape[0] * 2 - 1
    y = hull_vertices[:, 1] / im_thresh.shape[1] * 2 - 1
    found_superellipse = False
    try:
        t, d, s_a, s_b, c_x, c_y = get_superellipse_hull(x, y, n)
        d *= 0.9
        well_mask = gen_superellipse_mask(t, d, s_a, s_b, c_x, c_y, n, im_thresh.shape)
        found_superellipse = True
    except Exception:   # noqa: BLE001 -- the reference's own fallback (:218-220)
        traceback.print_exc()
        print("Falling back to convex hull well mask.", flush=True)
    well_mask = well_mask.astype(np.uint8) * mask_val
    well_mask = resize(well_mask, image.shape, order=0, preserve_range=True)
    if found_superellipse and return_superellipse_params:
        return well_mask, t, d, s_a, s_b, c_x, c_y, n
    return well_mask
