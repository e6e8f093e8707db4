"""Oracle (test infrastructure): Gaussian pre-filter + Otsu threshold + erosion.

Restates ``well_mask_generation.py:236-277`` (``auto_threshold_well``) of the reference
with the scikit-image 0.22 calls it makes restated on scipy.ndimage / numpy
(**parity unpinned** at the scikit-image boundary -- not installable here):

  skimage.filters.gaussian(image, sigma=1)  -> integer input scaled to [0,1] float64,
        float32 stays float32; ndi.gaussian_filter(mode='nearest', truncate=4.0)
  skimage.exposure.rescale_intensity(x, out_range=(0,255)) -> (clip(x)-min)/(max-min)*255
        in x's float type, then ``.astype(uint8)`` truncates (``:246``)
  skimage.filters.threshold_otsu            -> between-class variance argmax over the
        [min..max] integer histogram (float32 counts, as skimage 0.22 casts them)
  skimage.morphology.binary_erosion(disk(5)) -> ndi.binary_erosion(border_value=True)
"""
import numpy as np
from scipy import ndimage as ndi
from .morphology import disk, binary_erosion

_INT_SCALE = {np.dtype(np.uint8): 255.0, np.dtype(np.uint16): 65535.0}


def gaussian_kernel1d(sigma=1.0, truncate=4.0):
    """scipy.ndimage._filters._gaussian_kernel1d(sigma, 0, radius)."""
    radius = int(truncate * float(sigma) + 0.5)
    x = np.arange(-radius, radius + 1)
    phi = np.exp(-0.5 / (sigma * sigma) * x ** 2)
    return phi / phi.sum()


def gaussian(image, sigma=1.0):
    image = np.asarray(image)
    if image.dtype in _INT_SCALE:
        # skimage.util.dtype._convert (0.22): unsigned ints -> np.multiply(image, 1. / imax, dtype=float64)
        image = np.multiply(image, 1.0 / _INT_SCALE[image.dtype], dtype=np.float64)
    elif image.dtype == np.float16:
        image = image.astype(np.float32)
    elif image.dtype == bool:
        image = image.astype(np.float64)
    return ndi.gaussian_filter(image, sigma, mode="nearest", truncate=4.0)


def rescale_to_u8(x):
    """rescale_intensity(x, out_range=(0, 255)).astype(np.uint8) for a float image."""
    lo, hi = x.min(), x.max()
    y = np.clip(x, lo, hi)
    if lo != hi:
        y = (y - lo) / (hi - lo)
    return np.asarray(y * 255 + 0, dtype=x.dtype).astype(np.uint8)


def threshold_otsu_u8(img):
    lo, hi = int(img.min()), int(img.max())
    if lo == hi:
        return lo
    counts = np.bincount(img.ravel() - lo, minlength=hi - lo + 1).astype(np.float32)
    centers = np.arange(lo, hi + 1)
    w1 = np.cumsum(counts)
    w2 = np.cumsum(counts[::-1])[::-1]
    m1 = np.cumsum(counts * centers) / w1
    m2 = (np.cumsum((counts * centers)[::-1]) / w2[::-1])[::-1]
    var12 = w1[:-1] * w2[1:] * (m1[:-1] - m2[1:]) ** 2
    return int(centers[np.argmax(var12)])


def corner_invert_decision(u8):
    """well_mask_generation.py:248-270: True when the image must be inverted."""
    h, w = u8.shape
    xs, xe = int(h * 0.05), int(h * 0.95)
    ys, ye = int(w * 0.05), int(w * 0.95)
    meds = [np.median(u8[:xs, :ys]), np.median(u8[:xs, ye:]),
            np.median(u8[xe:, :ys]), np.median(u8[xe:, ye:])]
    d_min = np.abs(u8.min() - min(meds))
    d_max = np.abs(u8.max() - max(meds))
    return bool(d_min > d_max)


def auto_threshold_well(image, return_stages=False):
    blur = gaussian(image, 1.0)
    u8 = rescale_to_u8(blur)
    inv = corner_invert_decision(u8)
    if inv:
        u8 = 255 - u8
    thr = threshold_otsu_u8(u8)
    mask = binary_erosion(u8 >= thr, disk(5))
    if return_stages:
        return mask, {"blur": blur, "u8": u8, "invert": inv, "otsu": thr}
    return mask


# ---- superellipse fit / hull mask: well_mask_generation.py:16-140 (numpy-only in the reference) ------------------
# PINNED: tests/golden/wellmask.npz holds outputs of the reference's own functions (imported by
# tests/golden/make_golden.py with np.random.seed set); these restatements are checked against them on the CPU.
_SE_BOUNDS = np.array([(-np.pi / 20, np.pi / 20), (0.67, 1.33), (0.9, 1.1), (0.9, 1.1), (-0.3, 0.3), (-0.3, 0.3)])


def get_superellipse_hull(x, y, n, num_iters=25000):
    """:16-91.  Draws from numpy's global RNG, like the reference."""
    from scipy.special import gamma
    w = np.random.rand(num_iters, 6)
    pv = (_SE_BOUNDS[:, 1] - _SE_BOUNDS[:, 0]) * w + _SE_BOUNDS[:, 0]
    t, d, s_a, s_b, c_x, c_y = pv.T[..., np.newaxis]
    if n == 2:
        val = ((x - c_x) / (d * s_a)) ** 2 + ((y - c_y) / (d * s_b)) ** 2
    elif n % 2 == 0:
        val = ((((x - c_x) * np.cos(t) - ((y - c_y) * np.sin(t))) / (d * s_a)) ** n
               + (((x - c_x) * np.sin(t) + (y - c_y) * np.cos(t)) / (d * s_b)) ** n)
    else:
        val = (np.abs(((x - c_x) * np.cos(t) - ((y - c_y) * np.sin(t))) / (d * s_a)) ** n
               + np.abs(((x - c_x) * np.sin(t) + (y - c_y) * np.cos(t)) / (d * s_b)) ** n)
    cand = np.where(np.max(val, axis=1) < 1)[0]
    area = 4 * d[cand] ** 2 * s_a[cand] * s_b[cand] * gamma(1 + 1 / n) ** 2 / gamma(1 + 2 / n)
    k = cand[np.argmin(area)]            # ValueError on an empty candidate set, as in the reference
    return tuple(pv[k])


def gen_superellipse_mask(t, d, s_a, s_b, c_x, c_y, n, shape):
    """:94-118."""
    X, Y = np.meshgrid(np.linspace(-1, 1, shape[0]), np.linspace(-1, 1, shape[1]))
    mask = ((np.abs(((X - c_x) * np.cos(t) - (Y - c_y) * np.sin(t)) / (d * s_a))) ** n
            + (np.abs(((X - c_x) * np.sin(t) + (Y - c_y) * np.cos(t)) / (d * s_b))) ** n < 1)
    return np.swapaxes(mask, 0, 1)


def create_convex_hull_mask(array_shape, hull_vertices):
    """:121-140 -- scipy's Delaunay, as the reference."""
    from scipy.spatial import Delaunay
    coords = np.indices(array_shape).reshape(2, -1).T
    return (Delaunay(hull_vertices).find_simplex(coords) >= 0).reshape(array_shape)
