"""Mask cleanup on the B200: drop-in for the inference-side functions of the reference's
``fl_tissue_model_tools/transforms.py`` (``remove_small_islands:170``, ``regionprops_image:291``,
``filter_branch_seg_mask:306``) plus the scikit-image primitives they are built from, all running in
``csrc/morph.cu``.  Integer work, bit-exact against the reference."""
import ctypes as C

import numpy as np

from . import _ffi


def disk(radius):
    """skimage.morphology.disk."""
    y, x = np.mgrid[-radius:radius + 1, -radius:radius + 1]
    return (x * x + y * y <= radius * radius).astype(np.uint8)


def _mask_u8(mask):
    m = np.asarray(mask)
    if m.ndim != 2:
        raise ValueError(f"expected a 2-D mask, got shape {m.shape}")
    return np.ascontiguousarray(m != 0, dtype=np.uint8)


def _is_disk(footprint, r):
    fp = np.asarray(footprint)
    return fp.shape == (2 * r + 1, 2 * r + 1) and np.array_equal(fp != 0, disk(r) != 0)


def median_filter(mask, footprint=None):
    """skimage.filters.median(bool mask, disk(2)) (transforms.py:323)."""
    if footprint is not None and not _is_disk(footprint, 2):
        raise NotImplementedError("libtmat_b200 implements the reference's disk(2) median footprint")
    m = _mask_u8(mask)
    out = np.empty_like(m)
    _ffi.check(_ffi.lib().tmat_median_disk2_host(_ffi.default_context().handle, _ffi.ptr(m), *m.shape, _ffi.ptr(out)))
    return out.astype(bool)


def binary_erosion(mask, footprint):
    """skimage.morphology.binary_erosion(mask, disk(r)) (well_mask_generation.py:275)."""
    fp = np.asarray(footprint)
    r = (fp.shape[0] - 1) // 2
    if not _is_disk(fp, r):
        raise NotImplementedError("libtmat_b200 erodes with disk footprints")
    m = _mask_u8(mask)
    out = np.empty_like(m)
    _ffi.check(_ffi.lib().tmat_erode_disk_host(_ffi.default_context().handle, _ffi.ptr(m), *m.shape, r, 1, _ffi.ptr(out)))
    return out.astype(bool)


def measure_label(mask, connectivity=2, return_num=False):
    """skimage.measure.label for 2-D input: labels numbered in raster order of first pixel."""
    m = _mask_u8(mask)
    lab = np.empty(m.shape, dtype=np.int32)
    n = C.c_int()
    _ffi.check(_ffi.lib().tmat_label_host(_ffi.default_context().handle, _ffi.ptr(m), *m.shape, int(connectivity),
                                          _ffi.ptr(lab), C.byref(n)))
    return (lab, n.value) if return_num else lab


def skeletonize(mask, return_iterations=False):
    """skimage.morphology.skeletonize, 2-D (transforms.py:331)."""
    m = _mask_u8(mask)
    out = np.empty_like(m)
    it = C.c_int()
    _ffi.check(_ffi.lib().tmat_skeletonize_host(_ffi.default_context().handle, _ffi.ptr(m), *m.shape, _ffi.ptr(out),
                                                C.byref(it)))
    return (out.astype(bool), it.value) if return_iterations else out.astype(bool)


def remove_small_islands(mask, min_area0=100, min_area1=100, connectivity0=1, connectivity1=1):
    """transforms.py:170-206.  Same dtype out as in; ValueError for non-{0,1} input (:183-184)."""
    m = np.asarray(mask)
    if np.min(m) != 0 or np.max(m) > 1:
        raise ValueError("this function expects a binary mask of values 0 and 1")
    u = np.ascontiguousarray(m, dtype=np.uint8)
    out = np.empty_like(u)
    _ffi.check(_ffi.lib().tmat_remove_small_islands_host(_ffi.default_context().handle, _ffi.ptr(u), *u.shape,
                                                         int(min_area0), int(min_area1), int(connectivity0),
                                                         int(connectivity1), _ffi.ptr(out)))
    return out.astype(m.dtype)


def regionprops_image(mask, prop):
    """transforms.py:291-303.  ``area`` is computed on the GPU; other properties are outside the hot path."""
    if prop != "area":
        raise NotImplementedError(f"regionprops_image(prop={prop!r}): only 'area' is on the accelerated path")
    m = _mask_u8(mask)
    out = np.empty(m.shape, dtype=np.float64)
    _ffi.check(_ffi.lib().tmat_area_image_host(_ffi.default_context().handle, _ffi.ptr(m), *m.shape, _ffi.ptr(out)))
    return out


_DISK2 = object()


def filter_branch_seg_mask(mask, footprint=_DISK2, remove_isolated=True):
    """transforms.py:306-361.  Returns the cleaned boolean mask (the reference mutates and returns the
    median-filtered copy; with ``footprint=None`` it mutates the caller's array -- here the input is
    never modified)."""
    use_median = footprint is not None
    if use_median and footprint is not _DISK2 and not _is_disk(footprint, 2):
        raise NotImplementedError("libtmat_b200 implements the reference's default disk(2) footprint")
    m = _mask_u8(mask)
    out = np.empty_like(m)
    _ffi.check(_ffi.lib().tmat_filter_branch_seg_mask_host(_ffi.default_context().handle, _ffi.ptr(m), *m.shape,
                                                           int(use_median), int(bool(remove_isolated)), _ffi.ptr(out)))
    return out.astype(bool)
