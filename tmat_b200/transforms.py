"""Mask cleanup on the B200: drop-in for the inference-side functions of the reference's
``fl_tissue_model_tools/transforms.py`` (``remove_small_islands:170``, ``regionprops_image:291``,
``filter_branch_seg_mask:306``) plus the scikit-image primitives they are built from, all running in
``csrc/morph.cu``.  Integer work, bit-exact against ``oracle/morphology.py``; the scikit-image primitives that
oracle restates (notably the thinning table, ``csrc/skeleton_lut.inc``) are PARITY UNPINNED -- scikit-image is
installable neither in the build container nor on the GPU box."""
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
    if m.dtype == np.bool_ and m.flags["C_CONTIGUOUS"]:
        return m.view(np.uint8)                  # bool is stored as 0 / 1: no copy (the library never writes its inputs)
    return np.ascontiguousarray(m != 0, dtype=np.uint8)


def _is_disk(footprint, r):
    fp = np.asarray(footprint)
    return fp.shape == (2 * r + 1, 2 * r + 1) and np.array_equal(fp != 0, disk(r) != 0)


def _footprint_filter(m, footprint, op):
    fp = np.ascontiguousarray(np.asarray(footprint) != 0, dtype=np.uint8)
    if fp.ndim != 2 or fp.shape[0] % 2 == 0 or fp.shape[1] % 2 == 0:
        raise NotImplementedError(f"footprint must be 2-D with odd sides, got shape {fp.shape}")
    out = np.empty_like(m)
    _ffi.check(_ffi.lib().tmat_footprint_filter_host(_ffi.default_context().handle, _ffi.ptr(m), *m.shape,
                                                     _ffi.ptr(fp), fp.shape[0], fp.shape[1], op, 1, _ffi.ptr(out)))
    return out


def median_filter(mask, footprint=None):
    """skimage.filters.median(bool mask, footprint) (transforms.py:323); disk(2) -- the reference's default -- runs
    the specialised kernel, any other odd-sized footprint the generic rank kernel."""
    m = _mask_u8(mask)
    if footprint is not None and not _is_disk(footprint, 2):
        return _footprint_filter(m, footprint, 0).astype(bool)
    out = np.empty_like(m)
    _ffi.check(_ffi.lib().tmat_median_disk2_host(_ffi.default_context().handle, _ffi.ptr(m), *m.shape, _ffi.ptr(out)))
    return out.astype(bool)


def binary_erosion(mask, footprint):
    """skimage.morphology.binary_erosion(mask, footprint) (well_mask_generation.py:275; border_value=True)."""
    fp = np.asarray(footprint)
    m = _mask_u8(mask)
    r = (fp.shape[0] - 1) // 2
    if fp.ndim != 2 or not _is_disk(fp, r) or r > 8:
        return _footprint_filter(m, fp, 1).astype(bool)
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


_REGION_PROPS = {"area": 0, "equivalent_diameter_area": 1, "equivalent_diameter": 1, "eccentricity": 2}


def regionprops_image(mask, prop):
    """transforms.py:291-303 for the properties the reference's callers ask for (compute_branches.py:276-277):
    ``area``, ``equivalent_diameter_area`` and ``eccentricity`` (exact integer moments; a tolerance against
    scikit-image's float64 dot products).  Anything else raises NotImplementedError."""
    if prop not in _REGION_PROPS:
        raise NotImplementedError(f"regionprops_image(prop={prop!r}): built properties are {sorted(_REGION_PROPS)}")
    m = _mask_u8(mask)
    out = np.empty(m.shape, dtype=np.float64)
    _ffi.check(_ffi.lib().tmat_region_prop_image_host(_ffi.default_context().handle, _ffi.ptr(m), *m.shape,
                                                      _REGION_PROPS[prop], _ffi.ptr(out)))
    return out


_DISK2 = object()


def filter_branch_seg_mask(mask, footprint=_DISK2, remove_isolated=True):
    """transforms.py:306-361.  Returns the cleaned boolean mask.  Ownership as in the reference: with a
    footprint the input is left alone (the reference mutates the median-filtered copy, :323); with
    ``footprint=None`` the CALLER'S array is cleaned in place and returned (:356 writes into ``mask``)."""
    use_median = footprint is not None
    m = _mask_u8(mask)
    if use_median and footprint is not _DISK2 and not _is_disk(footprint, 2):
        m = _footprint_filter(m, footprint, 0)          # caller-supplied footprint: generic rank kernel first
        use_median = False
    out = np.empty_like(m)
    _ffi.check(_ffi.lib().tmat_filter_branch_seg_mask_host(_ffi.default_context().handle, _ffi.ptr(m), *m.shape,
                                                           int(use_median), int(bool(remove_isolated)), _ffi.ptr(out)))
    if footprint is None and isinstance(mask, np.ndarray) and mask.flags.writeable:
        mask[...] = out.astype(mask.dtype)
        return mask
    return out.view(np.bool_)


def set_skeletonize_lut(lut):
    """Replace the thinning table process-wide (256 entries in 0..3, layout of csrc/skeleton_lut.inc) -- for a
    maintainer who can read scikit-image's shipped table (scripts/extract_skimage_lut.py)."""
    t = np.ascontiguousarray(lut, dtype=np.uint8)
    if t.shape != (256,):
        raise ValueError("expected 256 entries")
    _ffi.check(_ffi.lib().tmat_skeleton_lut_set(_ffi.ptr(t)))


def distance_transform_edt(mask, invert=False):
    """scipy.ndimage.distance_transform_edt(mask) (compute_branches.py:342), exact: float64 distance of every
    non-zero pixel to the nearest zero pixel.  ``invert=True`` transforms ``logical_not(mask)``."""
    m = _mask_u8(mask)
    out = np.empty(m.shape, dtype=np.float64)
    _ffi.check(_ffi.lib().tmat_distance_transform_edt_host(_ffi.default_context().handle, _ffi.ptr(m), *m.shape,
                                                           int(bool(invert)), _ffi.ptr(out)))
    return out


def centerline_weighting(pred, seg_mask, skel):
    """compute_branches.py:341-344 given the medial-axis skeleton: ``pred * dist / (dist + centerline_dt)`` with
    ``dist = distance_transform_edt(seg_mask)`` and ``centerline_dt = distance_transform_edt(~skel)``."""
    p = np.ascontiguousarray(pred, dtype=np.float64)
    m, s = _mask_u8(seg_mask), _mask_u8(skel)
    if p.shape != m.shape or s.shape != m.shape:
        raise ValueError("pred, seg_mask and skel must have the same shape")
    out = np.empty(m.shape, dtype=np.float64)
    _ffi.check(_ffi.lib().tmat_centerline_weight_host(_ffi.default_context().handle, _ffi.ptr(p), _ffi.ptr(m), _ffi.ptr(s),
                                                      *m.shape, _ffi.ptr(out)))
    return out
