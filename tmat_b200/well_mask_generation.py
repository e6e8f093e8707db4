"""Threshold + erosion part of well-mask generation on the B200: drop-in for
``auto_threshold_well`` of the reference's ``fl_tissue_model_tools/well_mask_generation.py:236-277``.
(The superellipse fit of ``generate_well_mask`` works on a <=200 px thumbnail with an unseeded RNG
and stays host-side / out of scope.)"""
import ctypes as C

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
