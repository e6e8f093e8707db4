"""Z-stack projections on the B200: drop-in for the projection functions of the reference's
``fl_tissue_model_tools/zstacks.py`` (``proj_focus_stacking:153``, ``proj_avg:192``,
``proj_med:206``, ``proj_max:221``, ``proj_min:237``).  Same names, arguments and result
dtypes; the work happens in ``csrc/zproj.cu`` behind ``tmat_zproj_host``.
"""
import numpy as np

from . import _ffi


def _project(stack, axis, method, kernel_size=5):
    stack = np.asarray(stack)
    if axis != 0:
        stack = np.moveaxis(stack, axis, 0)
    if stack.ndim != 3:
        raise ValueError(f"expected a 3-D (Z, H, W) stack, got shape {stack.shape}")
    if stack.dtype not in (np.uint8, np.uint16):
        raise TypeError(f"libtmat_b200 projects uint8 / uint16 stacks, got {stack.dtype}")
    stack = np.ascontiguousarray(stack)
    z, h, w = stack.shape
    out_dtype = np.float64 if method in ("avg", "med") else stack.dtype
    out = np.empty((h, w), dtype=out_dtype)
    ctx = _ffi.default_context()
    _ffi.check(_ffi.lib().tmat_zproj_host(ctx.handle, _ffi.ptr(stack), _ffi.DTYPES[stack.dtype],
                                          z, h, w, _ffi.ZPROJ[method], kernel_size, _ffi.ptr(out)))
    return out


def proj_focus_stacking(stack, axis=0, kernel_size=5):
    """zstacks.py:153-189."""
    return _project(stack, axis, "fs", kernel_size)


def proj_avg(stack, axis=0):
    """zstacks.py:192-203."""
    return _project(stack, axis, "avg")


def proj_med(stack, axis=0):
    """zstacks.py:206-218."""
    return _project(stack, axis, "med")


def proj_max(stack, axis=0):
    """zstacks.py:221-234."""
    return _project(stack, axis, "max")


def proj_min(stack, axis=0):
    """zstacks.py:237-249."""
    return _project(stack, axis, "min")
