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
    if stack.dtype == np.float64 and method != "fs":
        # numpy would stay float64; the kernels compute float stacks in float32
        raise TypeError("libtmat_b200 projects uint8 / uint16 / float32 stacks; cast a float64 stack to float32 first")
    if stack.dtype not in (np.uint8, np.uint16) and not (stack.dtype == np.float32 and method != "fs"):
        raise TypeError(f"libtmat_b200 projects uint8 / uint16 stacks (and float32 for min/max/avg/med), got "
                        f"{stack.dtype} for method {method!r}")
    stack = np.ascontiguousarray(stack)
    z, h, w = stack.shape
    # numpy's result dtypes: mean/median of an integer stack are float64, of a float32 stack float32
    out_dtype = np.float64 if (method in ("avg", "med") and stack.dtype != np.float32) else stack.dtype
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
