"""Oracle (test infrastructure): Z-stack projections.

Restates ``zstacks.py:134-249`` of the reference: ``proj_max:221``, ``proj_min:237``,
``proj_avg:192``, ``proj_med:206``, ``proj_focus_stacking:153`` with
``_blur_and_lap:134-150``.  The focus-stacking arithmetic is written out in exact
integers (what OpenCV's 5x5 GaussianBlur(sigma=0) + Laplacian(CV_64F, ksize=5)
compute on uint8/uint16 input), so the oracle does not depend on cv2.
PINNED against the imported reference (which does call cv2) by
tests/golden/make_golden.py.
"""
import numpy as np


def proj_max(stack, axis=0):
    return np.max(stack, axis=axis)


def proj_min(stack, axis=0):
    return np.min(stack, axis=axis)


def proj_avg(stack, axis=0):
    return np.mean(stack, axis=axis)


def proj_med(stack, axis=0):
    return np.median(stack, axis=axis)


def _reflect101(idx, n):
    """cv2.BORDER_REFLECT_101 index map."""
    idx = np.abs(idx)
    idx = np.where(idx >= n, 2 * (n - 1) - idx, idx)
    return idx


def _sep5(img, ky, kx):
    """Exact int64 separable 5-tap correlation with BORDER_REFLECT_101."""
    h, w = img.shape
    rows = _reflect101(np.arange(-2, h + 2), h)
    cols = _reflect101(np.arange(-2, w + 2), w)
    p = img[rows][:, cols].astype(np.int64)
    t = sum(int(kx[d]) * p[:, d:d + w] for d in range(5))
    return sum(int(ky[d]) * t[d:d + h, :] for d in range(5))


_S = (1, 4, 6, 4, 1)      # cv2.getGaussianKernel(5, 0) * 16
_D2 = (1, 0, -2, 0, 1)    # cv2.getDerivKernels(2, 0, 5)


def blur5_int(img):
    """cv2.GaussianBlur(img, (5,5), 0) for uint8/uint16: one rounding (S+128)>>8."""
    return ((_sep5(img, _S, _S) + 128) >> 8).astype(img.dtype)


def abs_lap5_int(blurred):
    """|cv2.Laplacian(blurred, CV_64F, ksize=5)| as exact integers."""
    return np.abs(_sep5(blurred, _S, _D2) + _sep5(blurred, _D2, _S))


def proj_focus_stacking(stack, axis=0, kernel_size=5):
    """zstacks.py:153-189.  Strict '>' against a float32 running maximum that starts
    at -inf: the earliest slice wins ties; output takes the unblurred pixel."""
    if kernel_size != 5:
        raise NotImplementedError("oracle restates kernel_size=5 only")
    if axis != 0:
        stack = np.moveaxis(stack, axis, 0)
    if stack.dtype not in (np.uint8, np.uint16):
        raise NotImplementedError("integer restatement covers uint8/uint16")
    best = np.full(stack[0].shape, -np.inf, dtype=np.float32)
    out = stack[0].copy()
    for sl in stack:
        sharp = abs_lap5_int(blur5_int(sl)).astype(np.float64)
        take = sharp > best
        best[take] = sharp[take]
        out[take] = sl[take]
    return out


METHODS = {"min": proj_min, "max": proj_max, "avg": proj_avg, "med": proj_med,
           "fs": proj_focus_stacking}
