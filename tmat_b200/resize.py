"""The resizes either side of the segmentation path, on the B200 (``csrc/resize.cu``).

``resize`` has ``cv2.resize``'s signature -- including its positional ``dst`` slot, so that
``resize(img, dsize, cv2.INTER_AREA)`` (scripts/compute_cell_area.py:53-56) is, as with OpenCV, a bilinear resize --
and covers the two interpolations the reference's scripts use (INTER_LINEAR, INTER_LANCZOS4 of
scripts/compute_branches.py:309-312).  ``dsize`` is ``(width, height)`` as in OpenCV; the scripts pass
``(rows, cols)``-derived tuples into it, which is harmless only for square images and is
reproduced by keeping OpenCV's convention.  Parity with OpenCV is a tolerance (vendor code paths, see the header).

``pil_resize`` is ``Image.fromarray(arr).resize(size, resample)`` for float images (Pillow mode 'F'), bit-exact:
the pair of calls inside ``UNetXceptionPatchSegmentor.predict`` (models.py:630-651).
"""
import numpy as np

from . import _ffi

INTER_NEAREST, INTER_LINEAR, INTER_CUBIC, INTER_AREA, INTER_LANCZOS4 = 0, 1, 2, 3, 4
LANCZOS, NEAREST = 1, 0          # PIL.Image.Resampling values
_PIL_LANCZOS, _PIL_NEAREST, _CV_LINEAR, _CV_LANCZOS4 = 0, 1, 2, 3


def _run(img, w_out, h_out, mode):
    img = np.ascontiguousarray(img)
    if img.ndim != 2:
        raise ValueError(f"expected a 2-D image, got shape {img.shape}")
    if w_out < 1 or h_out < 1:
        raise ValueError(f"bad output size ({w_out}, {h_out})")
    out = np.empty((h_out, w_out), dtype=img.dtype)
    _ffi.check(_ffi.lib().tmat_resize_host(_ffi.default_context().handle, _ffi.ptr(img), _ffi.DTYPES[img.dtype],
                                           img.shape[0], img.shape[1], _ffi.ptr(out), h_out, w_out, mode))
    return out


def resize(src, dsize, dst=None, fx=None, fy=None, interpolation=INTER_LINEAR):
    """cv2.resize for 2-D uint8 / uint16 / float32 images, INTER_LINEAR and INTER_LANCZOS4."""
    src = np.asarray(src)
    if src.dtype == np.float64:
        raise TypeError("libtmat_b200 resizes uint8 / uint16 / float32 images")
    if dsize is None or tuple(dsize) == (0, 0):
        if not fx or not fy:
            raise ValueError("either dsize or both fx and fy must be given")
        dsize = (int(round(src.shape[1] * fx)), int(round(src.shape[0] * fy)))
    modes = {INTER_LINEAR: _CV_LINEAR, INTER_LANCZOS4: _CV_LANCZOS4}
    if interpolation not in modes:
        raise NotImplementedError(f"interpolation {interpolation}: INTER_LINEAR and INTER_LANCZOS4 are built")
    return _run(src, int(dsize[0]), int(dsize[1]), modes[interpolation])


def pil_resize(arr, size, resample=LANCZOS):
    """np.array(Image.fromarray(arr).resize(size, resample)) for float input (float64 becomes float32, as
    Image.fromarray does): ``size`` = (width, height); resample LANCZOS or NEAREST."""
    a = np.asarray(arr)
    if a.dtype.kind != "f":
        raise TypeError("pil_resize mirrors Pillow's mode-'F' path: float input only")
    a = a.astype(np.float32, copy=False)
    if resample not in (LANCZOS, NEAREST):
        raise NotImplementedError("resample must be LANCZOS (1) or NEAREST (0)")
    return _run(a, int(size[0]), int(size[1]), _PIL_LANCZOS if resample == LANCZOS else _PIL_NEAREST)
