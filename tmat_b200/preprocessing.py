"""Threshold helpers: signature-compatible with the reference's
``fl_tissue_model_tools/preprocessing.py`` (``apply_mask:13``, ``bin_thresh:26``).  These are single
element-wise passes that the reference runs on host arrays it already holds; the GPU versions used
inside the fused path are ``tmat_apply_mask_f32_dev`` / ``tmat_threshold_gt_f64_dev``."""
import numpy as np

from . import _ffi

MAX_UINT8 = 255


def apply_mask(img, mask):
    """preprocessing.py:13-23 on the GPU for float32 images (other dtypes: same result via a copy)."""
    a = np.asarray(img)
    m = np.asarray(mask)
    if a.dtype != np.float32 or a.ndim != 2:
        out = np.copy(a)
        out[m == 0] = 0
        return out
    import torch
    ctx = _ffi.default_context()
    d_img = torch.from_numpy(np.ascontiguousarray(a)).cuda()
    d_m = torch.from_numpy(np.ascontiguousarray(m != 0).astype(np.uint8)).cuda()
    d_out = torch.empty_like(d_img)
    s = _ffi.C.c_void_p(torch.cuda.current_stream().cuda_stream)
    _ffi.check(_ffi.lib().tmat_apply_mask_f32_dev(ctx.handle, _ffi.ptr(d_img), _ffi.ptr(d_m), a.size, _ffi.ptr(d_out), s))
    return d_out.cpu().numpy()


def bin_thresh(img, img_max, threshold=0):
    """preprocessing.py:26-41."""
    return np.where(np.copy(img) > threshold, img_max, 0)
