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


def blur(img, blur_itr, k_size=3, gs=True):
    """preprocessing.py:165-183 on the GPU: ``blur_itr`` x ``cv2.GaussianBlur((k, k), 0)`` then ``.round()``
    (``.astype(uint8)`` when ``gs``).  uint8 input is bit-exact (fixed-point kernels); float32 input is
    bit-exact for ``k_size=3`` and within 2 ulp of the image range for 5 and 7."""
    a = np.ascontiguousarray(img)
    if a.ndim != 2 or a.dtype not in (np.dtype(np.uint8), np.dtype(np.float32)):
        raise TypeError("blur: a 2-D uint8 or float32 image is required on the GPU path")
    if k_size not in (3, 5, 7):
        raise ValueError("blur: k_size must be 3, 5 or 7 (cv2's fixed sigma=0 kernels)")
    is_f = a.dtype == np.float32
    final_op = (2 if gs else 1) if is_f else 0
    out = np.empty(a.shape, np.uint8 if (gs or not is_f) else np.float32)
    _ffi.check(_ffi.lib().tmat_blur_host(_ffi.default_context().handle, _ffi.ptr(a), _ffi.DTYPES[a.dtype], a.shape[0],
                                         a.shape[1], int(k_size), int(blur_itr), final_op, _ffi.ptr(out)))
    return out


DIST_L2 = 2          # cv2.DIST_L2


def gen_circ_mask(center, radius, shape, mask_val=1):
    """preprocessing.py:96-116: ``cv2.circle(zeros(shape, uint8), center, radius, mask_val, -1)``; ``center`` is
    (column, row).  OpenCV's integer midpoint circle, bit-exact."""
    h, w = int(shape[0]), int(shape[1])
    out = np.empty((h, w), np.uint8)
    _ffi.check(_ffi.lib().tmat_circle_mask_host(_ffi.default_context().handle, h, w, int(center[0]), int(center[1]),
                                                int(radius), int(mask_val), _ffi.ptr(out)))
    return out


def _chamfer5(mask_u8, invert=False):
    m = np.ascontiguousarray(mask_u8, dtype=np.uint8)
    out = np.empty(m.shape, np.float32)
    _ffi.check(_ffi.lib().tmat_distance_transform_cv5_host(_ffi.default_context().handle, _ffi.ptr(m), m.shape[0],
                                                           m.shape[1], int(invert), _ffi.ptr(out)))
    return out


def dt_blur(img, blur_itr, dist_metric=DIST_L2, k_size=3):
    """preprocessing.py:118-135: ``cv2.distanceTransform(bin_thresh(img, 255).round().astype(uint8), DIST_L2, 5)`` then
    ``blur`` (uint8 out).  The chamfer transform matches OpenCV to ~1e-6 relative (see csrc/edt.cu)."""
    if dist_metric != DIST_L2:
        raise NotImplementedError("dt_blur: cv2.DIST_L2 (the reference's default) is built")
    mask = (np.asarray(img) > 0).astype(np.uint8)
    return blur(_chamfer5(mask), blur_itr, k_size)


def sdt_blur(img, blur_itr, dist_metric=DIST_L2, k_size=3):
    """preprocessing.py:138-163: signed distance (inside minus outside) then ``blur(..., gs=False)`` (float32 out)."""
    if dist_metric != DIST_L2:
        raise NotImplementedError("sdt_blur: cv2.DIST_L2 (the reference's default) is built")
    mask = (np.asarray(img) > 0).astype(np.uint8)
    return blur(_chamfer5(mask) - _chamfer5(mask, invert=True), blur_itr, k_size, gs=False)


def gmm_foreground_threshold(pixels, sd_coef, rand_state, return_params=False):
    """The scalar of preprocessing.py:75-87: 2-component GMM on ``pixels`` (1-D), threshold =
    ``min(255, mean_fg + sd_coef * sd_fg)``.  The three uniform draws scikit-learn's k-means++ seeding would
    take from ``rand_state`` are taken here in the same order; everything that touches the samples runs on
    the GPU (csrc/gmm.cu)."""
    x = np.ascontiguousarray(np.asarray(pixels).reshape(-1), dtype=np.float32)
    rnd = np.empty(3, np.float64)
    rnd[0] = rand_state.random_sample()          # random_state.choice(n_samples, p=uniform)
    rnd[1:] = rand_state.uniform(size=2)         # the two local seeding trials
    res = np.empty(10, np.float64)
    _ffi.check(_ffi.lib().tmat_gmm2_threshold_host(_ffi.default_context().handle, _ffi.ptr(x), x.size, float(sd_coef),
                                                   _ffi.ptr(rnd), _ffi.ptr(res)))
    if return_params:
        return res[0], {"weights": res[1:3].copy(), "means": res[3:5].copy(), "variances": res[5:7].copy(),
                        "n_iter": int(res[7]), "kmeans_iter": int(res[8]), "lower_bound": res[9]}
    return res[0]


def exec_threshold(masked, mask_idx, sd_coef, rand_state):
    """preprocessing.py:44-93.  ``masked``: 2-D image (the reference passes float32 in [0, 1]); ``mask_idx``:
    index tuple of the pixels inside the mask, or None for all pixels.  Returns a copy with background pixels
    (``<=`` threshold) set to 0."""
    a = np.asarray(masked)
    pixels = a[mask_idx] if mask_idx is not None else a.reshape(-1)
    thr = gmm_foreground_threshold(pixels, sd_coef, rand_state)
    if a.dtype != np.float32:
        return np.where(np.copy(a) <= thr, 0, a)
    src = np.ascontiguousarray(a)
    out = np.empty_like(src)
    _ffi.check(_ffi.lib().tmat_threshold_le_zero_f32_host(_ffi.default_context().handle, _ffi.ptr(src), src.size,
                                                          float(thr), _ffi.ptr(out)))
    return out
