"""Oracle (test infrastructure): threshold helpers and the iterated blur.

Restates ``preprocessing.py:13-23`` (``apply_mask``), ``:26-41`` (``bin_thresh``),
``:44-93`` (``exec_threshold``: 2-component GMM, threshold = mean_fg + sd_coef*sd_fg
capped at 255, ``<=`` comparison), ``:165-183`` (``blur``) and the call-site wrapper
``scripts/compute_cell_area.py:60-87`` (``mask_and_threshold``).
PINNED against the imported reference module (tests/golden/make_golden.py).
"""
import numpy as np

MAX_UINT8 = 255  # defs.py:11


def apply_mask(img, mask):
    out = np.copy(img)
    out[mask == 0] = 0
    return out


def bin_thresh(img, img_max, threshold=0):
    return np.where(np.copy(img) > threshold, img_max, 0)


def gmm_foreground_threshold(pixels, sd_coef, rand_state):
    """preprocessing.py:75-87 -- the scalar threshold only."""
    from sklearn.mixture import GaussianMixture
    gm = GaussianMixture(n_components=2, random_state=rand_state).fit(pixels[:, None])
    means = gm.means_.squeeze()
    sds = np.sqrt(gm.covariances_.squeeze())
    k = int(np.argmax(means))
    return min(MAX_UINT8, means[k] + sds[k] * sd_coef)


def exec_threshold(masked, mask_idx, sd_coef, rand_state):
    if mask_idx is None:
        mask_idx = tuple(np.indices(masked.shape).reshape(2, -1))
    thr = gmm_foreground_threshold(masked[mask_idx], sd_coef, rand_state)
    return np.where(np.copy(masked) <= thr, 0, masked)


def rescale_unit_f32(img):
    """skimage rescale_intensity(img, out_range=(0, 1)).astype(np.float32)
    (compute_cell_area.py:80, compute_branches.py:316): float64 arithmetic for integer
    input, the input's own float type otherwise."""
    img = np.asarray(img)
    lo, hi = img.min(), img.max()
    ft = img.dtype if img.dtype.kind == "f" and img.dtype.itemsize >= 4 else np.float64
    y = np.clip(img, lo, hi)
    if lo != hi:
        y = (y - lo) / (hi - lo)
        return np.asarray(y * 1 + 0, dtype=ft).astype(np.float32)
    return np.clip(y, 0, 1).astype(ft).astype(np.float32)     # constant image: skimage clips to the output range


def mask_and_threshold(img, sd_coef, rand_state, well_mask=None, well_idx=None):
    img = rescale_unit_f32(img)
    masked = apply_mask(img, well_mask) if well_mask is not None else img
    out = exec_threshold(masked, well_idx, sd_coef, rand_state)
    return (out > 0).astype(np.uint8) * MAX_UINT8


def blur(img, blur_itr, k_size=3, gs=True):
    """preprocessing.py:165-183; cv2.GaussianBlur is the arithmetic being restated, and
    cv2 ships in the image on both boxes."""
    import cv2
    out = img.copy()
    for _ in range(blur_itr):
        out = cv2.GaussianBlur(out, (k_size, k_size), 0)
    return out.round() if not gs else out.round().astype(np.uint8)
