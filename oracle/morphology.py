"""Oracle (test infrastructure): binary morphology / component cleanup.

Restates ``transforms.py:170-206`` (``remove_small_islands``), ``:209-288``
(``nx_graph_from_binary_skeleton``), ``:291-303`` (``regionprops_image``) and
``:306-361`` (``filter_branch_seg_mask``) of the reference, with the scikit-image
0.22 primitives it calls restated on scipy.ndimage / numpy:

  skimage.filters.median(bool, footprint)    -> ndi.median_filter(mode='nearest')
  skimage.measure.label(connectivity=c)      -> ndi.label(structure of rank c)  (raster-ordered)
  regionprops.area / .perimeter              -> bincount / 4-neighbourhood border-code histogram
  skimage.morphology.remove_small_objects    -> bincount filter (labels with size < min_size)
  skimage.morphology.skeletonize (2-D)       -> Zhang-Suen two-sub-pass parallel thinning [Zha84]
  skimage.morphology.binary_erosion          -> ndi.binary_erosion(border_value=True)
  skimage.morphology.disk(r)                 -> x^2 + y^2 <= r^2

scikit-image is not installable in the build container and is absent from the GPU box's
image (probed in round 2: ``import skimage`` fails there too), so these primitives are
**parity unpinned**; the *composition* in ``filter_branch_seg_mask`` is pinned by
running the reference's own ``transforms.py`` over these primitives
(tests/golden/make_golden.py).  The thinning table in particular is DERIVED from the
published Zhang-Suen conditions, not transcribed from scikit-image's hard-coded table
(``_skeletonize_cy.pyx``); it lives in one data file shared with the CUDA path
(``tmat_b200/csrc/skeleton_lut.inc``) so that a corrected table changes kernel, oracle
and fixtures together.
"""
import numpy as np
from scipy import ndimage as ndi


def disk(r):
    y, x = np.mgrid[-r:r + 1, -r:r + 1]
    return (x * x + y * y <= r * r).astype(np.uint8)


def median_bool(mask, footprint):
    """skimage.filters.median on a boolean image (transforms.py:323)."""
    m = ndi.median_filter(np.asarray(mask).astype(np.uint8), footprint=footprint, mode="nearest")
    return m.astype(bool)


def label(mask, connectivity=2):
    """skimage.measure.label for 2-D input (transforms.py:188,196,301,325)."""
    st = ndi.generate_binary_structure(2, connectivity)
    lab, n = ndi.label(np.asarray(mask) != 0, structure=st)
    return lab.astype(np.int32), n


def binary_erosion(mask, footprint):
    """skimage.morphology.binary_erosion (well_mask_generation.py:275)."""
    return ndi.binary_erosion(np.asarray(mask).astype(bool), structure=footprint, border_value=True)


PERIM_WEIGHTS = np.zeros(50, dtype=np.float64)
PERIM_WEIGHTS[[5, 7, 15, 17, 25, 27]] = 1
PERIM_WEIGHTS[[21, 33]] = np.sqrt(2)
PERIM_WEIGHTS[[13, 23]] = (1 + np.sqrt(2)) / 2


def perimeter_code_hist(labels, n):
    """Per-label histogram (n, 50) of skimage.measure.perimeter(neighborhood=4) codes.
    For 8-connected labels two touching foreground pixels always share a label, so the
    per-region computation on the cropped region image equals this global one."""
    fg = labels > 0
    p = np.pad(fg, 1)
    interior = p[:-2, 1:-1] & p[2:, 1:-1] & p[1:-1, :-2] & p[1:-1, 2:] & fg
    border = fg & ~interior
    k = np.array([[10, 2, 10], [2, 1, 2], [10, 2, 10]])
    code = ndi.convolve(border.astype(np.int64), k, mode="constant", cval=0)
    hist = np.zeros((n, 50), dtype=np.int64)
    sel = border
    np.add.at(hist, (labels[sel] - 1, code[sel]), 1)
    return hist


def region_areas(labels, n):
    return np.bincount(labels.ravel(), minlength=n + 1)[1:]


def circularities(labels, n):
    """transforms.py:326-329: 4*pi*area / (perimeter^2 + 1e-7)."""
    area = region_areas(labels, n)
    perim = perimeter_code_hist(labels, n) @ PERIM_WEIGHTS
    return 4 * np.pi * area / (perim ** 2 + 1e-7)


def _zs_lut():
    """Zhang-Suen deletability: bit0 NW, 1 N, 2 NE, 3 E, 4 SE, 5 S, 6 SW, 7 W.
    value 1: deletable in sub-pass 1 only, 2: sub-pass 2 only, 3: both."""
    lut = np.zeros(256, dtype=np.uint8)
    for v in range(256):
        nw, n_, ne, e, se, s, sw, w_ = [(v >> b) & 1 for b in range(8)]
        ring = [n_, ne, e, se, s, sw, w_, nw]            # P2..P9
        b = sum(ring)
        a = sum(1 for i in range(8) if ring[i] == 0 and ring[(i + 1) % 8] == 1)
        if not (2 <= b <= 6 and a == 1):
            continue
        first = (n_ * e * s == 0) and (e * s * w_ == 0)
        second = (n_ * e * w_ == 0) and (n_ * s * w_ == 0)
        lut[v] = (1 if first else 0) | (2 if second else 0)
    return lut


def _shared_lut():
    """The ONE table constant shared with the CUDA path (tmat_b200/csrc/skeleton_lut.inc): the oracle reads the
    data file, it does not import product code.  tests/test_oracle_golden.py checks it equals _zs_lut()."""
    import os
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tmat_b200", "csrc",
                        "skeleton_lut.inc")
    body = open(path).read().split("*/", 1)[1]
    vals = [int(t) for t in body.replace(",", " ").split()]
    assert len(vals) == 256 and all(0 <= v <= 3 for v in vals), len(vals)
    return np.array(vals, dtype=np.uint8)


ZS_LUT = _shared_lut()


def neighbour_code(img):
    p = np.pad(img.astype(np.int32), 1)
    return (p[:-2, :-2] + 2 * p[:-2, 1:-1] + 4 * p[:-2, 2:] + 8 * p[1:-1, 2:]
            + 16 * p[2:, 2:] + 32 * p[2:, 1:-1] + 64 * p[2:, :-2] + 128 * p[1:-1, :-2])


def skeletonize(mask):
    """2-D skimage.morphology.skeletonize (transforms.py:331): repeat {sub-pass 1,
    sub-pass 2} -- each a fully parallel deletion on a zero-padded image -- until an
    iteration deletes nothing."""
    sk = np.asarray(mask).astype(bool).copy()
    iters = 0
    while True:
        removed = False
        for bit in (1, 2):
            kill = sk & ((ZS_LUT[neighbour_code(sk)] & bit) != 0)
            if kill.any():
                sk &= ~kill
                removed = True
        iters += 1
        if not removed:
            break
    skeletonize.last_iterations = iters
    return sk


def skeleton_degree(sk):
    """Degree of each skeleton pixel in the 8-neighbour pixel graph
    (transforms.py:248 shifts (1,0),(0,1),(1,1),(1,-1))."""
    p = np.pad(sk.astype(np.int32), 1)
    deg = sum(p[1 + dy:p.shape[0] - 1 + dy, 1 + dx:p.shape[1] - 1 + dx]
              for dy in (-1, 0, 1) for dx in (-1, 0, 1) if (dy, dx) != (0, 0))
    return deg * sk


def filter_branch_seg_mask(mask, footprint="disk2", remove_isolated=True):
    """transforms.py:306-361.  Returns a new bool mask (the reference mutates the
    median-filtered copy, or the caller's array when footprint is None)."""
    if isinstance(footprint, str):
        footprint = disk(2)
    if footprint is not None:
        mask = median_bool(mask, footprint)
    else:
        mask = np.array(mask).astype(bool)
    labels, n = label(mask, 2)
    circ = circularities(labels, n)
    sk = skeletonize(mask)
    fork = skeleton_degree(sk) > 2
    sk_lab, n_sk = label(sk, 2)
    has_fork = np.zeros(n_sk + 1, dtype=bool)
    has_fork[np.unique(sk_lab[fork])] = True
    first = ndi.minimum_position(sk_lab, sk_lab, index=np.arange(1, n_sk + 1)) if n_sk else []
    kill = np.zeros(n + 1, dtype=bool)
    for s, pos in enumerate(first, start=1):
        lab = labels[pos]
        if (remove_isolated and not has_fork[s]) or circ[lab - 1] > 0.8:
            kill[lab] = True
    out = mask.copy()
    out[kill[labels]] = False
    return out


def remove_small_objects(labels, min_size):
    sizes = np.bincount(labels.ravel())
    small = sizes < min_size
    small[0] = False
    out = labels.copy()
    out[small[labels]] = 0
    return out


def remove_small_islands(mask, min_area0=100, min_area1=100, connectivity0=1, connectivity1=1):
    """transforms.py:170-206."""
    if np.min(mask) != 0 or np.max(mask) > 1:
        raise ValueError("this function expects a binary mask of values 0 and 1")
    mask = mask.copy()
    inv_lab, _ = label(1 - mask, connectivity0)
    inv_lab = remove_small_objects(inv_lab, min_area0)
    mask[inv_lab == 0] = 1
    lab, _ = label(mask, connectivity1)
    lab = remove_small_objects(lab, min_area1)
    mask[lab == 0] = 0
    return mask


def regionprops_image_area(mask):
    """transforms.py:291-303 for prop='area' (float64 image, background 0)."""
    lab, n = label(mask, 2)
    props = np.insert(region_areas(lab, n).astype(np.float64), 0, 0)
    return props[lab]


def regionprops_image_eccentricity(mask):
    """transforms.py:291-303 with prop='eccentricity' (compute_branches.py:276), restating scikit-image 0.22's
    RegionProperties: moments_central of the bounding-box crop about its centroid (float64 dot products),
    inertia tensor [[mu02, -mu11], [-mu11, mu20]] / mu00, eigvalsh clipped at 0 and sorted descending,
    eccentricity = 0 if l1 == 0 else sqrt(1 - l2 / l1).  PARITY UNPINNED: scikit-image is not installable here."""
    lab, n = label(np.asarray(mask) != 0, 2)
    out = np.zeros(lab.shape, dtype=np.float64)
    for k, sl in enumerate(ndi.find_objects(lab), start=1):
        img = (lab[sl] == k).astype(np.float64)
        r = np.arange(img.shape[0], dtype=np.float64)
        c = np.arange(img.shape[1], dtype=np.float64)
        m00 = img.sum()
        cr, cc = (img.sum(axis=1) @ r) / m00, (img.sum(axis=0) @ c) / m00
        dr, dc = r - cr, c - cc
        mu20 = (dr ** 2) @ img @ np.ones_like(c)
        mu02 = np.ones_like(r) @ img @ (dc ** 2)
        mu11 = dr @ img @ dc
        T = np.array([[mu02, -mu11], [-mu11, mu20]]) / m00
        ev = np.clip(np.linalg.eigvalsh(T), 0, None)
        l1, l2 = max(ev), min(ev)
        out[lab == k] = 0.0 if l1 == 0 else np.sqrt(1 - l2 / l1)
    return out
