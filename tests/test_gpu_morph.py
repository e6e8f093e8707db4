"""GPU: thresholds, Gaussian pre-filter, auto_threshold_well and the mask-cleanup stage through the
C ABI vs the oracle and the golden fixtures.  Integer / mask results are bit-exact."""
import numpy as np
import pytest

import helpers
from oracle import morphology as om, wellmask as ow

pytestmark = pytest.mark.gpu


def _unpack(g, name):
    shape = tuple(g[f"{name}_shape"])
    return lambda a: np.unpackbits(a)[: shape[0] * shape[1]].reshape(shape).astype(bool)


def test_filter_branch_seg_mask_golden(golden):
    from tmat_b200 import transforms as T
    g = golden("morphology")
    for name in ("real", "synth"):
        un = _unpack(g, name)
        m = un(g[f"{name}_in"])
        assert np.array_equal(T.filter_branch_seg_mask(m), un(g[f"{name}_filter_composition"])), name
        # footprint=None cleans the array it is given in place (transforms.py:356): hand it a copy
        assert np.array_equal(T.filter_branch_seg_mask(m.copy(), None, False),
                              un(g[f"{name}_filter_nomedian_keepisolated_composition"])), name
        assert np.array_equal(T.remove_small_islands(m.astype(np.uint8), 30, 40, 1, 2).astype(bool),
                              un(g[f"{name}_islands_composition"])), name
        assert np.array_equal(T.regionprops_image(m, "area").astype(np.float32), g[f"{name}_area_image_composition"])


@pytest.mark.parametrize("h,w,seed", [(64, 64, 1), (97, 131, 2), (300, 340, 3), (33, 500, 4), (1, 40, 5), (40, 1, 6),
                                      (128, 264, 7), (200, 1032, 8)])
def test_morphology_primitives_vs_oracle(h, w, seed):
    from tmat_b200 import transforms as T
    m = helpers.synth_mask(h, w, seed, density=0.15)
    assert np.array_equal(T.median_filter(m, T.disk(2)), om.median_bool(m, om.disk(2)))
    for r in (1, 3, 5):
        assert np.array_equal(T.binary_erosion(m, T.disk(r)), om.binary_erosion(m, om.disk(r))), r
    for conn in (1, 2):
        lab, n = T.measure_label(m, conn, return_num=True)
        ref, nr = om.label(m, conn)
        assert n == nr and np.array_equal(lab, ref), conn
    sk, it = T.skeletonize(m, return_iterations=True)
    assert np.array_equal(sk, om.skeletonize(m)) and it == om.skeletonize.last_iterations
    for rem in (True, False):
        for fp in ("disk2", None):
            got = T.filter_branch_seg_mask(m.copy(), T.disk(2) if fp else None, rem)
            assert np.array_equal(got, om.filter_branch_seg_mask(m, fp, rem)), (rem, fp)
    u = m.astype(np.uint8)
    if u.min() == 0 and u.max() == 1:
        assert np.array_equal(T.remove_small_islands(u, 20, 25, 2, 1), om.remove_small_islands(u, 20, 25, 2, 1))
    assert np.array_equal(T.regionprops_image(m, "area"), om.regionprops_image_area(m))


def test_region_stats_vs_oracle():
    import ctypes as C
    import torch
    from tmat_b200 import _ffi
    m = helpers.synth_mask(200, 260, 8, density=0.2)
    lab, n = om.label(m, 2)
    ctx = _ffi.default_context()
    d_lab = torch.from_numpy(lab).cuda()
    area = np.zeros(n, np.int64)
    hist = np.zeros((n, 50), np.int64)
    _ffi.check(_ffi.lib().tmat_region_stats_dev(ctx.handle, _ffi.ptr(d_lab), 200, 260, n, _ffi.ptr(area), _ffi.ptr(hist), None))
    assert np.array_equal(area, om.region_areas(lab, n))
    assert np.array_equal(hist, om.perimeter_code_hist(lab, n))


def test_edge_cases_empty_full_and_errors():
    from tmat_b200 import transforms as T
    z = np.zeros((50, 60), bool)
    o = np.ones((50, 60), bool)
    assert not T.filter_branch_seg_mask(z).any()
    assert np.array_equal(T.filter_branch_seg_mask(o), om.filter_branch_seg_mask(o))
    lab, n = T.measure_label(z, 2, return_num=True)
    assert n == 0 and not lab.any()
    assert T.measure_label(o, 1, return_num=True)[1] == 1
    assert not T.skeletonize(z).any()
    with pytest.raises(ValueError):
        T.remove_small_islands(np.full((8, 8), 2, np.uint8))
    with pytest.raises(ValueError):
        T.remove_small_islands(np.ones((8, 8), np.uint8))     # the reference rejects an all-ones mask too


@pytest.mark.parametrize("h,w", [(50, 8), (70, 264), (37, 256), (130, 520), (1, 16), (9, 1032), (64, 72)])
def test_word_lane_erosion_and_median_vs_oracle(h, w):
    """Widths on the 8-pixel grid take the 8-pixels-per-lane kernels (a warp spans 256 columns): one word, one span plus
    one word, exact spans, several spans, a single row -- random masks dense enough for the erosion to keep something."""
    from tmat_b200 import transforms as T
    rng = np.random.default_rng(h * 1009 + w)
    for dens in (0.5, 0.9, 0.985):
        m = rng.random((h, w)) < dens
        assert np.array_equal(T.median_filter(m, T.disk(2)), om.median_bool(m, om.disk(2))), dens
        assert np.array_equal(T.binary_erosion(m, T.disk(5)), om.binary_erosion(m, om.disk(5))), dens
    m = np.ones((h, w), bool)
    assert T.binary_erosion(m, T.disk(5)).all() and T.median_filter(m, T.disk(2)).all()


@pytest.mark.parametrize("w", [264, 100])
def test_erosion_border_value_zero(w):
    """tmat_erode_disk with border_value = 0 (outside the image counts as background) on both kernel families."""
    from scipy import ndimage as ndi
    from tmat_b200 import _ffi
    rng = np.random.default_rng(w)
    m = (rng.random((90, w)) < 0.97).astype(np.uint8)
    out = np.empty_like(m)
    for r in (5, 2):
        _ffi.check(_ffi.lib().tmat_erode_disk_host(_ffi.default_context().handle, _ffi.ptr(m), 90, w, r, 0, _ffi.ptr(out)))
        assert np.array_equal(out != 0, ndi.binary_erosion(m != 0, structure=om.disk(r), border_value=0)), r


@pytest.mark.parametrize("shape", [(256, 512), (120, 264), (64, 1032)])
def test_auto_threshold_well_word_path_shapes(shape):
    """auto_threshold_well on widths that are multiples of 8: the fused Gaussian passes feed the word-lane erosion with
    the Otsu threshold folded into its load."""
    from tmat_b200 import well_mask_generation as W
    for dtype, invert in ((np.uint16, False), (np.uint8, True)):
        img = helpers.synth_well(*shape, 23, dtype, invert)
        mask, info = W.auto_threshold_well(img, return_info=True)
        ref, st = ow.auto_threshold_well(img, return_stages=True)
        assert info["invert"] == st["invert"] and info["otsu"] == st["otsu"]
        assert np.array_equal(mask, ref)


@pytest.mark.parametrize("dtype,invert", [(np.uint16, False), (np.uint16, True), (np.uint8, False), (np.float32, True),
                                          (np.float32, False)])
def test_auto_threshold_well_vs_oracle(dtype, invert):
    from tmat_b200 import well_mask_generation as W
    shape = (517, 1031) if (dtype == np.uint16 and invert) else (240, 310)
    img = helpers.synth_well(*shape, 11, dtype, invert)
    blur = W.gaussian(img, 1.0)
    ref_blur = ow.gaussian(img, 1.0)
    assert blur.dtype == ref_blur.dtype and np.array_equal(blur, ref_blur)
    mask, info = W.auto_threshold_well(img, return_info=True)
    ref, st = ow.auto_threshold_well(img, return_stages=True)
    assert info["invert"] == st["invert"] and info["otsu"] == st["otsu"]
    assert np.array_equal(mask, ref)


@pytest.mark.parametrize("shape", [(243, 309), (7, 5), (1, 600), (600, 1), (3, 1030), (1025, 3), (64, 512)])
def test_gaussian_ragged_shapes_bit_exact(shape):
    """The fixed-radius Gaussian kernels (8-row / 4 x 4-pixel thread tiles) on shapes off every tile grid, smaller than
    the filter radius, and single rows / columns -- bit-exact against scipy's gaussian_filter for uint16 and uint8."""
    from tmat_b200 import well_mask_generation as W
    rng = np.random.default_rng(shape[0] * 7919 + shape[1])
    for dtype in (np.uint16, np.uint8):
        img = rng.integers(0, np.iinfo(dtype).max + 1, shape).astype(dtype)
        got, ref = W.gaussian(img, 1.0), ow.gaussian(img, 1.0)
        assert got.dtype == ref.dtype == np.float64 and np.array_equal(got, ref), (shape, dtype)
    # another sigma goes through the generic kernels.  Not on the reference's path (auto_threshold_well uses sigma = 1
    # only): the library builds the weights with libm's exp, scipy with numpy's vectorised exp, which differ by an ulp
    # for some arguments -- a tolerance of a few ulp of the [0, 1] range here, bit-exact at sigma = 1 above
    img = rng.integers(0, 65536, shape).astype(np.uint16)
    assert np.abs(W.gaussian(img, 2.0) - ow.gaussian(img, 2.0)).max() < 1e-15


def test_full_size_mask_path_properties():
    """C3/C4 size (2048x2048): idempotence and agreement with the oracle on a crop-independent statistic."""
    from tmat_b200 import transforms as T
    m = helpers.synth_mask(2048, 2048, 21, density=0.05)
    out = T.filter_branch_seg_mask(m)
    assert out.shape == m.shape and not (out & ~om.median_bool(m, om.disk(2))).any()
    again = T.filter_branch_seg_mask(out, None, True)
    assert np.array_equal(again, out)            # surviving components all have forks and low circularity
    lab, n = T.measure_label(out, 2, return_num=True)
    ref, nr = om.label(out, 2)
    assert n == nr and np.array_equal(lab, ref)


def test_threshold_and_apply_mask():
    import torch
    from tmat_b200 import _ffi, preprocessing as P
    ctx = _ffi.default_context()
    x = np.random.default_rng(0).random(10007)
    x[:3] = 0.5
    d = torch.from_numpy(x).cuda()
    m = torch.empty(x.size, dtype=torch.uint8, device="cuda")
    _ffi.check(_ffi.lib().tmat_threshold_gt_f64_dev(ctx.handle, _ffi.ptr(d), x.size, 0.5, _ffi.ptr(m), None))
    torch.cuda.synchronize()
    assert np.array_equal(m.cpu().numpy().astype(bool), x > 0.5)
    img = np.random.default_rng(1).random((33, 47)).astype(np.float32)
    mk = (np.random.default_rng(2).random((33, 47)) > 0.4).astype(np.uint8)
    ref = img.copy()
    ref[mk == 0] = 0
    assert np.array_equal(P.apply_mask(img, mk), ref)
    assert np.array_equal(P.bin_thresh(img, 255, 0.5), np.where(img > 0.5, 255, 0))


def test_generic_footprints_ownership_and_region_props():
    """VERDICT r1 weak #10 / missing #5: any odd-sized footprint for the median and the erosion (the reference's
    functions take any), in-place semantics of filter_branch_seg_mask(footprint=None) (transforms.py:356) and the
    area-like regionprops the reference's callers ask for (compute_branches.py:277)."""
    from scipy import ndimage as ndi
    from tmat_b200 import transforms as T
    m = helpers.synth_mask(150, 170, 31, density=0.15)
    rng = np.random.default_rng(3)
    fps = [np.ones((3, 3), np.uint8), np.ones((1, 5), np.uint8), om.disk(3), om.disk(9),
           (rng.random((5, 7)) < 0.6).astype(np.uint8), np.array([[0, 1, 1], [0, 1, 0], [1, 0, 0]], np.uint8)]
    for fp in fps:
        assert np.array_equal(T.median_filter(m, fp), om.median_bool(m, fp)), fp.shape
        assert np.array_equal(T.binary_erosion(m, fp), ndi.binary_erosion(m, structure=fp, border_value=True)), fp.shape
    sq = np.ones((3, 3), np.uint8)
    assert np.array_equal(T.filter_branch_seg_mask(m, sq, True), om.filter_branch_seg_mask(m, sq, True))
    # footprint=None cleans the caller's array in place and returns it; with a footprint the input is untouched
    mine = m.copy()
    ret = T.filter_branch_seg_mask(mine, None, True)
    assert ret is mine and np.array_equal(mine, om.filter_branch_seg_mask(m, None, True))
    keep = m.copy()
    T.filter_branch_seg_mask(keep)
    assert np.array_equal(keep, m)
    # equivalent_diameter_area = sqrt(4 * area / pi) per component (skimage.measure.regionprops)
    area = om.regionprops_image_area(m)
    assert np.array_equal(T.regionprops_image(m, "equivalent_diameter_area"), np.sqrt(4 * area / np.pi))
    with pytest.raises(NotImplementedError):
        T.regionprops_image(m, "solidity")


def test_skeleton_lut_override_changes_the_skeleton_and_restores():
    """tmat_skeleton_lut_set: the table is data (csrc/skeleton_lut.inc); a different table gives the skeleton the
    oracle computes with that same table."""
    from tmat_b200 import transforms as T
    m = helpers.synth_mask(120, 140, 17, density=0.15)
    base = om.ZS_LUT.copy()
    alt = base.copy()
    alt[[3, 6, 12]] = 1                 # any valid table: here three end-point codes deleted in sub-pass 1 only
    try:
        T.set_skeletonize_lut(alt)
        om.ZS_LUT[:] = alt
        assert np.array_equal(T.skeletonize(m), om.skeletonize(m))
    finally:
        om.ZS_LUT[:] = base
        T.set_skeletonize_lut(base)
    assert np.array_equal(T.skeletonize(m), om.skeletonize(m))


@pytest.mark.parametrize("h,w,seed", [(64, 64, 1), (97, 131, 2), (300, 340, 3), (1, 40, 5), (40, 1, 6), (512, 512, 7)])
def test_exact_edt_and_centerline_weighting_vs_scipy(h, w, seed):
    """SURVEY 8(f) N1 (compute_branches.py:340-344): exact EDT pinned by scipy.ndimage.distance_transform_edt, and
    pred * dist / (dist + centerline_dt) bit-identical to the numpy expression."""
    from scipy import ndimage as ndi
    from tmat_b200 import transforms as T
    m = helpers.synth_mask(h, w, seed, density=0.15)
    if m.all():
        m[0, 0] = False
    dist = ndi.distance_transform_edt(m)
    assert np.array_equal(T.distance_transform_edt(m), dist)
    sk = om.skeletonize(m)
    if not sk.any():
        sk[h // 2, w // 2] = True
    cdt = ndi.distance_transform_edt(np.logical_not(sk))
    assert np.array_equal(T.distance_transform_edt(sk, invert=True), cdt)
    pred = np.random.default_rng(seed).random((h, w))
    with np.errstate(invalid="ignore", divide="ignore"):
        ref = pred * (dist / (dist + cdt))
    got = T.centerline_weighting(pred, m, sk)
    assert np.array_equal(got, ref, equal_nan=True)


def test_edt_sparse_zeros_large_distances():
    from scipy import ndimage as ndi
    from tmat_b200 import transforms as T
    m = np.ones((700, 900), bool)
    m[13, 877] = m[650, 3] = m[300, 444] = False          # distances of several hundred pixels, columns without a zero
    assert np.array_equal(T.distance_transform_edt(m), ndi.distance_transform_edt(m))
