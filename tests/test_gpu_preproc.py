"""GPU: preprocessing.blur (A10) and exec_threshold (A8) vs the oracle (cv2 / scikit-learn restatements
pinned by tests/golden/thresholds.npz) and the frozen reference outputs."""
import numpy as np
import pytest

from oracle import thresholds as ot

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("k,itr", [(3, 1), (3, 3), (5, 2), (7, 1)])
def test_blur_uint8_bit_exact(k, itr):
    from tmat_b200 import preprocessing as prep
    rng = np.random.default_rng(k * 10 + itr)
    img = rng.integers(0, 256, (97, 131), dtype=np.uint8)
    img[20:40, 30:80] = 255
    out = prep.blur(img, itr, k)
    ref = ot.blur(img, itr, k)
    assert out.dtype == ref.dtype and np.array_equal(out, ref)
    assert np.array_equal(prep.blur(img, itr, k, gs=False), ot.blur(img, itr, k, gs=False))


def test_blur_float32_and_golden(golden):
    from tmat_b200 import preprocessing as prep
    g = golden("thresholds")
    img = g["img"]
    assert np.array_equal(prep.blur((img * 255).astype(np.uint8), 3), g["blur_u8"])
    out = prep.blur(img, 2, gs=False)                     # k = 3: same arithmetic as cv2 -> bit-exact
    assert out.dtype == np.float32 and np.array_equal(out, g["blur_f32"])
    big = (np.random.default_rng(5).random((70, 90), dtype=np.float32) * 200).astype(np.float32)
    for k in (3, 5, 7):
        raw = prep.blur(big, 2, k, gs=False)
        ref = ot.blur(big, 2, k, gs=False)
        assert np.abs(raw - ref).max() <= 1.0              # rounded values: at most one step apart
        assert (raw != ref).mean() < 1e-3
    assert np.array_equal(prep.blur(big, 1, 3, gs=True), ot.blur(big, 1, 3, gs=True))
    tiny = np.arange(6, dtype=np.uint8).reshape(1, 6) * 40
    assert np.array_equal(prep.blur(tiny, 2, 5), ot.blur(tiny, 2, 5))


def _bimodal(shape, seed, frac=0.3):
    rng = np.random.default_rng(seed)
    fg = rng.random(shape) < frac
    return np.clip(np.where(fg, rng.normal(0.7, 0.08, shape), rng.normal(0.2, 0.05, shape)), 0, 1).astype(np.float32)


@pytest.mark.parametrize("shape,seed,sd", [((96, 80), 31, 0.0), ((256, 256), 7, -0.5), ((512, 384), 9, 1.0)])
def test_gmm_threshold_vs_sklearn(shape, seed, sd):
    """Tolerance: scikit-learn fits in float32 and stops EM at |d lower bound| < 1e-3; the threshold agrees to
    1e-3 of the intensity range and the thresholded masks on >= 99.9 % of pixels."""
    from tmat_b200 import preprocessing as prep
    img = _bimodal(shape, seed)
    thr, par = prep.gmm_foreground_threshold(img.reshape(-1), sd, np.random.RandomState(123), return_params=True)
    ref = ot.gmm_foreground_threshold(img.reshape(-1), sd, np.random.RandomState(123))
    assert par["n_iter"] > 0
    assert abs(thr - ref) < 1e-3, (thr, ref, par)
    out = prep.exec_threshold(img, None, sd, np.random.RandomState(123))
    oref = ot.exec_threshold(img, None, sd, np.random.RandomState(123))
    assert out.dtype == oref.dtype and ((out > 0) == (oref > 0)).mean() >= 0.999


def test_exec_threshold_golden_and_rng_stream(golden):
    from tmat_b200 import preprocessing as prep
    g = golden("thresholds")
    img, mask = g["img"], g["mask"]
    masked = ot.apply_mask(img, mask)
    rs = np.random.RandomState(123)
    a = prep.exec_threshold(masked, None, 0.0, rs)
    assert ((a > 0) == (g["exec_threshold_all"] > 0)).mean() >= 0.999
    # the caller's RandomState advanced exactly as scikit-learn advances it (3 uniform draws)
    rs_ref = np.random.RandomState(123)
    ot.exec_threshold(masked, None, 0.0, rs_ref)
    assert rs.random_sample() == rs_ref.random_sample()
    b = prep.exec_threshold(masked, np.nonzero(mask), -0.5, np.random.RandomState(123))
    assert ((b > 0) == (g["exec_threshold_idx"] > 0)).mean() >= 0.999
    assert np.array_equal(b[b > 0], masked[b > 0])


def test_gen_circ_mask_bit_exact_vs_cv2():
    """preprocessing.py:96-116: cv2.circle(..., thickness=-1), OpenCV's integer midpoint circle."""
    import cv2
    from tmat_b200 import preprocessing as prep
    for (cx, cy), r, shape, val in (((80, 75), 33, (160, 170), 1), ((10, 12), 0, (30, 40), 255), ((5, 40), 25, (64, 48), 7),
                                    ((-4, 10), 20, (50, 60), 3), ((30, 30), 200, (61, 61), 1), ((20, 7), 1, (15, 41), 9)):
        ref = np.zeros(shape, np.uint8)
        cv2.circle(ref, (cx, cy), r, val, -1)
        out = prep.gen_circ_mask((cx, cy), r, shape, val)
        assert out.dtype == np.uint8 and np.array_equal(out, ref), ((cx, cy), r, shape)


def test_dt_blur_and_sdt_blur_vs_cv2_within_tolerance():
    """preprocessing.py:118-163.  The 5x5 chamfer transform is a closed form here and a float32 raster scan in OpenCV:
    distances agree to 1e-5 relative; after the reference's rounding the uint8 image differs in < 0.1 % of the pixels
    by one grey level."""
    import cv2
    import helpers
    from tmat_b200 import preprocessing as prep
    m = helpers.synth_mask(200, 230, 5, density=0.3)
    m[60:140, 80:170] = True                                   # a large blob: distances up to ~40
    img = m.astype(np.float32) * 0.7
    mask = np.where(img > 0, 255, 0).round().astype(np.uint8)
    ref_dt = cv2.distanceTransform(mask, cv2.DIST_L2, 5)
    got_dt = prep._chamfer5(mask)
    assert got_dt.dtype == np.float32
    assert np.abs(got_dt - ref_dt).max() <= 1e-5 * max(1.0, float(ref_dt.max()))
    ref_in = cv2.distanceTransform(np.logical_not(mask).astype(np.uint8), cv2.DIST_L2, 5)
    assert np.abs(prep._chamfer5(mask, invert=True) - ref_in).max() <= 1e-5 * max(1.0, float(ref_in.max()))
    # dt_blur: uint8 result of the reference's own composition (cv2 transform + the oracle's blur restatement)
    ref = ot.blur(ref_dt, 2, 3)
    out = prep.dt_blur(img, 2)
    assert out.dtype == ref.dtype == np.uint8
    d = np.abs(out.astype(int) - ref.astype(int))
    assert d.max() <= 1 and (d > 0).mean() < 1e-3, (int(d.max()), float((d > 0).mean()))
    ref_s = ot.blur(ref_dt - ref_in, 1, 3, gs=False)
    out_s = prep.sdt_blur(img, 1)
    assert out_s.dtype == np.float32 and np.abs(out_s - ref_s).max() <= 1.0
    assert (out_s != ref_s).mean() < 1e-3
