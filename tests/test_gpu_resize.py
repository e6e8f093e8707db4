"""GPU: the resizes either side of the path against the libraries the reference calls.

* Pillow (models.py:630-651): bit-exact -- ImagingResample is scalar C with double taps.
* OpenCV (compute_branches.py:309-312, compute_cell_area.py:53-56): tolerance -- the wheel's cv2.resize goes through
  vendor (IPP) code whose rounding differs from OpenCV's own C++ loops (seen here: a 1-row image and a 2-row image
  with identical rows give different float32 results): <= 1 grey level on integer images, 1e-5 of the range on float32.
"""
import numpy as np
import pytest

import helpers

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("h,w,wo,ho", [(40, 52, 33, 25), (37, 41, 41, 60), (64, 64, 40, 40), (30, 50, 50, 19),
                                       (200, 200, 125, 125), (97, 131, 200, 150), (5, 7, 3, 2), (1, 9, 4, 1)])
def test_pil_lanczos_and_nearest_bit_exact(h, w, wo, ho):
    from PIL import Image
    from tmat_b200 import resize as rz
    a = np.random.default_rng(h * w).random((h, w)).astype(np.float32) * 3 - 1
    ref = np.array(Image.fromarray(a).resize((wo, ho), resample=Image.Resampling.LANCZOS))
    out = rz.pil_resize(a, (wo, ho), rz.LANCZOS)
    assert out.dtype == ref.dtype == np.float32 and out.shape == ref.shape
    assert np.array_equal(out, ref), float(np.abs(out - ref).max())
    ref = np.array(Image.fromarray(a).resize((wo, ho), resample=Image.Resampling.NEAREST))
    assert np.array_equal(rz.pil_resize(a, (wo, ho), rz.NEAREST), ref)
    # float64 input goes through Image.fromarray's float32 conversion (the prediction, models.py:647)
    a64 = a.astype(np.float64) + 1e-9
    ref = np.array(Image.fromarray(a64).resize((wo, ho), resample=Image.Resampling.NEAREST))
    assert np.array_equal(rz.pil_resize(a64, (wo, ho), rz.NEAREST), ref)


def test_pil_full_size_ds_ratio():
    """The shipped config: 0.625 x a 2048^2 field (unet_patch_segmentor_1.json) and back."""
    from PIL import Image
    from tmat_b200 import resize as rz
    a = helpers.synth_projection(2048, 2048, 77)
    ref = np.array(Image.fromarray(a).resize((1280, 1280), resample=Image.Resampling.LANCZOS))
    out = rz.pil_resize(a, (1280, 1280))
    assert np.array_equal(out, ref)
    back = np.array(Image.fromarray(ref).resize((2048, 2048), resample=Image.Resampling.NEAREST))
    assert np.array_equal(rz.pil_resize(out, (2048, 2048), rz.NEAREST), back)


@pytest.mark.parametrize("dtype", [np.uint8, np.uint16, np.float32])
@pytest.mark.parametrize("h,w,wo,ho", [(97, 131, 82, 61), (97, 131, 200, 150), (120, 120, 75, 75), (64, 200, 40, 97)])
def test_cv2_linear_and_lanczos4_within_tolerance(dtype, h, w, wo, ho):
    import cv2
    from tmat_b200 import resize as rz
    top = {np.uint8: 255, np.uint16: 60000, np.float32: 1.0}[dtype]
    base = helpers.synth_projection(h, w, 5) * 0.7 + np.random.default_rng(1).random((h, w)) * 0.3
    img = (base * top).astype(dtype)
    for interp, name in ((cv2.INTER_LINEAR, "linear"), (cv2.INTER_LANCZOS4, "lanczos4")):
        ref = cv2.resize(img, (wo, ho), interpolation=interp)
        out = rz.resize(img, (wo, ho), interpolation=interp)
        assert out.dtype == ref.dtype and out.shape == ref.shape
        d = np.abs(out.astype(np.float64) - ref.astype(np.float64))
        tol = 1.0 if dtype != np.float32 else 1e-5 * 2.0          # Lanczos overshoots [0, 1] a little: range ~ 2
        print(f"cv2 {name} {np.dtype(dtype).name} {(h, w)}->{(ho, wo)}: max diff {d.max():.3g}, differing {float((d > 0).mean()):.4f}")
        assert d.max() <= tol, (name, float(d.max()))
        if dtype != np.float32:
            # uint8 follows OpenCV's own fixed-point loops and is all but identical; uint16 goes through float taps
            assert (d > 0).mean() < (0.005 if dtype == np.uint8 else 0.02), (name, float((d > 0).mean()))
    # the positional quirk of compute_cell_area.py:56: the flag sits in the `dst` slot -> bilinear
    quirk = cv2.resize(img, (wo, ho), cv2.INTER_AREA)
    assert np.array_equal(quirk, cv2.resize(img, (wo, ho), interpolation=cv2.INTER_LINEAR))
    assert np.array_equal(rz.resize(img, (wo, ho), rz.INTER_AREA), rz.resize(img, (wo, ho), interpolation=rz.INTER_LINEAR))


def test_segmentor_auto_resample_matches_the_reference_flow():
    """UNetXceptionPatchSegmentor.predict(x) with the reference's default auto_resample=True (models.py:624-653):
    same Pillow round trip around the same tiled prediction; result float32 at the input's shape."""
    from PIL import Image
    from oracle import unet as ou
    from tmat_b200 import models
    from tmat_b200.smooth_tiled_predictions import predict_img_with_smooth_windowing
    fc = (64, 128, 256, 512)
    w = ou.make_weights(fc, seed=3)
    seg = models.UNetXceptionPatchSegmentor(64, w, fc, ds_ratio=0.625)
    x = helpers.synth_projection(200, 200, 12)
    got = seg.predict(x)
    small = np.array(Image.fromarray(x).resize((125, 125), resample=Image.Resampling.LANCZOS))
    pred = predict_img_with_smooth_windowing(small, 64, 2, seg.model.predict)
    ref = np.array(Image.fromarray(pred).resize((200, 200), resample=Image.Resampling.NEAREST))
    assert got.shape == (200, 200) and got.dtype == np.float32
    assert np.array_equal(got, ref)
    same = seg.predict(x, auto_resample=False)
    assert same.shape == x.shape and same.dtype == np.float64
    # a missing checkpoint raises instead of silently random-initialising
    with pytest.raises((FileNotFoundError, ValueError)):
        models.UNetXceptionPatchSegmentor(64, "no_such_checkpoint.h5", fc)
    with pytest.raises(FileNotFoundError):
        models.UNetXceptionPatchSegmentor(64, "no_such_checkpoint.npz", fc)
    assert models.UNetXceptionPatchSegmentor(64, "random", fc).ds_ratio == 0.5
