"""GPU: tile gather + blend (generic pred_func path) through the C ABI vs golden / oracle."""
import numpy as np
import pytest

import helpers
from oracle import stp as ostp

pytestmark = pytest.mark.gpu


def test_stp_generic_path_golden_bitexact(golden):
    from tmat_b200.smooth_tiled_predictions import predict_img_with_smooth_windowing
    g = golden("stp")
    for name in ("a", "b", "c"):
        ws, sd = (int(v) for v in g[f"{name}_cfg"])
        out = predict_img_with_smooth_windowing(g[f"{name}_in"], ws, sd, helpers.toy_pred_func)
        assert out.dtype == np.float64 and out.shape == g[f"{name}_in"].shape
        assert np.array_equal(out, g[f"{name}_out"]), (name, np.abs(out - g[f"{name}_out"]).max())


def test_stp_gather_matches_reference_tiles():
    from tmat_b200 import _ffi
    from tmat_b200.smooth_tiled_predictions import geometry
    img = helpers.synth_projection(90, 70, 3)
    ws, sd = 32, 2
    geo = geometry(90, 70, ws, sd)
    aug, step = geo["aug"], geo["step"]
    P = np.pad(img, aug, constant_values=img.min())
    ctx = _ffi.default_context()
    for g in range(8):
        rot = np.ascontiguousarray(ostp.d4_do(P, g))
        a, b = geo["tiles_even"] if g % 2 == 0 else geo["tiles_odd"]
        tiles = np.empty((a * b, ws, ws), np.float32)
        _ffi.check(_ffi.lib().tmat_stp_gather_host(ctx.handle, _ffi.ptr(img), 90, 70, ws, sd, g,
                                                   _ffi.ptr(tiles)))
        ref = np.stack([rot[i * step:i * step + ws, j * step:j * step + ws]
                        for i in range(a) for j in range(b)])
        assert np.array_equal(tiles, ref), g


@pytest.mark.parametrize("h,w,ws,sd", [(64, 64, 32, 2), (50, 81, 16, 4), (33, 47, 32, 2), (320, 200, 64, 2)])
def test_stp_generic_vs_oracle(h, w, ws, sd):
    from tmat_b200.smooth_tiled_predictions import predict_img_with_smooth_windowing
    img = helpers.synth_projection(h, w, h + w)
    out = predict_img_with_smooth_windowing(img, ws, sd, helpers.toy_pred_func)
    ref = ostp.predict_img_with_smooth_windowing(img, ws, sd, helpers.toy_pred_func)
    assert np.array_equal(out, ref)


def test_stp_partition_of_unity_interior():
    """An all-ones prediction blends to exactly 1 wherever every axis is multiply covered."""
    from tmat_b200.smooth_tiled_predictions import predict_img_with_smooth_windowing
    img = np.zeros((128, 128), np.float32)
    out = predict_img_with_smooth_windowing(
        img, 32, 2, lambda b, verbose=0: np.ones(b.shape + (1,), np.float32))
    assert np.abs(out - 1).max() < 1e-12


@pytest.mark.parametrize("world", [2, 3])
def test_row_strips_are_bit_identical_to_the_full_prediction(world):
    """Owner-computes row strips (one large image sharded over GPUs): concatenated strips == full image."""
    from oracle import unet as ou
    from tmat_b200 import models
    from tmat_b200.smooth_tiled_predictions import (predict_img_with_smooth_windowing, strip_rows,
                                                    predict_rows_with_smooth_windowing, patches_for_rows)
    fc, ws = (64, 64, 128, 128), 64
    net = models.CudaUNet(ou.make_weights(fc, seed=2, randomize_affine=True, head_gain=4.0), ws, fc, impl="tc")
    img = helpers.synth_projection(200, 150, 77)
    full = predict_img_with_smooth_windowing(img, ws, 2, net.predict)
    parts, n_eval = [], 0
    for r in range(world):
        row0, row1 = strip_rows(200, r, world)
        parts.append(predict_rows_with_smooth_windowing(img, ws, 2, net.predict, row0, row1))
        n_eval += len(patches_for_rows(200, 150, ws, 2, row0, row1))
    assert np.array_equal(np.concatenate(parts), full)
    n_total = len(patches_for_rows(200, 150, ws, 2, 0, 200))
    assert n_total < n_eval < 2 * n_total


@pytest.mark.gpu
def test_pinned_result_buffers_are_recycled_and_writable():
    """Results of the all-device path live in cudaHostAlloc memory (tmat_host_alloc): a dropped result's
    buffer is handed out again, a live one never is, and the arrays behave like ordinary numpy arrays."""
    import gc
    from tmat_b200 import _ffi
    a = _ffi.pinned_empty((64, 48), np.float64)
    assert a.shape == (64, 48) and a.dtype == np.float64 and a.flags["C_CONTIGUOUS"] and a.flags["WRITEABLE"]
    a[:] = 3.5
    addr_a = a.ctypes.data
    b = _ffi.pinned_empty((64, 48), np.float64)
    assert b.ctypes.data != addr_a          # a is still alive
    view = a[10:20]
    del a
    gc.collect()
    c = _ffi.pinned_empty((64, 48), np.float64)
    assert c.ctypes.data != addr_a          # the view keeps a's buffer alive
    assert float(view[0, 0]) == 3.5
    del view
    gc.collect()
    d = _ffi.pinned_empty((64, 48), np.float64)
    assert d.ctypes.data == addr_a          # recycled
    assert _ffi.pinned_empty((0, 4), np.float32).size == 0
