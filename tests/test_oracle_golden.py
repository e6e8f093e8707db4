"""CPU: the oracle restatements reproduce the frozen outputs of the reference's own code."""
import numpy as np

import helpers
from oracle import stp, zstacks, thresholds, morphology as om


def test_stp_matches_reference_bitwise(golden):
    g = golden("stp")
    for name in ("a", "b", "c"):
        ws, sd = (int(v) for v in g[f"{name}_cfg"])
        out = stp.predict_img_with_smooth_windowing(g[f"{name}_in"], ws, sd, helpers.toy_pred_func)
        assert out.dtype == np.float64
        assert np.array_equal(out, g[f"{name}_out"])
    for ws in (32, 256, 320):
        assert np.array_equal(stp.spline_window_1d(ws), g[f"window1d_{ws}"])


def test_stp_single_canvas_identity(golden):
    """one canvas with D4 folded into tile load/store == the reference."""
    g = golden("stp")
    img = g["a_in"]
    ws, sd = (int(v) for v in g["a_cfg"])
    aug, step = stp.aug_of(ws, sd), ws // sd
    P = np.pad(img, aug, constant_values=img.min())
    PM, PN = P.shape
    w1 = stp.spline_window_1d(ws)
    W = w1[:, None] * w1[None, :]
    canvas = np.zeros(P.shape)
    for gi in range(8):
        Lm, Ln = (PM, PN) if gi % 2 == 0 else (PN, PM)
        for i in range(0, Lm - ws + 1, step):
            for j in range(0, Ln - ws + 1, step):
                r0, c0 = stp.tile_origin(gi, i, j, PM, PN, ws)
                tile = np.ascontiguousarray(stp.d4_do(P[r0:r0 + ws, c0:c0 + ws], gi))
                y = helpers.toy_pred_func(tile[None])[0, :, :, 0]
                canvas[r0:r0 + ws, c0:c0 + ws] += stp.d4_undo(y * W, gi)
    out = (canvas / (8 * sd * sd))[aug:-aug, aug:-aug][: img.shape[0], : img.shape[1]]
    assert np.abs(out - g["a_out"]).max() < 1e-14


def test_zproj_matches_reference_bitwise(golden):
    g = golden("zproj")
    for name in ("u16_odd", "u16_even", "u8_even"):
        st = g[f"{name}_in"]
        for m, f in zstacks.METHODS.items():
            out = f(st)
            ref = g[f"{name}_{m}"]
            assert out.dtype == ref.dtype and np.array_equal(out, ref), (name, m)


def test_thresholds_match_reference(golden):
    g = golden("thresholds")
    img, mask = g["img"], g["mask"]
    assert np.array_equal(thresholds.apply_mask(img, mask), g["apply_mask"])
    assert np.array_equal(thresholds.bin_thresh(img, 255, 0.5), g["bin_thresh"])
    masked = thresholds.apply_mask(img, mask)
    a = thresholds.exec_threshold(masked, None, 0.0, np.random.RandomState(123))
    assert np.array_equal(a, g["exec_threshold_all"])
    b = thresholds.exec_threshold(masked, np.nonzero(mask), -0.5, np.random.RandomState(123))
    assert np.array_equal(b, g["exec_threshold_idx"])
    assert np.array_equal(thresholds.blur((img * 255).astype(np.uint8), 3), g["blur_u8"])
    assert np.array_equal(thresholds.blur(img, 2, gs=False), g["blur_f32"])


def test_morphology_composition_matches_reference(golden):
    g = golden("morphology")
    for name in ("real", "synth"):
        shape = tuple(g[f"{name}_shape"])
        n = shape[0] * shape[1]
        unpack = lambda a: np.unpackbits(a)[:n].reshape(shape).astype(bool)
        m = unpack(g[f"{name}_in"])
        assert np.array_equal(om.filter_branch_seg_mask(m), unpack(g[f"{name}_filter_composition"]))
        assert np.array_equal(om.filter_branch_seg_mask(m, None, False),
                              unpack(g[f"{name}_filter_nomedian_keepisolated_composition"]))
        assert np.array_equal(om.remove_small_islands(m.astype(np.uint8), 30, 40, 1, 2).astype(bool),
                              unpack(g[f"{name}_islands_composition"]))
        assert np.array_equal(om.regionprops_image_area(m).astype(np.float32),
                              g[f"{name}_area_image_composition"])


def test_zhang_suen_lut_properties():
    """The shared table (tmat_b200/csrc/skeleton_lut.inc) is what the Zhang-Suen conditions give.  This pins the
    data file to its stated derivation -- NOT to scikit-image's shipped table (parity unpinned, see the file)."""
    lut = om.ZS_LUT
    assert np.array_equal(lut, om._zs_lut())
    assert lut[0] == 0 and lut[255] == 0          # isolated / interior pixels survive
    rot = lambda v: ((v << 4) | (v >> 4)) & 255   # 180-degree rotation swaps the two sub-passes
    swap = {0: 0, 1: 2, 2: 1, 3: 3}
    assert all(swap[int(lut[v])] == lut[rot(v)] for v in range(256))
    sk = om.skeletonize(helpers.synth_mask(120, 130, 5))
    # idempotent, subset of the input, and thin (no 2x2 block survives)
    assert np.array_equal(om.skeletonize(sk), sk)
    assert not (sk[:-1, :-1] & sk[1:, :-1] & sk[:-1, 1:] & sk[1:, 1:]).any()


def test_unet_oracle_shapes_and_flops():
    from oracle import unet
    assert abs(unet.flops_per_patch(320) / 1e9 - 31.026) < 0.01
    assert abs(unet.flops_per_patch(256) / 1e9 - 19.856) < 0.01
    w = unet.make_weights((8, 16, 32, 64), seed=3, randomize_affine=True)
    y = unet.forward(w, np.random.default_rng(0).random((2, 64, 64), dtype=np.float32))
    assert y.shape == (2, 64, 64, 1) and y.dtype == np.float32
    assert sum(v.size for v in unet.make_weights().values()) == 8197313


def test_wellmask_superellipse_and_hull_match_reference(golden):
    """well_mask_generation.py:16-140 (numpy-only in the reference): the restatements in oracle/wellmask.py against
    outputs of the reference's own functions, same np.random seed."""
    from oracle import wellmask as ow
    g = golden("wellmask")
    for name in ("circ2", "sq8", "circ8", "odd3"):
        n, seed, iters = (int(v) for v in g[f"hull_{name}_cfg"])
        np.random.seed(seed)
        out = ow.get_superellipse_hull(g[f"hull_{name}_x"], g[f"hull_{name}_y"], n, iters)
        assert np.array_equal(np.array(out), g[f"hull_{name}_out"]), name
        t, d, s_a, s_b, c_x, c_y = out
        for shp in ((200, 200), (150, 200), (37, 53)):
            m = ow.gen_superellipse_mask(t, d * 0.9, s_a, s_b, c_x, c_y, n, shp)
            assert m.shape == shp
            assert np.array_equal(np.packbits(m), g[f"mask_{name}_{shp[0]}x{shp[1]}"]), (name, shp)
    for name in ("a", "b", "c", "tri"):
        H, W = (int(v) for v in g[f"chull_{name}_shape"])
        m = ow.create_convex_hull_mask((H, W), g[f"chull_{name}_verts"])
        assert np.array_equal(np.packbits(m), g[f"chull_{name}_mask"]), name


def test_eccentricity_oracle_known_shapes():
    """The restated regionprops eccentricity on shapes with a closed form: a disk / square -> 0, a 1 x L line -> 1,
    an a x b rectangle -> sqrt(1 - (b^2 - 1) / (a^2 - 1)) (variance of k equally spaced points = (k^2 - 1) / 12)."""
    m = np.zeros((80, 120), bool)
    m[5:25, 5:25] = True                      # 20 x 20 square
    m[40, 10:60] = True                       # 1 x 50 line
    m[50:58, 70:110] = True                   # 8 x 40 rectangle
    m[70, 5] = True                           # single pixel
    e = om.regionprops_image_eccentricity(m)
    assert abs(e[10, 10]) < 1e-6
    assert e[40, 30] == 1.0
    assert abs(e[52, 80] - np.sqrt(1 - (8 ** 2 - 1) / (40 ** 2 - 1))) < 1e-12
    assert e[70, 5] == 0.0 and e[0, 0] == 0.0
