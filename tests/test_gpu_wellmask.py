"""GPU: the numpy-only pieces of generate_well_mask (well_mask_generation.py:16-140) and
regionprops 'eccentricity' (compute_branches.py:276), through the C ABI.

* get_superellipse_hull / gen_superellipse_mask / create_convex_hull_mask: bit-exact against outputs of the
  reference's own functions (tests/golden/wellmask.npz, np.random.seed set) and against the oracle restatement on
  further seeds; the hull mask also against scipy's Delaunay directly.
* eccentricity: 1e-6 absolute against the restated scikit-image algorithm (float64 dot products there, exact integer
  moments here; parity with scikit-image itself is unpinned).
"""
import numpy as np
import pytest

import helpers
from oracle import wellmask as ow, morphology as om

pytestmark = pytest.mark.gpu


def test_superellipse_hull_and_mask_golden(golden):
    from tmat_b200 import well_mask_generation as W
    g = golden("wellmask")
    for name in ("circ2", "sq8", "circ8", "odd3"):
        n, seed, iters = (int(v) for v in g[f"hull_{name}_cfg"])
        np.random.seed(seed)
        out = W.get_superellipse_hull(g[f"hull_{name}_x"], g[f"hull_{name}_y"], n, iters)
        assert np.array_equal(np.array(out), g[f"hull_{name}_out"]), name
        t, d, s_a, s_b, c_x, c_y = out
        for shp in ((200, 200), (150, 200), (37, 53)):
            m = W.gen_superellipse_mask(t, d * 0.9, s_a, s_b, c_x, c_y, n, shp)
            assert m.dtype == np.bool_ and m.shape == shp
            assert np.array_equal(np.packbits(m), g[f"mask_{name}_{shp[0]}x{shp[1]}"]), (name, shp)


@pytest.mark.parametrize("n", [2, 8, 4, 3])
def test_superellipse_hull_vs_oracle_more_seeds(n):
    from tmat_b200 import well_mask_generation as W
    for seed in (101, 102, 103):
        x, y = helpers.wellmask_points("circle" if n in (2, 3) else "square", seed, npts=17 + seed % 5)
        np.random.seed(seed)
        ref = ow.get_superellipse_hull(x, y, n, 5000)
        np.random.seed(seed)
        out = W.get_superellipse_hull(x, y, n, 5000)
        assert np.array_equal(np.array(out), np.array(ref)), (n, seed)
        # the global RNG has advanced by the same amount in both
        m = W.gen_superellipse_mask(*out, n, (64, 96))
        assert np.array_equal(m, ow.gen_superellipse_mask(*ref, n, (64, 96)))


def test_superellipse_hull_without_candidate_raises():
    from tmat_b200 import well_mask_generation as W
    x = np.array([-3.0, 3.0, 0.0]); y = np.array([0.0, 0.0, 3.0])        # far outside every parameter set
    np.random.seed(1)
    with pytest.raises(ValueError):
        ow.get_superellipse_hull(x, y, 2, 500)
    np.random.seed(1)
    with pytest.raises(ValueError):
        W.get_superellipse_hull(x, y, 2, 500)


def test_convex_hull_mask_golden_and_scipy(golden):
    from scipy.spatial import ConvexHull
    from tmat_b200 import well_mask_generation as W
    g = golden("wellmask")
    for name in ("a", "b", "c", "tri"):
        H, Wd = (int(v) for v in g[f"chull_{name}_shape"])
        verts = g[f"chull_{name}_verts"]
        m = W.create_convex_hull_mask((H, Wd), verts)
        assert np.array_equal(np.packbits(m), g[f"chull_{name}_mask"]), name
        # vertex order must not matter (the reference triangulates the point set)
        assert np.array_equal(W.create_convex_hull_mask((H, Wd), verts[::-1]), m)
        assert np.array_equal(W.create_convex_hull_mask((H, Wd), np.roll(verts, 2, axis=0)), m)
    rng = np.random.default_rng(9)
    for H, Wd, npts in ((64, 64, 12), (200, 173, 500), (90, 140, 5)):
        pts = np.stack([rng.integers(0, H, npts), rng.integers(0, Wd, npts)], axis=1)
        hv = pts[ConvexHull(pts).vertices]
        assert np.array_equal(W.create_convex_hull_mask((H, Wd), hv), ow.create_convex_hull_mask((H, Wd), hv))
        # interior points in the list change nothing
        assert np.array_equal(W.create_convex_hull_mask((H, Wd), pts), ow.create_convex_hull_mask((H, Wd), hv))
    with pytest.raises(Exception):
        W.create_convex_hull_mask((10, 10), np.array([[1, 1], [2, 2], [3, 3]]))


def test_regionprops_eccentricity_vs_oracle():
    from tmat_b200 import transforms as T
    m = np.zeros((80, 120), bool)
    m[5:25, 5:25] = True
    m[40, 10:60] = True
    m[50:58, 70:110] = True
    m[70, 5] = True
    e = T.regionprops_image(m, "eccentricity")
    assert e.dtype == np.float64 and e.shape == m.shape
    assert abs(e[10, 10]) < 1e-6 and e[40, 30] == 1.0 and e[70, 5] == 0.0 and e[0, 0] == 0.0
    assert abs(e[52, 80] - np.sqrt(1 - (8 ** 2 - 1) / (40 ** 2 - 1))) < 1e-12
    for shape, seed, dens in (((200, 260), 8, 0.2), ((513, 381), 9, 0.1), ((64, 2100), 10, 0.3)):
        mm = helpers.synth_mask(*shape, seed, density=dens)
        got, ref = T.regionprops_image(mm, "eccentricity"), om.regionprops_image_eccentricity(mm)
        assert np.array_equal(got == 0, ref == 0) or np.abs(got - ref).max() < 1e-6
        assert np.abs(got - ref).max() < 1e-6, float(np.abs(got - ref).max())
    # the threshold the reference applies (compute_branches.py:279) decides the same pixels
    mm = helpers.synth_mask(300, 300, 21, density=0.15)
    keep = T.regionprops_image(mm, "eccentricity") * T.regionprops_image(mm, "equivalent_diameter_area") > 3.5
    lab_area = om.regionprops_image_area(mm)
    keep_ref = om.regionprops_image_eccentricity(mm) * np.sqrt(4 * lab_area / np.pi) > 3.5
    assert np.array_equal(keep, keep_ref)
