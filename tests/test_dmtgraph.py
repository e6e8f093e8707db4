"""The Morse graph: native host code of libtmat_b200 vs the reference's own numba implementation
(dmtgraph.py:38-99) -- golden fixtures written by tests/golden/make_golden.py from the imported reference, plus a live
comparison where the reference is installed (baseline/_ref).  Same vertices, same edges, same ORDER.  No GPU needed:
the algorithm is a sequential union-find sweep and runs on the host inside the library."""
import os
import sys

import numpy as np
import pytest

import helpers

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_dmt_graph_matches_reference_golden(golden):
    from tmat_b200 import dmtgraph
    g = golden("dmt")
    for name in ("a", "b", "c", "d", "flat"):
        d1, d2 = (float(x) for x in g[f"{name}_cfg"])
        v, e = dmtgraph.compute_dmt_graph(g[f"{name}_in"], d1, d2)
        assert v.dtype == np.int32 and e.dtype == np.int32
        assert np.array_equal(v, g[f"{name}_vertices"]), name
        assert np.array_equal(e, g[f"{name}_edges"]), name
    assert len(g["c_edges"]) > 1000          # the fixtures are not trivial


def test_dmt_graph_live_against_installed_reference():
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.isdir(os.path.join(ref_dir, "fl_tissue_model_tools")):
        pytest.skip("baseline/_ref not installed")
    sys.path.insert(0, ref_dir)
    try:
        from fl_tissue_model_tools import dmtgraph as ref
    except Exception as e:   # noqa: BLE001 -- numba missing
        pytest.skip(f"reference dmtgraph not importable: {e}")
    finally:
        sys.path.remove(ref_dir)
    from tmat_b200 import dmtgraph
    img = (helpers.synth_projection(120, 96, 17) * 255).astype(np.float32)
    for d1, d2 in ((3.0, 0.0), (25.0, 5.0)):
        rv, re_ = ref.compute_dmt_graph(img.copy(), d1, d2)
        v, e = dmtgraph.compute_dmt_graph(img, d1, d2)
        assert np.array_equal(v, rv) and np.array_equal(e, re_), (d1, d2)
    with pytest.raises(ValueError):
        dmtgraph.compute_dmt_graph(np.zeros((3, 4, 5), np.float32), 1.0)


def test_hull_order_matches_scipy_convex_hull():
    """tmat_b200.well_mask_generation._hull_order (the exact integer monotone chain that orders the hull vertices for
    tmat_convex_hull_mask_host) returns scipy's ConvexHull vertex set, as a simple polygon, for any point list."""
    from scipy.spatial import ConvexHull
    from tmat_b200.well_mask_generation import _hull_order
    rng = np.random.default_rng(3)
    for n, (H, W) in ((3, (50, 60)), (7, (40, 40)), (60, (200, 173)), (400, (64, 64))):
        pts = np.stack([rng.integers(0, H, n), rng.integers(0, W, n)], axis=1)
        poly = _hull_order(pts)
        ref = ConvexHull(pts)
        assert sorted(poly) == sorted(map(tuple, pts[ref.vertices].tolist()))
        # same orientation all the way round (strictly convex polygon) and the same area
        a = np.array(poly, dtype=np.int64)
        b = np.roll(a, -1, axis=0)
        c = np.roll(a, -2, axis=0)
        cross = (b[:, 0] - a[:, 0]) * (c[:, 1] - b[:, 1]) - (b[:, 1] - a[:, 1]) * (c[:, 0] - b[:, 0])
        assert (cross > 0).all() or (cross < 0).all()
        area2 = abs(int(np.sum(a[:, 0] * b[:, 1] - b[:, 0] * a[:, 1])))
        assert abs(area2 / 2 - ref.volume) < 1e-9
    assert len(_hull_order(np.array([[1, 1], [2, 2], [3, 3], [5, 5]]))) < 3      # collinear: no polygon
