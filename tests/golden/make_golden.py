"""Generate tests/golden/*.npz by running the REFERENCE's own code in the build container.

Run from the repo root:  python tests/golden/make_golden.py
Needs /root/reference (read-only); never runs on the GPU box -- the fixtures it writes
are what travels.  What is executed:

* ``smooth_tiled_predictions`` and ``zstacks`` are imported unchanged.
* ``preprocessing`` is imported unchanged with two stand-ins in ``sys.modules``: an empty
  ``dask`` (absent here; only used by training helpers) and a ``defs`` holding
  ``MAX_UINT8`` (the real ``defs.py`` needs a generated ``package.cfg``).
* ``transforms`` is imported unchanged over a stand-in ``skimage`` whose primitives are
  ``oracle.morphology``'s restatements (scikit-image is not installable here).  That pins
  the reference's *composition* (networkx graph, fork test, removal rule); the primitives
  stay unpinned and the fixture names say so (``*_composition``).
"""
import os
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
REF = "/root/reference"
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, REF)
OUT = os.path.join(ROOT, "tests", "golden")

import helpers  # noqa: E402
from oracle import morphology as om  # noqa: E402


def _install_stubs():
    dask = types.ModuleType("dask")
    dask.delayed = lambda f: f
    sys.modules["dask"] = dask
    pkg = types.ModuleType("fl_tissue_model_tools")
    pkg.__path__ = [os.path.join(REF, "fl_tissue_model_tools")]
    sys.modules["fl_tissue_model_tools"] = pkg
    defs = types.ModuleType("fl_tissue_model_tools.defs")
    defs.MAX_UINT8 = np.iinfo(np.uint8).max
    defs.MAX_UINT16 = np.iinfo(np.uint16).max
    sys.modules["fl_tissue_model_tools.defs"] = defs
    pkg.defs = defs

    sk = types.ModuleType("skimage")
    filters = types.ModuleType("skimage.filters")
    filters.median = lambda image, footprint=None: om.median_bool(image, footprint)
    filters.apply_hysteresis_threshold = None
    measure = types.ModuleType("skimage.measure")
    measure.label = lambda m, connectivity=None: om.label(m, 2 if connectivity is None else connectivity)[0]

    class _Region:
        def __init__(self, area, perim):
            self.area, self.perimeter = area, perim

    def regionprops(lab):
        n = int(lab.max())
        areas = om.region_areas(lab, n)
        perims = om.perimeter_code_hist(lab, n) @ om.PERIM_WEIGHTS
        return [_Region(a, p) for a, p in zip(areas, perims)]

    def regionprops_table(lab, properties):
        assert list(properties) == ["area"]
        return {"area": om.region_areas(lab, int(lab.max())).astype(np.float64)}

    measure.regionprops = regionprops
    measure.regionprops_table = regionprops_table
    morph = types.ModuleType("skimage.morphology")
    morph.remove_small_objects = lambda lab, min_size=64: om.remove_small_objects(lab, min_size)
    morph.disk = om.disk
    morph.skeletonize = om.skeletonize
    # well_mask_generation.py imports these at module level; the functions pinned from it (get_superellipse_hull,
    # gen_superellipse_mask, create_convex_hull_mask) are numpy / scipy only and never call them
    filters.gaussian = filters.threshold_otsu = None
    morph.binary_erosion = None
    feature = types.ModuleType("skimage.feature")
    feature.canny = None
    exposure = types.ModuleType("skimage.exposure")
    exposure.rescale_intensity = None
    transform = types.ModuleType("skimage.transform")
    transform.rescale = transform.resize = None
    draw = types.ModuleType("skimage.draw")
    draw.disk = None
    for name, mod in (("skimage", sk), ("skimage.filters", filters), ("skimage.measure", measure),
                      ("skimage.morphology", morph), ("skimage.feature", feature), ("skimage.exposure", exposure),
                      ("skimage.transform", transform), ("skimage.draw", draw)):
        sys.modules[name] = mod


def gen_stp():
    from fl_tissue_model_tools import smooth_tiled_predictions as r
    cases = {}
    for name, (h, w, ws, sd, seed) in {"a": (100, 72, 32, 2, 11), "b": (40, 53, 16, 4, 12),
                                       "c": (64, 64, 32, 2, 13)}.items():
        img = helpers.synth_projection(h, w, seed)
        out = r.predict_img_with_smooth_windowing(img, ws, sd, helpers.toy_pred_func)
        cases[f"{name}_in"] = img
        cases[f"{name}_cfg"] = np.array([ws, sd])
        cases[f"{name}_out"] = out
    cases["window1d_320"] = r._spline_window(320)
    cases["window1d_256"] = r._spline_window(256)
    cases["window1d_32"] = r._spline_window(32)
    np.savez_compressed(os.path.join(OUT, "stp.npz"), **cases)


def gen_zproj():
    from fl_tissue_model_tools import zstacks as r
    cases = {}
    for name, (z, h, w, dt, seed) in {"u16_odd": (7, 61, 83, np.uint16, 21), "u16_even": (8, 48, 50, np.uint16, 22),
                                      "u8_even": (6, 40, 45, np.uint8, 23)}.items():
        st = helpers.synth_stack(z, h, w, seed, dt)
        cases[f"{name}_in"] = st
        for m, f in (("min", r.proj_min), ("max", r.proj_max), ("avg", r.proj_avg), ("med", r.proj_med),
                     ("fs", r.proj_focus_stacking)):
            cases[f"{name}_{m}"] = f(st)
    # float32 stacks: the reference's np.max/min/mean/median take any dtype and stay float32 (zstacks.py:192-249)
    for name, (z, h, w, seed) in {"f32_odd": (5, 37, 41, 24), "f32_even": (12, 30, 52, 25), "f32_z32": (32, 24, 36, 26)}.items():
        rng = np.random.default_rng(seed)
        st = (helpers.synth_stack(z, h, w, seed).astype(np.float32) / np.float32(65535.0)
              + rng.normal(0, 1e-3, (z, h, w)).astype(np.float32))
        st[:, :4, :4] = np.float32(0.25)          # ties
        cases[f"{name}_in"] = st
        for m, f in (("min", r.proj_min), ("max", r.proj_max), ("avg", r.proj_avg), ("med", r.proj_med)):
            cases[f"{name}_{m}"] = f(st)
            assert cases[f"{name}_{m}"].dtype == np.float32
    np.savez_compressed(os.path.join(OUT, "zproj.npz"), **cases)


def gen_thresholds():
    from fl_tissue_model_tools import preprocessing as r
    rng = np.random.default_rng(31)
    img = np.clip(np.where(rng.random((96, 80)) < 0.3, rng.normal(0.7, 0.08, (96, 80)),
                           rng.normal(0.2, 0.05, (96, 80))), 0, 1).astype(np.float32)
    mask = (np.hypot(*np.mgrid[-48:48, -40:40]) < 38).astype(np.uint8)
    cases = {"img": img, "mask": mask}
    cases["apply_mask"] = r.apply_mask(img, mask)
    cases["bin_thresh"] = r.bin_thresh(img, 255, 0.5)
    masked = r.apply_mask(img, mask)
    idx = np.nonzero(mask)
    cases["exec_threshold_all"] = r.exec_threshold(masked, None, 0.0, np.random.RandomState(123))
    cases["exec_threshold_idx"] = r.exec_threshold(masked, idx, -0.5, np.random.RandomState(123))
    u8 = (img * 255).astype(np.uint8)
    cases["blur_u8"] = r.blur(u8, 3)
    cases["blur_f32"] = r.blur(img, 2, gs=False)
    np.savez_compressed(os.path.join(OUT, "thresholds.npz"), **cases)


def gen_morphology():
    from PIL import Image
    from fl_tissue_model_tools import transforms as r
    cases = {}
    real = np.array(Image.open(os.path.join(REF, "notebooks/topology/sample_data/mask.tif"))) > 0
    crops = {"real": real[700:1212, 600:1112], "synth": helpers.synth_mask(300, 340, 41)}
    for name, m in crops.items():
        cases[f"{name}_in"] = np.packbits(m)
        cases[f"{name}_shape"] = np.array(m.shape)
        cases[f"{name}_filter_composition"] = np.packbits(r.filter_branch_seg_mask(m.copy(), om.disk(2), True))
        cases[f"{name}_filter_nomedian_keepisolated_composition"] = np.packbits(
            r.filter_branch_seg_mask(m.copy(), None, False))
        cases[f"{name}_islands_composition"] = np.packbits(
            r.remove_small_islands(m.astype(np.uint8), 30, 40, 1, 2).astype(bool))
        cases[f"{name}_area_image_composition"] = r.regionprops_image(m, "area").astype(np.float32)
    np.savez_compressed(os.path.join(OUT, "morphology.npz"), **cases)


def gen_dmt():
    """The reference's own numba implementation (dmtgraph.py:38-99) on small weighted-prediction-like images."""
    from fl_tissue_model_tools import dmtgraph as r
    cases = {}
    for name, (h, w, d1, d2, seed) in {"a": (40, 52, 5.0, 0.0, 1), "b": (64, 64, 20.0, 10.0, 2), "c": (97, 131, 12.0, 0.0, 3),
                                       "d": (192, 160, 8.0, 2.0, 4), "flat": (16, 16, 1.0, 0.0, 5)}.items():
        img = (helpers.synth_projection(h, w, seed) * 255).astype(np.float32)
        img[:5, :7] = 0                           # exact zeros: their edges are dropped (np.isclose, :73-79)
        if name == "flat":
            img[:] = 7.0                          # every edge ties: the index tie-breaks decide everything
        if name == "d":
            img = np.round(img / 8) * 8           # heavy value ties
        v, e = r.compute_dmt_graph(img.copy(), d1, d2)
        cases[f"{name}_in"] = img
        cases[f"{name}_cfg"] = np.array([d1, d2], np.float32)
        cases[f"{name}_vertices"] = v
        cases[f"{name}_edges"] = e
    np.savez_compressed(os.path.join(OUT, "dmt.npz"), **cases)


def gen_wellmask():
    """The reference's own numpy-only well-mask functions (well_mask_generation.py:16-140), np.random.seed set."""
    from fl_tissue_model_tools import well_mask_generation as r
    from scipy.spatial import ConvexHull
    cases = {}
    for name, (kind, n, seed, iters) in {"circ2": ("circle", 2, 11, 25000), "sq8": ("square", 8, 12, 25000),
                                         "circ8": ("circle", 8, 13, 6000), "odd3": ("circle", 3, 14, 4000)}.items():
        x, y = helpers.wellmask_points(kind, seed)
        np.random.seed(seed)
        out = r.get_superellipse_hull(x, y, n, iters)
        cases[f"hull_{name}_cfg"] = np.array([n, seed, iters], np.int64)
        cases[f"hull_{name}_x"], cases[f"hull_{name}_y"] = x, y
        cases[f"hull_{name}_out"] = np.array(out, np.float64)
        for shp in ((200, 200), (150, 200), (37, 53)):
            t, d, s_a, s_b, c_x, c_y = out
            m = r.gen_superellipse_mask(t, d * 0.9, s_a, s_b, c_x, c_y, n, shp)
            cases[f"mask_{name}_{shp[0]}x{shp[1]}"] = np.packbits(m)
    rng = np.random.default_rng(5)
    for name, (H, W, npts) in {"a": (120, 160, 40), "b": (200, 200, 300), "c": (33, 47, 6), "tri": (50, 60, 3)}.items():
        pts = np.stack([rng.integers(0, H, npts), rng.integers(0, W, npts)], axis=1)
        hv = pts[ConvexHull(pts).vertices]
        cases[f"chull_{name}_shape"] = np.array([H, W], np.int64)
        cases[f"chull_{name}_verts"] = hv.astype(np.int64)
        cases[f"chull_{name}_mask"] = np.packbits(r.create_convex_hull_mask((H, W), hv))
    np.savez_compressed(os.path.join(OUT, "wellmask.npz"), **cases)


if __name__ == "__main__":
    _install_stubs()
    gen_wellmask()
    gen_dmt()
    gen_stp()
    gen_zproj()
    gen_thresholds()
    gen_morphology()
    for f in sorted(os.listdir(OUT)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(OUT, f)))
