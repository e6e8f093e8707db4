"""Put the B200 path behind the reference's own import names, so that the reference's scripts
(``scripts/compute_branches.py:38-49``, ``compute_zproj.py:17-23``, ``compute_cell_area.py:17-22``) run unchanged:

    import tmat_b200.shim; tmat_b200.shim.install()          # before the script's own imports
    from fl_tissue_model_tools import zstacks as zs           # -> proj_* now run on the GPU

What ``install()`` does, per reference module:

* the module imports in this environment (its own third-party dependencies are present): the hot-path functions are
  replaced IN it, everything else (``zstacks.find_zstack_files``, argument parsing, I/O ...) stays the reference's;
* the module does not import (TensorFlow / scikit-image / dask absent): the ``tmat_b200`` mirror is registered under
  the reference's name wholesale -- it holds the hot-path functions only, which is all the scripts' hot path uses.

``fl_tissue_model_tools`` itself must be importable (an installed reference, or ``reference_root`` pointing at a
directory that contains the package, e.g. this repository's git-ignored ``baseline/_ref``).  ``uninstall()`` restores
every replaced attribute and module.
"""
import importlib
import sys

_PKG = "fl_tissue_model_tools"

# reference module -> (tmat_b200 mirror, hot-path names that are replaced)
PATCHES = {
    "zstacks": ("tmat_b200.zstacks", ("proj_focus_stacking", "proj_avg", "proj_med", "proj_max", "proj_min")),
    "smooth_tiled_predictions": ("tmat_b200.smooth_tiled_predictions", ("predict_img_with_smooth_windowing",)),
    "transforms": ("tmat_b200.transforms", ("filter_branch_seg_mask", "remove_small_islands", "regionprops_image")),
    "well_mask_generation": ("tmat_b200.well_mask_generation", ("auto_threshold_well", "get_superellipse_hull",
                                                                "gen_superellipse_mask", "create_convex_hull_mask")),
    "preprocessing": ("tmat_b200.preprocessing", ("apply_mask", "bin_thresh", "exec_threshold", "blur", "dt_blur", "sdt_blur",
                                                   "gen_circ_mask")),
    "dmtgraph": ("tmat_b200.dmtgraph", ("compute_dmt_graph",)),
    "models": ("tmat_b200.models", ("UNetXceptionPatchSegmentor", "get_unet_patch_segmentor_from_cfg",
                                    "predict_img_with_smooth_windowing")),
}

_undo = []          # (kind, target, name, old) in application order
_installed = False


def _set_attr(mod, name, new):
    _undo.append(("attr", mod, name, getattr(mod, name, _undo)))       # _undo doubles as the "was absent" marker
    setattr(mod, name, new)


def _set_module(fullname, new):
    _undo.append(("module", None, fullname, sys.modules.get(fullname, _undo)))
    sys.modules[fullname] = new


def _with_fallback(ours, theirs):
    """regionprops_image: properties this library does not compute fall through to the reference's own function."""
    def regionprops_image(mask, prop):
        try:
            return ours(mask, prop)
        except NotImplementedError:
            return theirs(mask, prop)
    regionprops_image.__doc__ = ours.__doc__
    return regionprops_image


def install(reference_root=None, modules=None):
    """Install the mirrors; returns {reference module: "patched" | "replaced"}.  ``modules`` limits the set
    (names of PATCHES); ``reference_root``: directory containing ``fl_tissue_model_tools`` when it is not installed."""
    global _installed
    if _installed:
        uninstall()
    if reference_root and reference_root not in sys.path:
        sys.path.insert(0, reference_root)
        _undo.append(("path", None, reference_root, None))
    try:
        pkg = importlib.import_module(_PKG)
    except Exception as e:   # noqa: BLE001
        raise ImportError(f"tmat_b200.shim: '{_PKG}' is not importable ({type(e).__name__}: {e}); install the reference "
                          f"or pass reference_root=<directory containing {_PKG}>") from e
    report = {}
    for leaf in (modules or PATCHES):
        ours_name, names = PATCHES[leaf]
        ours = importlib.import_module(ours_name)
        full = f"{_PKG}.{leaf}"
        try:
            ref = importlib.import_module(full)
        except Exception:   # noqa: BLE001 -- the reference module's own dependencies are missing here
            sys.modules.pop(full, None)
            _set_module(full, ours)
            _set_attr(pkg, leaf, ours)
            report[leaf] = "replaced"
            continue
        for name in names:
            if not hasattr(ours, name):
                continue
            new = getattr(ours, name)
            if name == "regionprops_image" and hasattr(ref, name):
                new = _with_fallback(new, getattr(ref, name))
            _set_attr(ref, name, new)
        report[leaf] = "patched"
    # topology.py:12 binds compute_dmt_graph by name: rebind it there too when that module is already loaded
    topo = sys.modules.get(f"{_PKG}.topology")
    if topo is not None and hasattr(topo, "compute_dmt_graph") and "dmtgraph" in report:
        _set_attr(topo, "compute_dmt_graph", importlib.import_module("tmat_b200.dmtgraph").compute_dmt_graph)
    _installed = True
    return report


def uninstall():
    global _installed
    while _undo:
        kind, target, name, old = _undo.pop()
        if kind == "attr":
            if old is _undo:
                try:
                    delattr(target, name)
                except AttributeError:
                    pass
            else:
                setattr(target, name, old)
        elif kind == "module":
            if old is _undo:
                sys.modules.pop(name, None)
            else:
                sys.modules[name] = old
        elif kind == "path":
            if name in sys.path:
                sys.path.remove(name)
    _installed = False
