"""CPU: the C-ABI library builds, loads and exports every symbol include/tmat_b200.h declares;
host-only entry points (geometry, spline window) match the golden vectors; compute entry points
fail loudly without a GPU (no CPU fallback)."""
import os
import re

import numpy as np
import pytest

from tmat_b200 import _ffi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared(header="tmat_b200.h"):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(tmat_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported_and_bound():
    L = _ffi.lib()
    names = _declared()
    assert len(names) >= 40
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/tmat_b200.h but not exported"
    bound = set(_ffi.SIGNATURES) | set(_ffi._SPECIAL)
    assert set(names) == bound, set(names) ^ bound
    assert L.tmat_version() >= 100
    # the test hooks live in their own header and are not part of the boundary
    dbg = _declared("tmat_b200_debug.h")
    assert all("debug" in n for n in dbg) and set(dbg) == set(_ffi.DEBUG_SIGNATURES), dbg
    assert not any("debug" in n for n in names)
    for n in dbg:
        assert hasattr(L, n), n


def test_skeleton_lut_is_one_shared_constant():
    """csrc/skeleton_lut.inc is compiled into the library and parsed by the oracle: one table, two readers."""
    from oracle import morphology as om
    got = np.zeros(256, np.uint8)
    _ffi.check(_ffi.lib().tmat_skeleton_lut_get(_ffi.ptr(got)))
    assert np.array_equal(got, om.ZS_LUT)
    alt = got.copy()
    alt[3] = 1
    _ffi.check(_ffi.lib().tmat_skeleton_lut_set(_ffi.ptr(alt)))
    back = np.zeros(256, np.uint8)
    _ffi.check(_ffi.lib().tmat_skeleton_lut_get(_ffi.ptr(back)))
    assert np.array_equal(back, alt)
    _ffi.check(_ffi.lib().tmat_skeleton_lut_set(_ffi.ptr(got)))
    bad = got.copy()
    bad[7] = 9
    with pytest.raises(_ffi.TmatError):
        _ffi.check(_ffi.lib().tmat_skeleton_lut_set(_ffi.ptr(bad)))


def test_spline_window_matches_reference_golden(golden):
    g = golden("stp")
    from tmat_b200.smooth_tiled_predictions import _spline_window
    for ws in (32, 256, 320):
        out = np.empty(ws, dtype=np.float64)
        _ffi.check(_ffi.lib().tmat_spline_window(ws, _ffi.ptr(out)))
        assert np.array_equal(out, g[f"window1d_{ws}"])
        assert np.array_equal(_spline_window(ws), g[f"window1d_{ws}"])


def test_geometry_matches_reference_tiling():
    from tmat_b200.smooth_tiled_predictions import geometry
    g = geometry(1024, 1024, 320, 2)
    assert (g["aug"], g["PM"], g["step"], g["tiles_even"]) == (160, 1344, 160, (7, 7))
    g = geometry(2048, 2048, 320, 2)
    assert g["tiles_even"] == (13, 13)
    g = geometry(8192, 8192, 256, 2)
    assert g["tiles_even"] == (65, 65)
    g = geometry(100, 72, 32, 2)
    assert g["tiles_even"] == (7, 5) and g["tiles_odd"] == (5, 7)
    with pytest.raises(_ffi.TmatError):
        geometry(10, 10, 32, 64)


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(_ffi.TmatError) as e:
        _ffi.Context(0)
    assert e.value.code == _ffi.ERR_NO_DEVICE
    from tmat_b200 import zstacks
    with pytest.raises(_ffi.TmatError):
        zstacks.proj_max(np.zeros((2, 4, 4), np.uint16))
