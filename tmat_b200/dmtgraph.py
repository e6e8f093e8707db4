"""Discrete-Morse-theory graph: drop-in for ``fl_tissue_model_tools/dmtgraph.py::compute_dmt_graph`` (:38-99),
computed by the native (host C++) implementation inside libtmat_b200 -- the algorithm is a sequential union-find sweep
in filtration order, so there is no kernel; the result has the reference's content and ordering."""
import ctypes as C

import numpy as np

from . import _ffi


def compute_dmt_graph(img, delta1, delta2=0.0):
    """(vertices int32 (n, 2) as (row, column), edges int32 (m, 2)) of the DMT graph of ``img``."""
    a = np.ascontiguousarray(img, dtype=np.float32)
    if a.ndim != 2:
        raise ValueError(f"expected a 2-D image, got shape {a.shape}")
    L = _ffi.lib()
    nv, ne = C.c_int64(), C.c_int64()
    h, w = a.shape
    _ffi.check(L.tmat_dmt_graph_host(_ffi.ptr(a), h, w, float(delta1), float(delta2), None, 0, C.byref(nv), None, 0,
                                     C.byref(ne)))
    verts = np.empty((nv.value, 2), np.int32)
    edges = np.empty((ne.value, 2), np.int32)
    if nv.value or ne.value:
        _ffi.check(L.tmat_dmt_graph_host(_ffi.ptr(a), h, w, float(delta1), float(delta2), _ffi.ptr(verts), nv.value,
                                         C.byref(nv), _ffi.ptr(edges), ne.value, C.byref(ne)))
    return verts, edges
