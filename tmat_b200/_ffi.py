"""ctypes binding of libtmat_b200.so (include/tmat_b200.h).

The library is the product; there is no Python or CPU fallback.  If the shared object is
missing it is built in-tree (nvcc); if no sm_100 device is present every compute call raises
``TmatError``.
"""
import ctypes as C
import os
import threading
import weakref

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libtmat_b200.so")

OK, ERR_INVALID, ERR_CUDA, ERR_UNSUPPORTED, ERR_NOMEM, ERR_NO_DEVICE = 0, -1, -2, -3, -4, -5
U8, U16, F32, F64 = 0, 1, 2, 3
DTYPES = {np.dtype(np.uint8): U8, np.dtype(np.uint16): U16, np.dtype(np.float32): F32,
          np.dtype(np.float64): F64}
ZPROJ = {"min": 0, "max": 1, "avg": 2, "med": 3, "fs": 4}


class TmatError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libtmat_b200 error {code}: {msg}")
        self.code = code


class UnetDesc(C.Structure):
    _fields_ = [("patch", C.c_int), ("channels", C.c_int), ("filter_counts", C.c_int * 4),
                ("impl", C.c_int)]


_lib = None
_lib_lock = threading.Lock()

_vp, _i, _i64 = C.c_void_p, C.c_int, C.c_int64
_pi = C.POINTER(C.c_int)

# name -> argtypes (restype is always int unless listed in _SPECIAL)
SIGNATURES = {
    "tmat_ctx_create": [_i, C.POINTER(_vp)],
    "tmat_ctx_destroy": [_vp],
    "tmat_ctx_sync": [_vp],
    "tmat_ctx_launch_count": [_vp, C.POINTER(_i64)],
    "tmat_host_alloc": [_i64, C.POINTER(_vp)],
    "tmat_host_free": [_vp],
    "tmat_ctx_profile_enable": [_vp, _i],
    "tmat_ctx_profile_collect": [_vp, C.POINTER(C.c_double)],
    "tmat_zproj_dev": [_vp, _vp, _i, _i, _i, _i, _i, _i, _vp, _vp],
    "tmat_zproj_host": [_vp, _vp, _i, _i, _i, _i, _i, _i, _vp],
    "tmat_stp_geometry": [_i, _i, _i, _i, _pi],
    "tmat_stp_gather_dev": [_vp, _vp, _i, _i, _i, _i, _i, _vp, _vp],
    "tmat_stp_blend_dev": [_vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp],
    "tmat_stp_gather_host": [_vp, _vp, _i, _i, _i, _i, _i, _vp],
    "tmat_stp_blend_host": [_vp, _vp, _i, _i, _i, _i, _vp, _vp],
    "tmat_spline_window": [_i, _vp],
    "tmat_unet_load": [_vp, C.POINTER(UnetDesc), _i, C.POINTER(C.c_char_p), C.POINTER(_vp),
                       C.POINTER(_i64), C.POINTER(_vp)],
    "tmat_unet_free": [_vp],
    "tmat_unet_forward_dev": [_vp, _vp, _vp, _i, _vp, _i, _vp],
    "tmat_unet_forward_host": [_vp, _vp, _vp, _i, _vp, _i],
    "tmat_stp_predict_dev": [_vp, _vp, _vp, _i, _i, _i, _vp, _vp],
    "tmat_stp_predict_host": [_vp, _vp, _vp, _i, _i, _i, _vp],
    "tmat_stp_predict_rows_dev": [_vp, _vp, _vp, _i, _i, _i, _i, _i, _vp, C.POINTER(_i), _vp],
    "tmat_stp_predict_rows_host": [_vp, _vp, _vp, _i, _i, _i, _i, _i, _vp, C.POINTER(_i)],
    "tmat_stp_rows_patches": [_i, _i, _i, _i, _i, _i, _vp, _i, C.POINTER(_i)],
    "tmat_unet_flops_per_patch": [C.POINTER(UnetDesc), C.POINTER(C.c_double)],
    "tmat_gmm2_threshold_dev": [_vp, _vp, _i64, C.c_double, _vp, _vp, _vp],
    "tmat_gmm2_threshold_host": [_vp, _vp, _i64, C.c_double, _vp, _vp],
    "tmat_threshold_le_zero_f32_host": [_vp, _vp, _i64, C.c_double, _vp],
    "tmat_blur_dev": [_vp, _vp, _i, _i, _i, _i, _i, _i, _vp, _vp],
    "tmat_blur_host": [_vp, _vp, _i, _i, _i, _i, _i, _i, _vp],
    "tmat_threshold_gt_f64_dev": [_vp, _vp, _i64, C.c_double, _vp, _vp],
    "tmat_apply_mask_f32_dev": [_vp, _vp, _vp, _i64, _vp, _vp],
    "tmat_gaussian_dev": [_vp, _vp, _i, _i, _i, C.c_double, _vp, _vp],
    "tmat_gaussian_host": [_vp, _vp, _i, _i, _i, C.c_double, _vp],
    "tmat_auto_threshold_well_dev": [_vp, _vp, _i, _i, _i, _vp, _pi, _vp],
    "tmat_auto_threshold_well_host": [_vp, _vp, _i, _i, _i, _vp, _pi],
    "tmat_median_disk2_dev": [_vp, _vp, _i, _i, _vp, _vp],
    "tmat_erode_disk_dev": [_vp, _vp, _i, _i, _i, _i, _vp, _vp],
    "tmat_label_dev": [_vp, _vp, _i, _i, _i, _vp, _pi, _vp],
    "tmat_region_stats_dev": [_vp, _vp, _i, _i, _i, _vp, _vp, _vp],
    "tmat_skeletonize_dev": [_vp, _vp, _i, _i, _vp, _pi, _vp],
    "tmat_filter_branch_seg_mask_dev": [_vp, _vp, _i, _i, _i, _i, _vp, _vp],
    "tmat_filter_branch_seg_mask_host": [_vp, _vp, _i, _i, _i, _i, _vp],
    "tmat_remove_small_islands_host": [_vp, _vp, _i, _i, _i, _i, _i, _i, _vp],
    "tmat_area_image_host": [_vp, _vp, _i, _i, _vp],
    "tmat_region_prop_image_host": [_vp, _vp, _i, _i, _i, _vp],
    "tmat_footprint_filter_dev": [_vp, _vp, _i, _i, _vp, _i, _i, _i, _i, _vp, _vp],
    "tmat_footprint_filter_host": [_vp, _vp, _i, _i, _vp, _i, _i, _i, _i, _vp],
    "tmat_skeleton_lut_set": [_vp],
    "tmat_skeleton_lut_get": [_vp],
    "tmat_median_disk2_host": [_vp, _vp, _i, _i, _vp],
    "tmat_erode_disk_host": [_vp, _vp, _i, _i, _i, _i, _vp],
    "tmat_label_host": [_vp, _vp, _i, _i, _i, _vp, _pi],
    "tmat_skeletonize_host": [_vp, _vp, _i, _i, _vp, _pi],
    "tmat_superellipse_hull_host": [_vp, _vp, _vp, _i, _i, _vp, _vp, _vp, _i, C.c_double, C.c_double, _pi],
    "tmat_superellipse_mask_host": [_vp, _vp, _i, _vp, _i, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double,
                                    C.c_double, _i, _vp],
    "tmat_convex_hull_mask_host": [_vp, _i, _i, _vp, _i, _vp],
    "tmat_distance_transform_edt_dev": [_vp, _vp, _i, _i, _i, _vp, _vp],
    "tmat_distance_transform_edt_host": [_vp, _vp, _i, _i, _i, _vp],
    "tmat_distance_transform_edt_static_dev": [_vp, _i, _i, _i, _vp, _vp],
    "tmat_centerline_weight_dev": [_vp, _vp, _vp, _vp, _i, _i, _vp, _vp],
    "tmat_centerline_weight_host": [_vp, _vp, _vp, _vp, _i, _i, _vp],
    "tmat_distance_transform_cv5_dev": [_vp, _vp, _i, _i, _i, _vp, _vp],
    "tmat_distance_transform_cv5_host": [_vp, _vp, _i, _i, _i, _vp],
    "tmat_circle_mask_host": [_vp, _i, _i, _i, _i, _i, _i, _vp],
    "tmat_dmt_graph_host": [_vp, _i, _i, C.c_float, C.c_float, _vp, _i64, C.POINTER(_i64), _vp, _i64, C.POINTER(_i64)],
    "tmat_resize_dev": [_vp, _vp, _i, _i, _i, _vp, _i, _i, _i, _vp],
    "tmat_resize_host": [_vp, _vp, _i, _i, _i, _vp, _i, _i, _i],
    "tmat_branch_mask_dev": [_vp, _vp, _vp, _i, _i, _i, _i, _vp, _vp, _vp],
}
# include/tmat_b200_debug.h: test hooks, not part of the drop-in boundary
DEBUG_SIGNATURES = {
    "tmat_unet_debug_tap": [_vp, _vp, C.c_char_p, _vp, _i64, C.POINTER(_i64)],
    "tmat_debug_conv_selftest": [_vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, C.c_uint64,
                                 C.POINTER(C.c_double), C.POINTER(C.c_double)],
    "tmat_debug_set_halo_mode": [_i],
    "tmat_debug_dwpw_selftest": [_vp, _i, _i, _i, _i, _i, _i, C.c_uint64, C.POINTER(C.c_double),
                                 C.POINTER(C.c_double)],
    "tmat_debug_dwpw_pool_selftest": [_vp, _i, _i, _i, _i, _i, _i, C.c_uint64, C.POINTER(C.c_double),
                                      C.POINTER(C.c_double)],
}
_SPECIAL = {"tmat_version": ([], _i), "tmat_last_error": ([], C.c_char_p)}


def lib():
    """Load (building if necessary) the shared library and declare every prototype."""
    global _lib
    with _lib_lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                from . import build as _build
                _build.build()
            L = C.CDLL(LIB_PATH)
            for name, args in list(SIGNATURES.items()) + list(DEBUG_SIGNATURES.items()):
                f = getattr(L, name)
                f.argtypes = args
                f.restype = _i
            for name, (args, res) in _SPECIAL.items():
                f = getattr(L, name)
                f.argtypes = args
                f.restype = res
            _lib = L
    return _lib


def check(code):
    if code != OK:
        raise TmatError(code, lib().tmat_last_error().decode("utf-8", "replace"))


class Context:
    """One libtmat_b200 context (device scratch + launch counter).  Not thread-safe: use
    ``default_context()`` for a per-thread instance."""

    def __init__(self, device=0):
        self.handle = _vp()
        check(lib().tmat_ctx_create(device, C.byref(self.handle)))
        self.device = device

    def close(self):
        if self.handle:
            lib().tmat_ctx_destroy(self.handle)
            self.handle = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def sync(self):
        check(lib().tmat_ctx_sync(self.handle))

    def profile(self, on):
        check(lib().tmat_ctx_profile_enable(self.handle, int(on)))

    def profile_collect(self):
        """(kernel ms, launches, executed FLOPs) of the tcgen05 conv launches since the last call."""
        out = (C.c_double * 3)()
        check(lib().tmat_ctx_profile_collect(self.handle, out))
        return out[0], int(out[1]), out[2]

    def launch_count(self):
        n = _i64()
        check(lib().tmat_ctx_launch_count(self.handle, C.byref(n)))
        return n.value


_tls = threading.local()


def default_context():
    ctx = getattr(_tls, "ctx", None)
    if ctx is None:
        dev = int(os.environ.get("TMAT_DEVICE", os.environ.get("LOCAL_RANK", "0")))
        ctx = Context(dev)
        _tls.ctx = ctx
    return ctx


# ---- page-locked result buffers ---------------------------------------------------------------------------
# Results handed back to the caller are numpy arrays over cudaHostAlloc memory, so the device->host copy of a
# result runs at PCIe speed.  Buffers return to a small per-size pool when the array (and every view of it) is
# garbage-collected; beyond _PINNED_LIVE_CAP live bytes the result is an ordinary pageable array.
_PINNED_LIVE_CAP = 1 << 30
_pinned_lock = threading.Lock()
_pinned_pool = {}        # nbytes -> [address, ...]
_pinned_live = 0


def _pinned_release(addr, nbytes):
    global _pinned_live
    with _pinned_lock:
        _pinned_live -= nbytes
        free = _pinned_pool.setdefault(nbytes, [])
        if len(free) < 2:
            free.append(addr)
            return
    try:
        lib().tmat_host_free(_vp(addr))
    except Exception:
        pass


def pinned_empty(shape, dtype):
    """np.empty(shape, dtype) over page-locked memory (pageable when the live cap is reached)."""
    global _pinned_live
    dtype = np.dtype(dtype)
    nbytes = int(np.prod(shape)) * dtype.itemsize
    if nbytes == 0:
        return np.empty(shape, dtype)
    addr = None
    with _pinned_lock:
        free = _pinned_pool.get(nbytes)
        if free:
            addr = free.pop()
        elif _pinned_live + nbytes > _PINNED_LIVE_CAP:
            return np.empty(shape, dtype)
        _pinned_live += nbytes
    if addr is None:
        p = _vp()
        try:
            check(lib().tmat_host_alloc(nbytes, C.byref(p)))
        except Exception:
            with _pinned_lock:
                _pinned_live -= nbytes
            raise
        addr = p.value
        # a caller that rebinds `res = f(...)` in a loop holds the previous result while the next one is
        # produced: keep one spare of a size the first time it is seen, so the second call does not page-lock
        spare = _vp()
        with _pinned_lock:
            want_spare = not _pinned_pool.get(nbytes) and _pinned_live + nbytes <= _PINNED_LIVE_CAP
        if want_spare and lib().tmat_host_alloc(nbytes, C.byref(spare)) == 0:
            with _pinned_lock:
                _pinned_pool.setdefault(nbytes, []).append(spare.value)
    buf = (C.c_char * nbytes).from_address(addr)
    weakref.finalize(buf, _pinned_release, addr, nbytes)
    return np.frombuffer(buf, dtype=dtype).reshape(shape)


def ptr(a):
    """void* of a C-contiguous numpy array, or of anything exposing data_ptr() (torch)."""
    if a is None:
        return None
    if hasattr(a, "data_ptr"):
        return _vp(a.data_ptr())
    assert a.flags["C_CONTIGUOUS"]
    return _vp(a.ctypes.data)


def as_c(a, dtype=None):
    a = np.asarray(a)
    if dtype is not None and a.dtype != dtype:
        a = a.astype(dtype)
    return np.ascontiguousarray(a)
