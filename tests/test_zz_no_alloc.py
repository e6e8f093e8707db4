"""GPU: the allocation-free EDT entry point (tmat_distance_transform_edt_static_dev) against the existing one.

No counterpart in the reference: the entry point exists for callers that cannot call cudaMalloc (CUDA graph capture,
serving loops); its scratch plane is a file-scope __device__ array of TMAT_EDT_STATIC_MAX_PIXELS uint32.  Collected
last on purpose.  Bit-exact, like every integer / exact-EDT comparison in this suite.
"""
import numpy as np
import pytest

import helpers

pytestmark = pytest.mark.gpu
MAX_PIXELS = 1 << 22
_warm = []          # the first call on a device may load the module (driver-side, once); later calls must not move memory


def _edt_both(m, invert):
    import torch
    from tmat_b200 import _ffi
    H, W = m.shape
    d_m = torch.from_numpy(np.ascontiguousarray(m, np.uint8)).cuda()
    d_ref = torch.empty((H, W), dtype=torch.float64, device="cuda")
    d_out = torch.full((H, W), -1.0, dtype=torch.float64, device="cuda")
    ctx = _ffi.default_context()
    _ffi.check(_ffi.lib().tmat_distance_transform_edt_dev(ctx.handle, _ffi.ptr(d_m), H, W, int(invert), _ffi.ptr(d_ref), None))
    torch.cuda.synchronize()
    before, free_before = torch.cuda.memory_allocated(), torch.cuda.mem_get_info()[0]
    _ffi.check(_ffi.lib().tmat_distance_transform_edt_static_dev(_ffi.ptr(d_m), H, W, int(invert), _ffi.ptr(d_out), None))
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    if _warm:
        assert torch.cuda.mem_get_info()[0] == free_before      # the device's free memory did not move either
    _warm.append(1)
    return d_out.cpu().numpy(), d_ref.cpu().numpy()


def test_static_edt_matches_arena_edt_two_calls_in_a_row():
    """Two calls back to back on different images and sizes (the second smaller, so the array still holds the first
    call's values beyond its prefix), then the first again: each equals the existing entry point bit for bit."""
    from scipy import ndimage as ndi
    a = helpers.synth_mask(300, 420, 11, density=0.3)
    b = helpers.synth_mask(97, 131, 12, density=0.05)
    for m, invert in ((a, False), (b, True), (a, False), (np.ones((40, 50), bool), False)):
        out, ref = _edt_both(m, invert)
        assert np.array_equal(out, ref)
        if not m.all():
            src = np.logical_not(m) if invert else m
            assert np.array_equal(out, ndi.distance_transform_edt(src))
        else:
            assert np.isinf(out).all()


def test_static_edt_full_size_and_limit():
    import torch
    from tmat_b200 import _ffi
    m = helpers.synth_mask(2048, 2048, 13, density=0.2)          # exactly TMAT_EDT_STATIC_MAX_PIXELS
    assert m.size == MAX_PIXELS
    out, ref = _edt_both(m, False)
    assert np.array_equal(out, ref)
    d_m = torch.zeros((2048, 2049), dtype=torch.uint8, device="cuda")
    d_o = torch.empty((2048, 2049), dtype=torch.float64, device="cuda")
    rc = _ffi.lib().tmat_distance_transform_edt_static_dev(_ffi.ptr(d_m), 2048, 2049, 0, _ffi.ptr(d_o), None)
    assert rc == _ffi.ERR_INVALID
    with pytest.raises(_ffi.TmatError, match="TMAT_EDT_STATIC_MAX_PIXELS"):
        _ffi.check(rc)
    rc = _ffi.lib().tmat_distance_transform_edt_static_dev(None, 8, 8, 0, _ffi.ptr(d_o), None)
    assert rc == _ffi.ERR_INVALID
