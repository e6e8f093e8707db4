"""GPU: Z projections through the C ABI vs the oracle / golden fixtures, bit-exact."""
import numpy as np
import pytest

import helpers
from oracle import zstacks as oz

pytestmark = pytest.mark.gpu

METHODS = ("min", "max", "avg", "med", "fs")


def _fn(m):
    from tmat_b200 import zstacks as z
    return {"min": z.proj_min, "max": z.proj_max, "avg": z.proj_avg, "med": z.proj_med,
            "fs": z.proj_focus_stacking}[m]


def test_zproj_golden_bitexact(golden):
    g = golden("zproj")
    for name in ("u16_odd", "u16_even", "u8_even"):
        st = g[f"{name}_in"]
        for m in METHODS:
            out = _fn(m)(st)
            ref = g[f"{name}_{m}"]
            assert out.dtype == ref.dtype, (name, m)
            assert np.array_equal(out, ref), (name, m, int((out != ref).sum()))


def test_zproj_float32_golden_bitexact(golden):
    """float32 stacks through min/max/avg/med (numpy keeps float32); focus stacking of a float stack raises."""
    from tmat_b200 import zstacks as z
    g = golden("zproj")
    for name in ("f32_odd", "f32_even", "f32_z32"):
        st = g[f"{name}_in"]
        for m in ("min", "max", "avg", "med"):
            out = _fn(m)(st)
            ref = g[f"{name}_{m}"]
            assert out.dtype == ref.dtype == np.float32, (name, m)
            assert np.array_equal(out, ref), (name, m, int((out != ref).sum()))
    with pytest.raises(TypeError):
        z.proj_focus_stacking(g["f32_odd_in"])
    st = np.random.default_rng(1).random((70, 9, 13), dtype=np.float32)       # Z > 64: rank-selection kernel
    for m in ("min", "max", "avg", "med"):
        assert np.array_equal(_fn(m)(st), oz.METHODS[m](st)), m


@pytest.mark.parametrize("shape,dtype", [((1, 5, 7), np.uint16), ((2, 33, 65), np.uint16),
                                         ((3, 64, 64), np.uint8), ((5, 70, 131), np.uint16),
                                         ((16, 96, 200), np.uint16), ((31, 40, 48), np.uint8),
                                         ((33, 24, 40), np.uint16), ((70, 16, 24), np.uint16),
                                         # the warp-tile focus kernel: several 124-column tiles, ragged last tile,
                                         # heights off the 32-row grid, one slice, image smaller than one tile
                                         ((9, 70, 252), np.uint16), ((4, 40, 128), np.uint16), ((6, 100, 500), np.uint16),
                                         ((1, 33, 124), np.uint16), ((3, 8, 8), np.uint16), ((12, 65, 376), np.uint16),
                                         ((257, 9, 12), np.uint16)])
def test_zproj_ragged_shapes_vs_oracle(shape, dtype):
    rng = np.random.default_rng(sum(shape))
    top = np.iinfo(dtype).max
    st = rng.integers(0, top + 1, size=shape, dtype=dtype)
    st[:, :3, :3] = top                   # saturated + tied block
    st[:, -2:, -2:] = 0
    for m in METHODS:
        out = _fn(m)(st)
        ref = oz.METHODS[m](st)
        assert out.dtype == ref.dtype and np.array_equal(out, ref), (shape, m)


def test_zproj_axis_argument():
    st = helpers.synth_stack(6, 40, 44, 5)
    moved = np.ascontiguousarray(np.moveaxis(st, 0, 2))
    assert np.array_equal(_fn("max")(moved, axis=2), st.max(0))
    assert np.array_equal(_fn("med")(moved, axis=2), np.median(st, axis=0))


def test_zproj_full_size_properties():
    """BASELINE config C2 (2048x2048x32 uint16): size-independent properties."""
    rng = np.random.default_rng(2)
    st = rng.integers(0, 65536, size=(32, 2048, 2048), dtype=np.uint16)
    mx, mn, av, md, fs = (_fn(m)(st) for m in ("max", "min", "avg", "med", "fs"))
    assert np.array_equal(mx, st.max(0)) and np.array_equal(mn, st.min(0))
    assert np.array_equal(av, st.mean(0))
    assert (mn <= md).all() and (md <= mx).all()
    sub = np.s_[:, 1000:1064, 500:564]
    assert np.array_equal(md[sub[1:]], np.median(st[sub], axis=0))
    # every focus-stacked pixel is one of the stack's values at that pixel
    assert (st == fs[None]).any(0).all()
    crop = st[:, :96, :96]
    assert np.array_equal(fs[:80, :80], oz.proj_focus_stacking(crop)[:80, :80])
