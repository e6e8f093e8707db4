"""Probe which UMMA-descriptor variant of the haloed-tile 3x3 kernel is correct on this GPU (run on a B200).
usage: halo_probe.py MODE -> prints 'mode M ok' / 'mode M FAIL ...' and exits 0 / 1"""
import ctypes as C
import sys
sys.path.insert(0, ".")
from tmat_b200 import _ffi

mode = int(sys.argv[1])
ctx = _ffi.default_context()
L = _ffi.lib()
L.tmat_debug_set_halo_mode(mode)
CASES = [
    (2, 160, 160, 64, 0, 64, 9, 0, 0, 0, 2), (2, 80, 80, 128, 0, 128, 9, 0, 0, 1, 2),
    (1, 160, 160, 64, 0, 4, 9, 3, 0, 0, 0), (2, 47, 31, 128, 0, 64, 9, 0, 1, 0, 0),
    (3, 32, 32, 256, 0, 128, 9, 0, 1, 0, 0), (1, 16, 8, 64, 0, 64, 9, 0, 0, 0, 0),
]
ok = True
for case in CASES:
    d, m = C.c_double(), C.c_double()
    r = L.tmat_debug_conv_selftest(ctx.handle, *case, 7, C.byref(d), C.byref(m))
    good = (r == 0) and m.value > 0.1 and d.value <= m.value * 2 ** -7
    print(f"mode {mode} case {case}: rc={r} diff={d.value:.4g} ref={m.value:.4g} {'ok' if good else 'FAIL'}", flush=True)
    if r != 0:
        print(L.tmat_last_error())
    ok = ok and good
print(f"mode {mode} {'ok' if ok else 'FAIL'}")
sys.exit(0 if ok else 1)
