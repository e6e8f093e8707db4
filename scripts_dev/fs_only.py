"""Focus stacking of the C2 stack once (ncu target)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from tmat_b200 import _ffi
ctx = _ffi.Context(0); L = _ffi.lib()
S, Z = 2048, 32
g = torch.Generator(device="cuda").manual_seed(1)
st = torch.randint(0, 30000, (Z, S, S), generator=g, device="cuda", dtype=torch.int32).to(torch.int16)
out = torch.empty((S, S), dtype=torch.int16, device="cuda")
method = _ffi.ZPROJ[sys.argv[1] if len(sys.argv) > 1 else "fs"]
if method in (2, 3):
    out = torch.empty((S, S), dtype=torch.float64, device="cuda")
for _ in range(3):
    _ffi.check(L.tmat_zproj_dev(ctx.handle, _ffi.ptr(st), _ffi.U16, Z, S, S, method, 5, _ffi.ptr(out), None))
torch.cuda.synchronize()
