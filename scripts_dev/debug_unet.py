"""Bring-up script (GPU): conv self-tests and layer-by-layer comparison with the oracle."""
import ctypes as C
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
from tmat_b200 import _ffi, models
from oracle import unet as ou

ctx = _ffi.default_context()
L = _ffi.lib()

def selftest(B, H, W, C0, C1, Cout, ntaps, epi, relu=0, dual=0, seed=1):
    d, m = C.c_double(), C.c_double()
    t = time.time()
    try:
        _ffi.check(L.tmat_debug_conv_selftest(ctx.handle, B, H, W, C0, C1, Cout, ntaps, epi, relu, dual, seed,
                                              C.byref(d), C.byref(m)))
        print(f"selftest B={B} H={H} W={W} C0={C0} C1={C1} Cout={Cout} taps={ntaps} epi={epi} relu={relu} dual={dual}: "
              f"maxdiff={d.value:.4g} maxref={m.value:.4g}  ({time.time()-t:.2f}s)", flush=True)
    except Exception as e:
        print("selftest FAILED", (B, H, W, C0, C1, Cout, ntaps, epi), e, flush=True)
        raise

which = sys.argv[1] if len(sys.argv) > 1 else "all"
if which in ("all", "conv"):
    selftest(1, 8, 16, 64, 0, 64, 1, 0)          # exactly one tile, one k-step
    selftest(1, 8, 16, 64, 0, 64, 9, 0)
    selftest(2, 16, 16, 128, 0, 128, 9, 0, relu=1)
    selftest(2, 20, 20, 512, 0, 512, 9, 0, relu=1)
    selftest(3, 20, 20, 512, 512, 512, 9, 1, dual=1)
    selftest(2, 40, 40, 256, 512, 256, 9, 1, dual=1)
    selftest(2, 80, 80, 64, 0, 128, 1, 0, relu=1)
    selftest(2, 40, 40, 64, 0, 128, 1, 2, dual=1)
    selftest(1, 24, 24, 64, 0, 4, 9, 3)
    selftest(1, 24, 24, 64, 0, 4, 9, 3, relu=1)
    selftest(5, 13, 27, 64, 64, 64, 9, 0)

def compare(name, a, b):
    d = np.abs(a - b)
    print(f"  {name:14s} shape={a.shape} max|ref|={np.abs(b).max():.4g} maxabs={d.max():.4g} mean={d.mean():.4g}", flush=True)

def net_test(fc, ws, B, impl, affine=True):
    print(f"net fc={fc} ws={ws} B={B} impl={impl}", flush=True)
    w = ou.make_weights(fc, seed=3, randomize_affine=affine)
    x = np.random.default_rng(0).random((B, ws, ws), dtype=np.float32)
    taps = {}
    ref = ou.forward(w, x, return_logits=True, taps=taps)
    net = models.CudaUNet(w, ws, fc, impl=impl)
    t = time.time()
    out = net.predict(x, logits=True)
    print(f"  forward {time.time()-t:.3f}s")
    for name in ["stem", "down0.r1", "down0.b2", "down0", "down1", "down2", "up0.r1", "up0", "up1", "up2"]:
        got = net.debug_tap(name).reshape(taps[name].shape)
        compare(name, got, taps[name])
    z3 = net.debug_tap("up3.lowres")
    r = taps["up3"][:, ::2, ::2, :]
    compare("up3.lowres", z3.reshape(r.shape), r)
    compare("logits", out, ref)
    p = net.predict(x)
    compare("probs", p, ou.forward(w, x))
    return net

if which in ("all", "simt"):
    net_test((8, 16, 32, 64), 64, 2, "simt")
if which in ("all", "tc"):
    net_test((64, 128, 256, 512), 64, 3, "simt")
    net_test((64, 128, 256, 512), 64, 3, "tc")
    net_test((64, 128, 256, 512), 320, 2, "tc", affine=False)
print("done")
