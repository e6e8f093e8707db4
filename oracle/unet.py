"""Oracle (test infrastructure): UNet-Xception forward pass in torch-CPU fp32.

Restates ``models.py:110-171`` (``build_UNetXception``) with Keras 2.14 / TensorFlow
layer semantics written out by hand (TensorFlow is not installable in the build
container: **parity unpinned** at the TF boundary):

* Conv2D / SeparableConv2D / Conv2DTranspose all ``use_bias=True``.
* TF ``SAME`` padding: stride-2 3x3 conv and MaxPool(3, 2) on even sizes pad 0
  before and 1 after; pool padding never wins the max; 1x1 stride-2 samples even
  indices.
* ``Conv2DTranspose(3, stride 1, same)``:
  ``out[q, co] = b[co] + sum in[q + k - 1, ci] * K[2-ky, 2-kx, co, ci]`` (kernel stored
  ``(kh, kw, out, in)``).
* BatchNormalization in inference mode, ``epsilon = 1e-3``.
* ``UpSampling2D(2)`` nearest; sigmoid head.

Weights are a flat ``dict[str, np.ndarray]`` in Keras array layouts; ``make_weights``
is the seeded glorot-uniform generator handed identically to this oracle and to
the CUDA engine (the trained checkpoint is not in the reference tree,
``.MISSING_LARGE_BLOBS:6``).
"""
import numpy as np
import torch
import torch.nn.functional as F

BN_EPS = 1e-3


def _glorot(rng, shape, fan_in, fan_out, gain=1.0):
    lim = gain * np.sqrt(6.0 / (fan_in + fan_out))
    return rng.uniform(-lim, lim, size=shape).astype(np.float32)


def _bn(rng, c, randomize):
    if not randomize:
        return {"gamma": np.ones(c, np.float32), "beta": np.zeros(c, np.float32),
                "mean": np.zeros(c, np.float32), "var": np.ones(c, np.float32)}
    return {"gamma": rng.uniform(0.7, 1.3, c).astype(np.float32),
            "beta": rng.uniform(-0.2, 0.2, c).astype(np.float32),
            "mean": rng.uniform(-0.2, 0.2, c).astype(np.float32),
            "var": rng.uniform(0.6, 1.5, c).astype(np.float32)}


def make_weights(filter_counts=(64, 128, 256, 512), channels=1, seed=0,
                 randomize_affine=False, head_gain=1.0):
    """Keras-default initialisation (glorot-uniform kernels, zero bias, identity BN)
    when ``randomize_affine`` is False; otherwise biases and BN statistics are drawn
    too, so that bias/BN folding in the CUDA engine is actually exercised."""
    rng = np.random.default_rng(seed)
    fc = sorted(filter_counts)
    w = {}

    def bias(c):
        return (rng.uniform(-0.1, 0.1, c).astype(np.float32) if randomize_affine
                else np.zeros(c, np.float32))

    def put_bn(prefix, c):
        for k, v in _bn(rng, c, randomize_affine).items():
            w[f"{prefix}.{k}"] = v

    w["stem.kernel"] = _glorot(rng, (3, 3, channels, fc[0]), 9 * channels, 9 * fc[0])
    w["stem.bias"] = bias(fc[0])
    put_bn("stem.bn", fc[0])
    cin = fc[0]
    for i, f in enumerate(fc[1:]):
        p = f"down{i}"
        for s, ci in (("sep1", cin), ("sep2", f)):
            w[f"{p}.{s}.dw"] = _glorot(rng, (3, 3, ci, 1), 9 * ci, 9)
            w[f"{p}.{s}.pw"] = _glorot(rng, (1, 1, ci, f), ci, f)
            w[f"{p}.{s}.bias"] = bias(f)
            put_bn(f"{p}.bn{s[-1]}", f)
        w[f"{p}.res.kernel"] = _glorot(rng, (1, 1, cin, f), cin, f)
        w[f"{p}.res.bias"] = bias(f)
        cin = f
    for i, f in enumerate(reversed(fc)):
        p = f"up{i}"
        for s, ci in (("ct1", cin), ("ct2", f)):
            # Conv2DTranspose kernel layout (kh, kw, out, in)
            w[f"{p}.{s}.kernel"] = _glorot(rng, (3, 3, f, ci), 9 * f, 9 * ci)
            w[f"{p}.{s}.bias"] = bias(f)
            put_bn(f"{p}.bn{s[-1]}", f)
        w[f"{p}.res.kernel"] = _glorot(rng, (1, 1, cin, f), cin, f)
        w[f"{p}.res.bias"] = bias(f)
        cin = f
    w["head.kernel"] = _glorot(rng, (3, 3, cin, 1), 9 * cin, 9, gain=head_gain)
    w["head.bias"] = bias(1)
    return w


def _t(a):
    return torch.from_numpy(np.ascontiguousarray(a))


def _conv(x, k, b, stride=1):
    """Keras Conv2D, padding='same'.  k: (kh, kw, in, out).  x: NCHW."""
    kh = k.shape[0]
    wt = _t(k).permute(3, 2, 0, 1).contiguous()
    h = x.shape[-1]
    out = -(-h // stride)
    pad = max((out - 1) * stride + kh - h, 0)
    lo, hi = pad // 2, pad - pad // 2
    x = F.pad(x, (lo, hi, lo, hi))
    return F.conv2d(x, wt, _t(b), stride=stride)


def _sepconv(x, dw, pw, b):
    """Keras SeparableConv2D(3, same): depthwise (no bias) then pointwise + bias."""
    c = x.shape[1]
    wd = _t(dw).permute(2, 3, 0, 1).contiguous()  # (in, 1, kh, kw)
    x = F.conv2d(F.pad(x, (1, 1, 1, 1)), wd, None, groups=c)
    return F.conv2d(x, _t(pw).permute(3, 2, 0, 1).contiguous(), _t(b))


def _convT(x, k, b):
    """Keras Conv2DTranspose(3, stride 1, same).  k: (kh, kw, out, in)."""
    wt = _t(k).permute(3, 2, 0, 1).contiguous()  # torch wants (in, out, kh, kw)
    return F.conv_transpose2d(x, wt, _t(b), stride=1, padding=1)


def _bnorm(x, w, p):
    s = _t(w[p + ".gamma"]) / torch.sqrt(_t(w[p + ".var"]) + BN_EPS)
    return (x - _t(w[p + ".mean"])[None, :, None, None]) * s[None, :, None, None] \
        + _t(w[p + ".beta"])[None, :, None, None]


def _maxpool_same(x):
    return F.max_pool2d(F.pad(x, (0, 1, 0, 1), value=float("-inf")), 3, stride=2)


def _up2(x):
    return x.repeat_interleave(2, dim=2).repeat_interleave(2, dim=3)


@torch.no_grad()
def forward(w, patches, return_logits=False, taps=None):
    """patches: (n, ws, ws) or (n, ws, ws, 1) float32 -> (n, ws, ws, 1) float32.
    ``taps``: optional dict filled with intermediate NHWC activations by name."""
    x = np.asarray(patches, dtype=np.float32)
    if x.ndim == 3:
        x = x[..., None]
    x = _t(x).permute(0, 3, 1, 2).contiguous()
    nd = sum(1 for k in w if k.startswith("down") and k.endswith(".res.bias"))
    nu = sum(1 for k in w if k.startswith("up") and k.endswith(".res.bias"))

    def tap(name, t):
        if taps is not None:
            taps[name] = t.permute(0, 2, 3, 1).contiguous().numpy()

    x = F.relu(_bnorm(_conv(x, w["stem.kernel"], w["stem.bias"], 2), w, "stem.bn"))
    tap("stem", x)
    prev = x
    for i in range(nd):                                    # models.py:126-146
        p = f"down{i}"
        if i != 0:
            x = F.relu(x)
        x = _bnorm(_sepconv(x, w[p + ".sep1.dw"], w[p + ".sep1.pw"], w[p + ".sep1.bias"]), w, p + ".bn1")
        x = F.relu(x)
        tap(p + ".r1", x)
        x = _bnorm(_sepconv(x, w[p + ".sep2.dw"], w[p + ".sep2.pw"], w[p + ".sep2.bias"]), w, p + ".bn2")
        tap(p + ".b2", x)
        x = _maxpool_same(x)
        x = x + _conv(prev, w[p + ".res.kernel"], w[p + ".res.bias"], 2)
        tap(p, x)
        prev = x
    for i in range(nu):                                    # models.py:149-163
        p = f"up{i}"
        x = F.relu(x)
        x = _bnorm(_convT(x, w[p + ".ct1.kernel"], w[p + ".ct1.bias"]), w, p + ".bn1")
        x = F.relu(x)
        tap(p + ".r1", x)
        x = _bnorm(_convT(x, w[p + ".ct2.kernel"], w[p + ".ct2.bias"]), w, p + ".bn2")
        x = _up2(x)
        x = x + _conv(_up2(prev), w[p + ".res.kernel"], w[p + ".res.bias"], 1)
        tap(p, x)
        prev = x
    x = _conv(x, w["head.kernel"], w["head.bias"], 1)       # models.py:166
    if not return_logits:
        x = torch.sigmoid(x)
    return x.permute(0, 2, 3, 1).contiguous().numpy()


def make_pred_func(w, threads=None):
    """A Keras-``model.predict``-shaped callable for predict_img_with_smooth_windowing."""
    if threads:
        torch.set_num_threads(threads)

    def pred_func(batch, verbose=0):
        return forward(w, batch)
    return pred_func


def flops_per_patch(ws=320, filter_counts=(64, 128, 256, 512), channels=1):
    """2*MAC of every convolution as written in models.py:110-171 (BN/ReLU/pool/add
    excluded): 31.03 GFLOP for ws=320, 19.86 GFLOP for ws=256."""
    fc = sorted(filter_counts)
    s = ws // 2
    mac = s * s * 9 * channels * fc[0]
    cin = fc[0]
    for f in fc[1:]:
        mac += s * s * (9 * cin + cin * f) + s * s * (9 * f + f * f)
        s //= 2
        mac += s * s * cin * f
        cin = f
    for f in reversed(fc):
        mac += s * s * 9 * (cin * f + f * f)
        s *= 2
        mac += s * s * cin * f
        cin = f
    mac += s * s * 9 * cin
    return 2 * mac
