"""UNet-Xception patch segmentor on the B200: drop-in for the inference side of the reference's
``fl_tissue_model_tools/models.py`` (``build_UNetXception:85``, ``UNetXceptionPatchSegmentor:597``,
``get_unet_patch_segmentor_from_cfg:656``).  Forward pass only; training utilities are out of scope.

``CudaUNet`` plays the role of the Keras model: ``CudaUNet.predict(batch, verbose=0)`` has the
``model.predict`` contract that ``predict_img_with_smooth_windowing`` expects, and is recognised
there as the fused fast path.
"""
import ctypes as C
import json
import os

import numpy as np

from . import _ffi
from .smooth_tiled_predictions import predict_img_with_smooth_windowing


def weight_names(filter_counts=(64, 128, 256, 512)):
    """Names and Keras-layout shapes of every array of build_UNetXception (models.py:110-171)."""
    fc = sorted(filter_counts)
    out = []

    def bn(prefix, c):
        for k in ("gamma", "beta", "mean", "var"):
            out.append((f"{prefix}.{k}", (c,)))

    out.append(("stem.kernel", (3, 3, 1, fc[0])))
    out.append(("stem.bias", (fc[0],)))
    bn("stem.bn", fc[0])
    cin = fc[0]
    for i, f in enumerate(fc[1:]):
        p = f"down{i}"
        for s, ci in (("sep1", cin), ("sep2", f)):
            out.append((f"{p}.{s}.dw", (3, 3, ci, 1)))
            out.append((f"{p}.{s}.pw", (1, 1, ci, f)))
            out.append((f"{p}.{s}.bias", (f,)))
            bn(f"{p}.bn{s[-1]}", f)
        out.append((f"{p}.res.kernel", (1, 1, cin, f)))
        out.append((f"{p}.res.bias", (f,)))
        cin = f
    for i, f in enumerate(reversed(fc)):
        p = f"up{i}"
        for s, ci in (("ct1", cin), ("ct2", f)):
            out.append((f"{p}.{s}.kernel", (3, 3, f, ci)))     # Conv2DTranspose: (kh, kw, out, in)
            out.append((f"{p}.{s}.bias", (f,)))
            bn(f"{p}.bn{s[-1]}", f)
        out.append((f"{p}.res.kernel", (1, 1, cin, f)))
        out.append((f"{p}.res.bias", (f,)))
        cin = f
    out.append(("head.kernel", (3, 3, cin, 1)))
    out.append(("head.bias", (1,)))
    return out


def random_init_weights(filter_counts=(64, 128, 256, 512), seed=0):
    """Keras default initialisation of a freshly built model: glorot-uniform kernels, zero biases,
    identity BatchNorm statistics (what the reference has before ``load_weights``, models.py:615)."""
    rng = np.random.default_rng(seed)
    w = {}
    for name, shape in weight_names(filter_counts):
        leaf = name.rsplit(".", 1)[1]
        if leaf in ("kernel", "dw", "pw"):
            kh, kw, a, b = shape
            if leaf == "dw":
                fan_in, fan_out = kh * kw * a, kh * kw * b
            elif ".ct" in name:
                fan_in, fan_out = kh * kw * b, kh * kw * a
            else:
                fan_in, fan_out = kh * kw * a, kh * kw * b
            lim = np.sqrt(6.0 / (fan_in + fan_out))
            w[name] = rng.uniform(-lim, lim, size=shape).astype(np.float32)
        elif leaf in ("gamma", "var"):
            w[name] = np.ones(shape, np.float32)
        else:
            w[name] = np.zeros(shape, np.float32)
    return w


class CudaUNet:
    """The compiled UNet living on the GPU.  ``impl``: 'auto' | 'simt' | 'tc'."""

    def __init__(self, weights, patch, filter_counts=(64, 128, 256, 512), channels=1, impl="auto", ctx=None):
        self.ctx = ctx or _ffi.default_context()
        self.patch = int(patch)
        self.filter_counts = tuple(sorted(int(f) for f in filter_counts))
        desc = _ffi.UnetDesc()
        desc.patch = self.patch
        desc.channels = int(channels)
        for i, f in enumerate(self.filter_counts):
            desc.filter_counts[i] = f
        desc.impl = {"auto": 0, "simt": 1, "tc": 2}[impl]
        self.desc = desc
        names = sorted(weights)
        arrays = [np.ascontiguousarray(weights[k], dtype=np.float32) for k in names]
        n = len(names)
        c_names = (C.c_char_p * n)(*[k.encode() for k in names])
        c_ptrs = (C.c_void_p * n)(*[a.ctypes.data for a in arrays])
        c_sizes = (C.c_int64 * n)(*[a.size for a in arrays])
        h = C.c_void_p()
        _ffi.check(_ffi.lib().tmat_unet_load(self.ctx.handle, C.byref(desc), n, c_names, c_ptrs, c_sizes,
                                             C.byref(h)))
        self._tmat_handle = h

    def close(self):
        if getattr(self, "_tmat_handle", None):
            _ffi.lib().tmat_unet_free(self._tmat_handle)
            self._tmat_handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def flops_per_patch(self):
        f = C.c_double()
        _ffi.check(_ffi.lib().tmat_unet_flops_per_patch(C.byref(self.desc), C.byref(f)))
        return f.value

    def predict(self, batch, verbose=0, logits=False):
        """Keras ``model.predict`` contract: (n, ws, ws[, 1]) float32 -> (n, ws, ws, 1) float32."""
        x = _ffi.as_c(batch, np.float32)
        if x.ndim == 4 and x.shape[-1] == 1:
            x = x[..., 0]
        if x.ndim != 3 or x.shape[1] != self.patch or x.shape[2] != self.patch:
            raise ValueError(f"expected (n, {self.patch}, {self.patch}) patches, got {x.shape}")
        x = np.ascontiguousarray(x)
        out = np.empty(x.shape, dtype=np.float32)
        _ffi.check(_ffi.lib().tmat_unet_forward_host(self.ctx.handle, self._tmat_handle, _ffi.ptr(x), x.shape[0],
                                                     _ffi.ptr(out), int(bool(logits))))
        return out[..., None]

    def debug_tap(self, name):
        n = C.c_int64()
        L = _ffi.lib()
        _ffi.check(L.tmat_unet_debug_tap(self.ctx.handle, self._tmat_handle, name.encode(), None, 0, C.byref(n)))
        out = np.empty(n.value, np.float32)
        _ffi.check(L.tmat_unet_debug_tap(self.ctx.handle, self._tmat_handle, name.encode(), _ffi.ptr(out), n.value,
                                         C.byref(n)))
        return out


RANDOM_INIT = "random"    # explicit opt-in for a freshly initialised model (bench, tests)


def load_checkpoint(checkpoint_file, filter_counts):
    """Weights as a name -> array dict.  ``.npz`` files use the names of :func:`weight_names`; a dict is taken
    as is; the string ``"random"`` (or ``"random:<seed>"``) is the explicit opt-in for a Keras-default random
    initialisation.  Anything else raises, as ``model.load_weights`` does in the reference (models.py:622) --
    a missing checkpoint must never silently give a random-init model."""
    if isinstance(checkpoint_file, dict):
        return checkpoint_file
    name = "" if checkpoint_file is None else os.fspath(checkpoint_file)
    if name == RANDOM_INIT or name.startswith(RANDOM_INIT + ":"):
        seed = int(name.split(":", 1)[1]) if ":" in name else 0
        return random_init_weights(filter_counts, seed=seed)
    how = ("convert the Keras checkpoint once: np.savez(path, **dict(zip([n for n, _ in "
           "tmat_b200.models.weight_names(filter_counts)], keras_model.get_weights()))) -- see INTEGRATION.md")
    if not name.endswith(".npz"):
        raise ValueError(f"checkpoint {name!r}: libtmat_b200 reads .npz checkpoints with "
                         f"tmat_b200.models.weight_names() keys (no h5py/TensorFlow here); {how}; "
                         f"pass checkpoint_file='random' for an explicitly random-initialised model")
    if not os.path.exists(name):
        raise FileNotFoundError(f"checkpoint {name!r} does not exist; {how}")
    with np.load(name) as z:
        return {k: z[k] for k in z.files}


class UNetXceptionPatchSegmentor:
    """models.py:597-653, same constructor and ``predict`` signature."""

    def __init__(self, patch_size, checkpoint_file, filter_counts, ds_ratio=0.5, norm_mean=None,
                 norm_std=None, channels=1):
        self.ds_ratio = ds_ratio
        self.norm_mean = norm_mean
        self.norm_std = norm_std
        self.patch_size = patch_size
        self.channels = channels
        self.model = CudaUNet(load_checkpoint(checkpoint_file, filter_counts), patch_size, filter_counts, channels)

    def predict(self, x, auto_resample=True):
        """models.py:624-653.  With ``auto_resample`` the image is Lanczos-resampled by ``ds_ratio`` before and
        nearest-resampled back after the tiled prediction, exactly as the reference does with Pillow -- including
        its (rows, cols) tuple handed to Pillow's (width, height) parameter and the float32 result of the
        round trip through ``Image.fromarray``."""
        from . import resize as rz
        x = np.asarray(x).astype(np.float32)
        original_shape = x.shape
        target_shape = tuple(np.round(np.multiply(original_shape[:2], self.ds_ratio)).astype(int))
        do_resampling = original_shape != target_shape and auto_resample
        if do_resampling:
            x = rz.pil_resize(x, target_shape, rz.LANCZOS)
        if self.norm_mean is not None and self.norm_std is not None:
            x = (x - self.norm_mean) / self.norm_std
        pred = predict_img_with_smooth_windowing(x, self.patch_size, 2, self.model.predict)
        if do_resampling:
            pred = rz.pil_resize(pred, original_shape, rz.NEAREST)
        return pred


def get_unet_patch_segmentor_from_cfg(cfg_json, checkpoint_file=None, models_dir=None):
    """models.py:656-684: build the segmentor from the reference's JSON config.  The checkpoint is looked up as
    the reference does, ``<MODEL_TRAINING_DIR>/binary_segmentation/checkpoints/<cfg["checkpoint_file"]>``
    (``models_dir`` stands in for ``defs.MODEL_TRAINING_DIR``; default: two levels above the config file, which is
    where the reference keeps ``configs/``), with the ``.h5`` suffix replaced by ``.npz``.  A missing file raises.
    ``checkpoint_file`` overrides the lookup (a path, a weight dict, or ``"random"``)."""
    with open(cfg_json, "r") as f:
        cfg = json.load(f)
    ckpt = checkpoint_file
    if ckpt is None:
        root = models_dir or os.path.join(os.path.dirname(os.path.abspath(cfg_json)), "..", "..")
        name = cfg["checkpoint_file"]
        if name.endswith((".h5", ".hdf5", ".keras")):
            name = name.rsplit(".", 1)[0] + ".npz"
        ckpt = os.path.normpath(os.path.join(root, "binary_segmentation", "checkpoints", name))
    return UNetXceptionPatchSegmentor(cfg["patch_size"], ckpt, cfg["filter_counts"], ds_ratio=cfg.get("ds_ratio", 1),
                                      norm_mean=cfg.get("norm_mean", None), norm_std=cfg.get("norm_std", None),
                                      channels=cfg.get("channels", 1))
