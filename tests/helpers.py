"""Shared synthetic inputs for the parity tests (seeded; no reference needed at run time)."""
import numpy as np


def toy_pred_func(batch, verbose=0):
    """A deliberately non-equivariant, nonlinear stand-in for Keras ``model.predict``:
    (n, ws, ws) float32 -> (n, ws, ws, 1) float32.  Pure numpy so the generic
    ``pred_func`` path of predict_img_with_smooth_windowing can be checked exactly."""
    x = np.asarray(batch, dtype=np.float32)
    y = x + np.float32(0.3) * np.roll(x, 1, axis=1) - np.float32(0.2) * np.roll(x, 2, axis=2)
    y = y * np.float32(1.7) - np.float32(0.4)
    return (np.float32(1) / (np.float32(1) + np.exp(-y)))[..., None].astype(np.float32)


def synth_projection(h, w, seed):
    """C1-style field: noisy background + bright anti-aliased polylines, in [0, 1]."""
    rng = np.random.default_rng(seed)
    img = rng.normal(0.15, 0.03, (h, w))
    yy, xx = np.mgrid[0:h, 0:w]
    for _ in range(max(4, (h * w) // 26000)):
        x0, y0, x1, y1 = rng.uniform(0, [w, h, w, h])
        width = rng.uniform(1.5, 4.5)
        inten = rng.uniform(0.5, 1.0)
        dx, dy = x1 - x0, y1 - y0
        t = np.clip(((xx - x0) * dx + (yy - y0) * dy) / (dx * dx + dy * dy + 1e-9), 0, 1)
        d = np.hypot(xx - (x0 + t * dx), yy - (y0 + t * dy))
        img = np.maximum(img, inten * np.clip(width - d + 0.5, 0, 1))
    return np.clip(img, 0, 1).astype(np.float32)


def synth_stack(z, h, w, seed, dtype=np.uint16):
    """C2-style Z stack: textured base, per-slice defocus, a constant block for ties."""
    from scipy import ndimage as ndi
    rng = np.random.default_rng(seed)
    top = np.iinfo(dtype).max
    base = rng.uniform(0.003, 0.06, (h, w))
    for _ in range(12):
        cy, cx = rng.uniform(0, [h, w])
        r = rng.uniform(3, max(4, min(h, w) / 8))
        yy, xx = np.mgrid[0:h, 0:w]
        base += rng.uniform(0.2, 0.9) * np.exp(-((yy - cy) ** 2 + (xx - cx) ** 2) / (2 * r * r))
    z0 = rng.uniform(0, z, (h, w))
    z0 = ndi.gaussian_filter(z0, 6)
    out = np.empty((z, h, w), dtype=dtype)
    for k in range(z):
        sl = ndi.gaussian_filter(base, 0.5 + abs(k - z0.mean()) / 4)
        sl = sl + rng.normal(0, 0.004, (h, w))
        out[k] = np.clip(sl, 0, 1) * top
    out[:, : min(16, h), : min(16, w)] = top // 3   # constant block: exercises the tie rule
    return out


def synth_mask(h, w, seed, density=0.1):
    """Branching-like binary mask: thick polylines + round blobs + salt noise."""
    rng = np.random.default_rng(seed)
    m = np.zeros((h, w), dtype=bool)
    yy, xx = np.mgrid[0:h, 0:w]
    n_lines = max(3, int(density * h * w / 900))
    for _ in range(n_lines):
        x0, y0 = rng.uniform(0, [w, h])
        ang = rng.uniform(0, 2 * np.pi)
        ln = rng.uniform(20, 120)
        x1, y1 = x0 + ln * np.cos(ang), y0 + ln * np.sin(ang)
        wd = rng.uniform(1.0, 4.0)
        dx, dy = x1 - x0, y1 - y0
        t = np.clip(((xx - x0) * dx + (yy - y0) * dy) / (dx * dx + dy * dy + 1e-9), 0, 1)
        m |= np.hypot(xx - (x0 + t * dx), yy - (y0 + t * dy)) <= wd
        if rng.random() < 0.6:   # a side branch -> fork
            ang2 = ang + rng.uniform(0.6, 1.4)
            xm, ym = (x0 + x1) / 2, (y0 + y1) / 2
            x2, y2 = xm + 0.5 * ln * np.cos(ang2), ym + 0.5 * ln * np.sin(ang2)
            dx, dy = x2 - xm, y2 - ym
            t = np.clip(((xx - xm) * dx + (yy - ym) * dy) / (dx * dx + dy * dy + 1e-9), 0, 1)
            m |= np.hypot(xx - (xm + t * dx), yy - (ym + t * dy)) <= wd
    for _ in range(max(2, n_lines // 2)):
        cy, cx = rng.uniform(0, [h, w])
        m |= (yy - cy) ** 2 + (xx - cx) ** 2 <= rng.uniform(2, 9) ** 2
    m ^= rng.random((h, w)) < 0.004
    return m


def synth_well(h, w, seed, dtype=np.uint16, invert=False):
    """C3-style well image: bright disc with cell blobs inside a darker surround."""
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w]
    r = np.hypot(yy - h / 2 + rng.uniform(-3, 3), xx - w / 2 + rng.uniform(-3, 3))
    img = np.where(r < 0.42 * min(h, w), 0.55, 0.12) + rng.normal(0, 0.03, (h, w))
    for _ in range(10):
        cy, cx = rng.uniform(0.2, 0.8, 2) * [h, w]
        img += 0.3 * np.exp(-((yy - cy) ** 2 + (xx - cx) ** 2) / (2 * rng.uniform(2, 6) ** 2))
    img = np.clip(img, 0, 1)
    if invert:
        img = 1 - img
    if np.dtype(dtype).kind == "f":
        return img.astype(dtype)
    return (img * np.iinfo(dtype).max).astype(dtype)


def wellmask_points(kind, seed, npts=28):
    """Hull-vertex-like points in [-1, 1]^2 as generate_well_mask hands them over (:208-209): integer pixel
    coordinates of a <= 200-px thumbnail, scaled."""
    rng = np.random.default_rng(seed)
    size = 200
    ang = np.sort(rng.uniform(0, 2 * np.pi, npts))
    if kind == "circle":
        rad = 0.78 + rng.uniform(-0.02, 0.0, npts)
        px, py = rad * np.cos(ang), rad * np.sin(ang)
    else:                                         # rounded square: |x|^8 + |y|^8 = r^8
        rad = 0.8
        c, s_ = np.cos(ang), np.sin(ang)
        k = rad / (np.abs(c) ** 8 + np.abs(s_) ** 8) ** 0.125
        px, py = k * c, k * s_
    px, py = px + 0.04, py - 0.03
    rows = np.round((px + 1) / 2 * size).astype(np.int64)
    cols = np.round((py + 1) / 2 * size).astype(np.int64)
    return rows / size * 2 - 1, cols / size * 2 - 1
