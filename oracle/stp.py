"""Oracle (test infrastructure): smooth tiled prediction with D4 augmentation.

Restates ``smooth_tiled_predictions.py`` of the reference:
  window          -> ``_spline_window:26-41``, ``_window_2D:47-65``
  padding         -> ``_pad_img:68-79`` (constant = global image minimum)
  D4 do / undo    -> ``_rotate_mirror_do:95-113``, ``_rotate_mirror_undo:116-133``
  tiling+predict  -> ``_windowed_subdivs:136-192`` (batches of 16, ``:23``)
  overlap-add     -> ``_recreate_from_subdivs:195-217``
  driver          -> ``predict_img_with_smooth_windowing:220-267``
PINNED: bit-identical to the imported reference on the golden cases
(tests/golden/make_golden.py, tests/test_oracle_golden.py).
"""
import numpy as np
from numpy.lib.stride_tricks import sliding_window_view

BATCH = 16  # smooth_tiled_predictions.py:23


def triang_even_odd(m):
    """scipy.signal.windows.triang(m) (sym=True) written out."""
    n = np.arange(1, (m + 1) // 2 + 1)
    if m % 2 == 0:
        w = (2 * n - 1.0) / m
        return np.r_[w, w[::-1]]
    w = 2 * n / (m + 1.0)
    return np.r_[w, w[-2::-1]]


def spline_window_1d(ws, power=2):
    """smooth_tiled_predictions.py:26-41."""
    q = int(ws / 4)
    t = triang_even_odd(ws)
    outer = (abs(2 * t) ** power) / 2
    outer[q:-q] = 0
    inner = 1 - (abs(2 * (t - 1)) ** power) / 2
    inner[:q] = 0
    inner[-q:] = 0
    w = inner + outer
    return w / np.average(w)


def window_2d(ws, power=2):
    """smooth_tiled_predictions.py:47-65 -> (ws, ws, 1) float64."""
    w = spline_window_1d(ws, power)[:, None, None]
    return w * w.transpose(1, 0, 2)


def aug_of(ws, subdiv):
    return int(round(ws * (1 - 1.0 / subdiv)))


def d4_do(x, g):
    """Orientation g of the list built at smooth_tiled_predictions.py:103-112."""
    if g >= 4:
        x = x[:, ::-1]
    return np.rot90(x, k=g % 4, axes=(0, 1))


def d4_undo(y, g):
    """Inverse used at smooth_tiled_predictions.py:125-132."""
    y = np.rot90(y, k=(4 - g % 4) % 4, axes=(0, 1))
    if g >= 4:
        y = y[:, ::-1]
    return y


def _predict_one_orientation(padded, ws, subdiv, pred_func):
    step = int(ws / subdiv)
    tiles = sliding_window_view(padded, (ws, ws))[::step, ::step]
    a, b = tiles.shape[:2]
    flat = np.ascontiguousarray(tiles).reshape(a * b, ws, ws)
    if len(flat) > BATCH:
        preds = np.concatenate(
            [pred_func(flat[s:s + BATCH], verbose=0) for s in range(0, len(flat), BATCH)]
        )
    else:
        preds = pred_func(flat, verbose=0)
    w2 = window_2d(ws)
    preds = np.array([p * w2 for p in preds]).reshape(a, b, ws, ws)
    canvas = np.zeros(padded.shape)
    for ia in range(a):
        for ib in range(b):
            i, j = ia * step, ib * step
            canvas[i:i + ws, j:j + ws] = canvas[i:i + ws, j:j + ws] + preds[ia, ib]
    return canvas / (subdiv ** 2)


def predict_img_with_smooth_windowing(input_img, window_size, subdivisions, pred_func):
    """smooth_tiled_predictions.py:220-267."""
    aug = aug_of(window_size, subdivisions)
    padded = np.pad(input_img, ((aug, aug), (aug, aug)), mode="constant",
                    constant_values=input_img.min())
    per_orient = []
    for g in range(8):
        y = _predict_one_orientation(np.array(d4_do(padded, g)), window_size,
                                     subdivisions, pred_func)
        per_orient.append(np.array(d4_undo(y, g)))
    merged = np.mean(per_orient, axis=0)
    out = merged[aug:-aug, aug:-aug]
    return out[: input_img.shape[0], : input_img.shape[1]]


def tile_origin(g, i, j, PM, PN, ws):
    """top-left corner, in the ORIGINAL padded frame, of the
    tile that sits at (i, j) in the frame of orientation g."""
    k = g % 4
    if k == 0:
        r0, c0 = i, j
    elif k == 1:
        r0, c0 = j, PN - ws - i
    elif k == 2:
        r0, c0 = PM - ws - i, PN - ws - j
    else:
        r0, c0 = PM - ws - j, i
    if g >= 4:
        c0 = PN - ws - c0
    return r0, c0
