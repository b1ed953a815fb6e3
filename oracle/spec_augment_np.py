"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's SpecAugment.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU legs may import this module.

Restates, with every random draw made EXPLICIT (so that a GPU kernel fed the same draws can be compared element by
element), what the reference applies to a training spectrogram between extraction and pack_pathway_output
(/root/reference/audio_slowfast/datasets/epickitchens.py:160-168):

  spec_augment   /root/reference/audio_slowfast/datasets/spec_augment.py:144-191  = time_mask(freq_mask(time_warp(x)))
  time_warp      spec_augment.py:9-44          one control point -> sparse_image_warp
  freq_mask      spec_augment.py:47-91         2 masks, width < 27, filled with the CURRENT mean of the tensor
  time_mask      spec_augment.py:94-141        2 masks, width < 25, same fill
  sparse_image_warp / interpolate_spline / dense_image_warp / interpolate_bilinear
                 /root/reference/audio_slowfast/datasets/sparse_image_warp.py:5-354

What the reference's time warp really computes (read off sparse_image_warp.py:5-125 for ONE control point, and
checked numerically against it): the spline is fitted at the DESTINATION point c = (y_c, x_c + d) with y_c = F // 2,
x_c = the spectrogram VALUE spec[0][y_c][i] at a random column i (spec_augment.py:35 indexes the row with the random
column and uses the value as a coordinate) and d = randrange(-W, W); its flow is f = (0, d).  The order-2 polyharmonic
system for n = 1 is
    [ phi(0) = 0   b^T ] [w]   [f]
    [ b          E   ] [v] = [0] ,   b = (y_c, x_c + d, 1),  E = randn(3, 3) / 1e10  (sparse_image_warp.py:111)
with the solution v = u f / (b . u), w = -f / (b . u), u = E^-1 b.  cross_squared_distance_matrix sums |q|^2 over ALL
query points (sparse_image_warp.py:146-147), so the "distance" of every grid point to c is the same huge number
r0 = sum_q |q|^2 (+ O(1e4), below fp32 resolution) and the radial term is the constant K w, K = 0.5 r0 ln r0.  The
dense flow is therefore zero along frequency and, along time, the AFFINE function
    flow_x(y, x) = d * (u' . (y, x, 1) - 1e-10 K) / (u' . b),        u' = randn(3, 3)^-1 b
-- a random plane, roughly d at the control point.  dense_image_warp then resamples out[y][x] = img[y][x - flow_x(y, x)]
with linear interpolation and clamping to the border (sparse_image_warp.py:293-354).  tests/test_oracle_specaug.py
pins this restatement against the reference functions themselves, replaying their Python / torch random streams.
"""
from __future__ import annotations

import numpy as np


def radial_constant(F, T):
    """K = phi_2(r0) with r0 = sum over the F x T grid of (y^2 + x^2) + |c|^2 ~ the first term."""
    r0 = T * (F - 1) * F * (2 * F - 1) / 6.0 + F * (T - 1) * T * (2 * T - 1) / 6.0
    return 0.5 * r0 * np.log(r0)


def warp_coefficients(d, xc, E, F, T):
    """(a_y, a_x, a_0) of flow_x(y, x) = a_y y + a_x x + a_0 for displacement d, control value xc and the UNSCALED
    3x3 normal block E."""
    b = np.array([F // 2, xc + d, 1.0], dtype=np.float64)
    u = np.linalg.solve(np.asarray(E, dtype=np.float64).reshape(3, 3), b)
    den = float(u @ b)
    a = d * u / den
    a[2] -= d * 1e-10 * radial_constant(F, T) / den
    return a


def draw_params(rand, randn33, row_values, F, T, W=5, Fmax=27, Tmax=25, num_freq_masks=2, num_time_masks=2):
    """Replays the reference's draws IN ITS ORDER.  ``rand``: object with ``randrange(a, b)`` (Python's ``random``
    semantics); ``randn33()``: returns the 3x3 standard-normal block of sparse_image_warp.py:111 (before the / 1e10);
    ``row_values``: spec[0][F // 2] (length T).  Returns a dict of plain numbers:
      i, d          warp column / displacement;  xc = row_values[i]
      coef          (a_y, a_x, a_0): flow_x(y, x) = a_y y + a_x x + a_0
      fmask, tmask  lists of (start, end) half-open intervals actually overwritten (possibly empty)."""
    i = rand.randrange(W, T - W)
    d = rand.randrange(-W, W)
    xc = float(row_values[i])
    out = {"i": int(i), "d": int(d), "xc": xc, "coef": warp_coefficients(d, xc, randn33(), F, T)}
    for key, size, mx, cnt in (("fmask", F, Fmax, num_freq_masks), ("tmask", T, Tmax, num_time_masks)):
        masks = []
        for _ in range(cnt):
            w = rand.randrange(0, mx)
            z = rand.randrange(0, size - w)
            if w == 0:  # "avoids randrange error": the reference RETURNS here, skipping the remaining masks
                break
            end = rand.randrange(z, z + w)
            masks.append((int(z), int(end)))
        out[key] = masks
    return out


def time_warp(spec_ft, coef):
    """spec_ft: float32 [F, T] (the reference's (C, F, T) layout without C).  Returns the warped [F, T] float32."""
    F, T = spec_ft.shape
    y = np.arange(F, dtype=np.float64)[:, None]
    x = np.arange(T, dtype=np.float64)[None, :]
    flow = (coef[0] * y + coef[1] * x + coef[2]).astype(np.float32)            # [F, T] float32 like the reference
    q = (x.astype(np.float32) - flow).astype(np.float32)
    fl = np.minimum(np.maximum(np.floor(q), 0.0), T - 2).astype(np.float32)     # interpolate_bilinear:313-318
    a = np.clip(q - fl, 0.0, 1.0).astype(np.float32)
    i0 = fl.astype(np.int64)
    rows = np.arange(F)[:, None]
    left, right = spec_ft[rows, i0], spec_ft[rows, i0 + 1]
    return (a * (right - left) + left).astype(np.float32)


def apply_masks(spec_ft, fmask, tmask):
    """freq masks then time masks, each filled with the mean of the tensor AS IT IS at that moment (float32 mean)."""
    out = spec_ft.copy()
    for z, e in fmask:
        out[z:e, :] = out.mean(dtype=np.float32)
    for z, e in tmask:
        out[:, z:e] = out.mean(dtype=np.float32)
    return out


def spec_augment(spec_tf, params):
    """spec_tf: float32 [T, F] (the layout of the extractor / the network input).  Returns the augmented [T, F]."""
    x = np.ascontiguousarray(spec_tf.T.astype(np.float32))
    x = time_warp(x, params["coef"])
    x = apply_masks(x, params["fmask"], params["tmask"])
    return np.ascontiguousarray(x.T)
