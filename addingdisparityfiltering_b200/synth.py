"""Synthetic stereo inputs.

G1 restates MakeArtificialExample of the reference's perf test
(modules/ximgproc/perf/perf_disparity_wls_filter.cpp:95-167): flat background and a centred
foreground rectangle at disparity 0.15*w, additive Gaussian noise (sigma 6) on the guide and on
both disparity maps, ROI = (0.15*w, 0, w-0.15*w, h).  numpy's generator replaces cv::RNG, so the
pixel values differ from an OpenCV run but the CPU checker and the GPU always see identical arrays.
"""
import numpy as np


def make_artificial_example(w, h, channels=3, seed=0):
    rng = np.random.default_rng(seed)
    bg = int(rng.uniform(0.0, 255.0))
    fg = int(rng.uniform(0.0, 255.0))
    rect_w = int(rng.uniform(w // 16, w // 2))
    rect_h = int(rng.uniform(h // 16, h // 2))
    rect_disp = int(0.15 * w)
    sigma = 6.0
    x0, y0 = (w - rect_w) // 2, (h - rect_h) // 2
    view = np.full((h, w, channels), bg, np.float64)
    view[y0:y0 + rect_h, x0:x0 + rect_w] = fg
    left = np.zeros((h, w), np.float64)
    right = np.zeros((h, w), np.float64)
    left[y0:y0 + rect_h, x0:x0 + rect_w] = 16 * rect_disp
    xr = x0 - rect_disp
    right[y0:y0 + rect_h, max(xr, 0):xr + rect_w] = -16 * rect_disp
    view = np.clip(np.rint(view + rng.normal(0, sigma, view.shape)), 0, 255).astype(np.uint8)
    left = np.clip(np.rint(left + rng.normal(0, sigma, left.shape)), -32768, 32767).astype(np.int16)
    right = np.clip(np.rint(right + rng.normal(0, sigma, right.shape)), -32768, 32767).astype(np.int16)
    if channels == 1:
        view = view[:, :, 0]
    return view, left, right, (rect_disp, 0, w - rect_disp, h)


def make_textured_scene(w, h, channels=3, seed=42, max_disp=160, n_rect=8):
    """G2-like scene without a stereo matcher: textured guide, piecewise-constant true disparity with
    matcher-style noise and occlusion garbage, consistent left/right maps (left d > 0, right -d)."""
    rng = np.random.default_rng(seed)
    view = rng.integers(0, 256, (h, w, channels)).astype(np.float32)
    for _ in range(2):                                   # cheap blur so weights are not all ~0
        view = (view + np.roll(view, 1, 0) + np.roll(view, 1, 1) + np.roll(view, -1, 0) + np.roll(view, -1, 1)) / 5
    disp = np.full((h, w), 16.0)
    for _ in range(n_rect):
        rw, rh = int(rng.integers(w // 16, w // 3)), int(rng.integers(h // 16, h // 3))
        x, y = int(rng.integers(0, w - rw)), int(rng.integers(0, h - rh))
        d = float(rng.integers(24, max(25, max_disp - 20)))
        disp[y:y + rh, x:x + rw] = d
        view[y:y + rh, x:x + rw] = 0.5 * view[y:y + rh, x:x + rw] + rng.integers(0, 128, channels)
    view = np.clip(view, 0, 255).astype(np.uint8)
    left = np.rint(16 * disp + rng.normal(0, 4, disp.shape))
    xs = np.arange(w)[None, :] - np.rint(disp).astype(int)
    right = np.full((h, w), 0.0)
    ys = np.repeat(np.arange(h)[:, None], w, 1)
    ok = xs >= 0
    right[ys[ok], xs[ok]] = -(16 * disp[ok])
    right = np.rint(right + rng.normal(0, 4, right.shape))
    bad = rng.random((h, w)) < 0.02                      # mismatches
    left[bad] = rng.integers(0, 16 * max_disp, int(bad.sum()))
    left = np.clip(left, -32768, 32767).astype(np.int16)
    right = np.clip(right, -32768, 32767).astype(np.int16)
    if channels == 1:
        view = view[:, :, 0]
    return view, left, right
