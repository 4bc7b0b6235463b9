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


def _gauss_blur(a, sigma):
    """Separable Gaussian blur (reflect borders), numpy only."""
    r = max(1, int(3 * sigma + 0.5))
    k = np.exp(-0.5 * (np.arange(-r, r + 1) / sigma) ** 2)
    k /= k.sum()
    for axis in (0, 1):
        pad = [(0, 0)] * a.ndim
        pad[axis] = (r, r)
        p = np.pad(a, pad, mode="reflect")
        acc = np.zeros_like(a)
        for i, kv in enumerate(k):
            sl = [slice(None)] * a.ndim
            sl[axis] = slice(i, i + a.shape[axis])
            acc += kv * p[tuple(sl)]
        a = acc
    return a


def make_stereo_pair(w, h, seed=42, n_rect=8, bg_disp=16, min_disp=24, max_disp=140):
    """Generator G2: a rectified colour stereo pair for REAL matchers (cv2.StereoBM / StereoSGBM).
    Left view = 3-channel uniform noise blurred with sigma 1.5 px plus n_rect rectangles of distinct mean colour;
    true disparity = bg_disp px background, rectangles at U[min_disp, max_disp] px; right view by forward warping the
    left one (nearer surfaces win, holes keep an independent noise texture).  Returns (left, right, true_disparity_px)."""
    rng = np.random.default_rng(seed)
    left = _gauss_blur(rng.integers(0, 256, (h, w, 3)).astype(np.float32), 1.5)
    left = 128 + (left - 128) * 2.5                                       # the blur flattens the contrast: stretch it back
    disp = np.full((h, w), bg_disp, np.int32)
    for _ in range(n_rect):
        rw, rh = int(rng.integers(w // 16, w // 3)), int(rng.integers(h // 16, h // 3))
        x, y = int(rng.integers(0, w - rw)), int(rng.integers(0, h - rh))
        disp[y:y + rh, x:x + rw] = int(rng.integers(min_disp, max_disp + 1))
        left[y:y + rh, x:x + rw] = 0.5 * left[y:y + rh, x:x + rw] + rng.integers(0, 128, 3)
    left = np.clip(np.rint(left), 0, 255).astype(np.uint8)
    hole = _gauss_blur(rng.integers(0, 256, (h, w, 3)).astype(np.float32), 1.5)
    right = np.clip(np.rint(128 + (hole - 128) * 2.5), 0, 255).astype(np.uint8)
    ys = np.repeat(np.arange(h)[:, None], w, 1)
    for d in np.unique(disp):                                            # increasing disparity: nearer surfaces overwrite
        m = disp == d
        xs = np.arange(w)[None, :] - d
        ok = m & (xs >= 0)
        right[ys[ok], np.broadcast_to(xs, (h, w))[ok]] = left[ok]
    return left, right, disp
