"""Parity statistics shared by the GPU tests, smoke() and bench.py."""
import numpy as np


def s16_stats(gpu, ref, roi=None):
    """|delta| histogram of two CV_16S maps (optionally inside roi = x,y,w,h)."""
    a, b = gpu.astype(np.int64), ref.astype(np.int64)
    if roi is not None:
        x, y, w, h = roi
        a, b = a[y:y + h, x:x + w], b[y:y + h, x:x + w]
    d = np.abs(a - b)
    n = d.size
    return dict(n=int(n), exact=float((d == 0).sum() / n), within1=float((d <= 1).sum() / n), max=int(d.max()),
                mean=float(d.mean()))


def assert_s16_parity(gpu, ref, roi=None, frac=0.999):
    """north_star gate: CV_16S filtered disparity within 1 LSB on at least 99.9 % of pixels."""
    st = s16_stats(gpu, ref, roi)
    assert st["within1"] >= frac, st
    return st
