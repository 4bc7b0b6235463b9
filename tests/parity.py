"""Parity statistics shared by the GPU tests, smoke() and bench.py."""
import json
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NAN_S16 = -32768          # saturate_cast<short>(NaN): where the filtered confidence is 0 (disparity_filters.cpp:295-296)


def s16_stats(gpu, ref, roi=None):
    """|delta| histogram of two CV_16S maps (optionally inside roi = x,y,w,h).  Pixels where exactly one side holds the
    NaN marker -32768 (the 0 * inf corner) are counted separately (`corner`) and excluded from `max`."""
    a, b = gpu.astype(np.int64), ref.astype(np.int64)
    if roi is not None:
        x, y, w, h = roi
        a, b = a[y:y + h, x:x + w], b[y:y + h, x:x + w]
    d = np.abs(a - b)
    n = d.size
    corner = (a == NAN_S16) != (b == NAN_S16)
    rest = d[~corner]
    return dict(n=int(n), exact=float((d == 0).sum() / n), within1=float((d <= 1).sum() / n), max=int(rest.max()) if rest.size else 0,
                mean=float(rest.mean()) if rest.size else 0.0, corner=int(corner.sum()))


def assert_s16_parity(gpu, ref, roi=None, frac=0.999, max_abs=2, corner_frac=1e-3):
    """north_star gate: CV_16S filtered disparity within 1 LSB on at least 99.9 % of pixels; on top of it no pixel may be
    off by more than `max_abs`, except where one side is the NaN marker (-32768), and those are bounded by `corner_frac`."""
    st = s16_stats(gpu, ref, roi)
    assert st["within1"] >= frac, st
    assert st["max"] <= max_abs, st
    assert st["corner"] <= corner_frac * st["n"], st
    return st


def record(name, **stats):
    """Appends one line of measured parity figures to gpurun_out/parity_stats.jsonl (read back into BASELINE.md section 3)."""
    path = os.environ.get("DWLS_PARITY_LOG", os.path.join(ROOT, "gpurun_out", "parity_stats.jsonl"))
    try:
        os.makedirs(os.path.dirname(path), exist_ok=True)
        with open(path, "a") as f:
            f.write(json.dumps(dict(case=name, **{k: (float(v) if isinstance(v, (np.floating, float)) else v) for k, v in stats.items()})) + "\n")
    except OSError:
        pass
    print("parity", name, stats)
