#!/usr/bin/env python
"""Zero confidence over half the ROI next to a noisy guide (weights ~ 0): the filtered confidence is 0 / subnormal over whole
regions.  Prints GPU-vs-oracle statistics and the time of the final column pass."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import addingdisparityfiltering_b200 as A
from oracle import oracle as O
rng = np.random.default_rng(3)
w, h = 640, 480
view = rng.integers(0, 256, (h, w, 3)).astype(np.uint8)
dl = np.full((h, w), 160, np.int16) + rng.integers(-3, 4, (h, w)).astype(np.int16)
dr = (-dl).astype(np.int16)
dr[:, : w // 2] = 5000                      # LRC fails wherever the right pixel falls into the left half
roi = (32, 0, w - 32, h)
ctx = A.Context(0)
f = A.createDisparityWLSFilterGeneric(True, ctx=ctx); f.setSigmaColor(1.5); f.setDepthDiscontinuityRadius(2)
got = f.filter(dl, view, None, dr, roi)
ref = O.wls_filter(dl, view, dr, roi=roi, lam=8000.0, sigma=1.5, disc_radius=2, want_mask=True)
x, y, ww, hh = roi
g, r = got[y:y + hh, x:x + ww].astype(np.int64), ref["out"][y:y + hh, x:x + ww].astype(np.int64)
d = np.abs(g - r)
print("zero-conf px", float((ref["conf"][y:y + hh, x:x + ww] == 0).mean()), "ref -32768 px", float((r == -32768).mean()), "gpu -32768 px", float((g == -32768).mean()))
print("equal", float((d == 0).mean()), "within1", float((d <= 1).mean()), "max", int(d.max()))
