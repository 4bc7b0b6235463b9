#!/usr/bin/env python
"""One DisparityWLSFilter::filter per call on pageable numpy arrays (what the cv::Mat facade does), 1080p, incl. the confidence map."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import addingdisparityfiltering_b200 as A
from addingdisparityfiltering_b200 import synth
ctx = A.Context(0)
view, dl, dr = synth.make_textured_scene(1920, 1080, 3, seed=42, max_disp=160)
f = A.createDisparityWLSFilterGeneric(True, ctx=ctx)
f.setSigmaColor(1.5); f.setDepthDiscontinuityRadius(2)
roi = (160, 0, 1760, 1080)
for _ in range(5):
    f.filter(dl, view, None, dr, roi)
ts = []
for _ in range(40):
    t0 = time.perf_counter(); f.filter(dl, view, None, dr, roi); ts.append(time.perf_counter() - t0)
ts.sort()
print(json.dumps({"lib": os.environ.get("DWLS_LIB", "default"), "median_ms": round(ts[len(ts) // 2] * 1e3, 3), "best_ms": round(ts[0] * 1e3, 3)}))
