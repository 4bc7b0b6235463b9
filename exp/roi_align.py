#!/usr/bin/env python
"""cfg-2 workload with ROI x offsets of different alignment: per-kernel times (does the CV_16S output alignment matter?)"""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import addingdisparityfiltering_b200 as A
from addingdisparityfiltering_b200 import synth
from addingdisparityfiltering_b200.batch import BatchFilter
dev = torch.device("cuda", 0); ctx = A.Context(0)
W, H, B = 1920, 1080, 32
view, dl, dr = synth.make_textured_scene(W, H, 3, seed=42, max_disp=160)
tv = torch.from_numpy(view).to(dev).expand(B, -1, -1, -1).contiguous()
tl = torch.from_numpy(dl).to(dev).expand(B, -1, -1).contiguous(); tr = torch.from_numpy(dr).to(dev).expand(B, -1, -1).contiguous()
out = torch.empty((B, H, W), dtype=torch.int16, device=dev); conf = torch.empty((B, H, W), dtype=torch.float32, device=dev)
bf = BatchFilter(True, 8000.0, 1.5, 24, 2, ctx=ctx)
for roi in ((160, 0, 1760, 1080), (166, 0, 1748, 1080), (168, 0, 1744, 1080), (161, 0, 1752, 1080), (160, 6, 1760, 1068), (166, 6, 1748, 1068)):
    for _ in range(3): bf.filter_device(tl, tv, out, tr, conf, roi)
    ctx.kernel_timing_report(); ctx.set_kernel_timing(True)
    for _ in range(5): bf.filter_device(tl, tv, out, tr, conf, roi)
    kt = ctx.kernel_timing_report(); ctx.set_kernel_timing(False)
    print(roi, {k: round(t / 5 / B * 1e3, 2) for k, (n, t) in kt.items()})
