#!/usr/bin/env python
"""Sweep of (frames per launch group, lanes) for the device batch entry point at configs[1] (1080p), device-resident:
small groups on several internal streams keep a group's planes in L2 between its kernels; large groups on one stream
stream everything through HBM.  usage: python tools/lanes_sweep.py [frames] ["g,l g,l ..."]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import addingdisparityfiltering_b200 as A
from addingdisparityfiltering_b200 import synth
from addingdisparityfiltering_b200.batch import BatchFilter

B = int(sys.argv[1]) if len(sys.argv) > 1 else 192
cfgs = [tuple(int(v) for v in c.split(",")) for c in (sys.argv[2].split() if len(sys.argv) > 2 else
        "32,1 1,1 1,2 1,3 1,4 2,1 2,2 2,3 2,4 3,3 4,2 4,3 4,4 8,2 8,3 16,2".split())]
W, H = 1920, 1080
ROI = (160, 0, 1760, 1080)
dev = torch.device("cuda", 0)
ctx = A.Context(0)
base = [synth.make_textured_scene(W, H, 3, seed=42 + i, max_disp=160) for i in range(2)]
tb = [[torch.from_numpy(a).to(dev) for a in fr] for fr in base]
view, dl, dr = (torch.stack([torch.roll(tb[f % 2][k], shifts=3 * f, dims=1) for f in range(B)]).contiguous() for k in range(3))
out = torch.empty((B, H, W), dtype=torch.int16, device=dev)
conf = torch.empty((B, H, W), dtype=torch.float32, device=dev)
bf = BatchFilter(True, 8000.0, 1.5, 24, 2, ctx=ctx)
ref = None
for g, l in cfgs:
    ctx.set_group_frames(g); ctx.set_lanes(l)
    out.fill_(7)
    for _ in range(2):
        bf.filter_device(dl, view, out, dr, conf, ROI)
    torch.cuda.synchronize()
    if ref is None:
        ref = out.clone()
    same = bool(torch.equal(out, ref))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        bf.filter_device(dl, view, out, dr, conf, ROI)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    print(json.dumps({"group": g, "lanes": l, "us_per_frame": round(ms * 1e3 / B, 2), "Gpix_per_s": round(B * W * H / ms / 1e6, 2), "identical": same}), flush=True)
