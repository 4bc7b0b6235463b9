#!/usr/bin/env python
"""Device StereoBM alone: per-kernel CUDA-event times for the sample's half-resolution matching step at 1080p
(960x540 views, numDisparities 80, window 7, left + right matcher in one launch).  usage: python tools/bm_bench.py [frames] [reps]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import addingdisparityfiltering_b200 as A
from addingdisparityfiltering_b200 import synth

B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
W, H, ND, WS = 960, 540, 80, 7
dev = torch.device("cuda", 0)
ctx = A.Context(0)
pairs = [synth.make_stereo_pair(W, H, seed=70 + i, max_disp=70) for i in range(2)]
import cv2
gl = torch.stack([torch.from_numpy(cv2.cvtColor(pairs[f % 2][0], cv2.COLOR_BGR2GRAY)).to(dev) for f in range(B)]).contiguous()
gr = torch.stack([torch.from_numpy(cv2.cvtColor(pairs[f % 2][1], cv2.COLOR_BGR2GRAY)).to(dev) for f in range(B)]).contiguous()
bm = A.StereoBM.create(ND, WS, ctx=ctx)
A.createDisparityWLSFilter(bm, ctx=ctx)
dl = torch.empty((B, H, W), dtype=torch.int16, device=dev); dr = torch.empty_like(dl)
for _ in range(2):
    bm.compute_device(gl, gr, dl, dr)
torch.cuda.synchronize()
# correctness of what is being timed: frame 0 against cv2
cm = cv2.StereoBM_create(ND, WS); A.createDisparityWLSFilter(cm)
g0, g1 = gl[0].cpu().numpy(), gr[0].cpu().numpy()
ok = bool((dl[0].cpu().numpy() == cm.compute(g0, g1)).all() and (dr[0].cpu().numpy() == A.createRightMatcher(cm).compute(g1, g0)).all())
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    bm.compute_device(gl, gr, dl, dr)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
ctx.kernel_timing_report(); ctx.set_kernel_timing(True)
for _ in range(reps):
    bm.compute_device(gl, gr, dl, dr)
kt = ctx.kernel_timing_report(); ctx.set_kernel_timing(False)
print(json.dumps({"frames": B, "equal_to_cv2": ok, "us_per_frame_both_maps": round(ms * 1e3 / B, 2),
                  "Gdisp_evals_per_s": round(2 * B * W * H * ND / ms / 1e6, 1),
                  "kernels_us_per_frame": {k: round(t / reps / B * 1e3, 2) for k, (n, t) in kt.items()}}))
