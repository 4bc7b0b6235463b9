#!/usr/bin/env python
"""Per-kernel CUDA-event times of ONE 1080p frame per call (configs[1]) next to the whole-call time with plain launches and
with the CUDA-graph replay: shows where a single call loses against the batched per-frame time."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import addingdisparityfiltering_b200 as A
from addingdisparityfiltering_b200 import synth
from addingdisparityfiltering_b200.batch import BatchFilter

W, H, ROI = 1920, 1080, (160, 0, 1760, 1080)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1
dev = torch.device("cuda", 0)
ctx = A.Context(0)
view, dl, dr = synth.make_textured_scene(W, H, 3, seed=42, max_disp=160)
tv, tl, tr = (torch.from_numpy(a[None]).to(dev).expand(n, *a.shape).contiguous() for a in (view, dl, dr))
out = torch.empty((n, H, W), dtype=torch.int16, device=dev)
conf = torch.empty((n, H, W), dtype=torch.float32, device=dev)
bf = BatchFilter(True, 8000.0, 1.5, 24, 2, ctx=ctx)
res = {"frames_per_call": n}
for graphs in (False, True):
    ctx.set_cuda_graphs(graphs)
    for _ in range(5):
        bf.filter_device(tl, tv, out, tr, conf, ROI)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50):
        bf.filter_device(tl, tv, out, tr, conf, ROI)
    e1.record(); torch.cuda.synchronize()
    res["call_us_graphs" if graphs else "call_us_plain"] = round(e0.elapsed_time(e1) * 1e3 / 50, 1)
ctx.kernel_timing_report(); ctx.set_kernel_timing(True)
for _ in range(20):
    bf.filter_device(tl, tv, out, tr, conf, ROI)
kt = ctx.kernel_timing_report(); ctx.set_kernel_timing(False)
res["kernel_us_per_call"] = {k: round(t / 20 * 1e3, 2) for k, (c, t) in kt.items()}
res["kernel_sum_us"] = round(sum(res["kernel_us_per_call"].values()), 1)
print(json.dumps(res))
