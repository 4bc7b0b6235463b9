#!/usr/bin/env python
"""End-to-end throughput of dwls_filter_batch_host (pinned host buffers, copies inside the timed region) at configs[1].
usage: python tools/e2e_bench.py [frames] [steps] [group]   (DWLS_E2E_WINDOWS=0: whole-frame contiguous uploads)"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import addingdisparityfiltering_b200 as A
from addingdisparityfiltering_b200 import synth
from addingdisparityfiltering_b200.batch import BatchFilter

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
group = int(sys.argv[3]) if len(sys.argv) > 3 else 0
W, H, ROI = 1920, 1080, (160, 0, 1760, 1080)
ctx = A.Context(0)
if group:
    ctx.set_group_frames(group)
base = [synth.make_textured_scene(W, H, 3, seed=42 + i, max_disp=160) for i in range(2)]
hv = torch.from_numpy(np.stack([base[f % 2][0] for f in range(B)])).pin_memory()
hl = torch.from_numpy(np.stack([base[f % 2][1] for f in range(B)])).pin_memory()
hr = torch.from_numpy(np.stack([base[f % 2][2] for f in range(B)])).pin_memory()
ho = torch.empty((B, H, W), dtype=torch.int16).pin_memory()
hc = torch.empty((B, H, W), dtype=torch.float32).pin_memory()
bf = BatchFilter(True, 8000.0, 1.5, 24, 2, ctx=ctx)
res = {"frames": B, "group": group, "windows": os.environ.get("DWLS_E2E_WINDOWS", "1")}
for name, conf in (("s16_only", None), ("with_conf", hc)):
    for _ in range(2):
        bf.filter_host(hl, hv, ho, hr, conf, ROI)
    x0 = ctx.transfer_bytes(); t0 = time.perf_counter()
    for _ in range(steps):
        bf.filter_host(hl, hv, ho, hr, conf, ROI)
    dt = time.perf_counter() - t0; x1 = ctx.transfer_bytes()
    res[name] = {"Gpix_per_s": round(B * steps * W * H / dt / 1e9, 3), "h2d_GBps": round((x1[0] - x0[0]) / dt / 1e9, 1), "d2h_GBps": round((x1[1] - x0[1]) / dt / 1e9, 1)}
res["checksum"] = int(ho.to(torch.int64).sum())
print(json.dumps(res))
