"""Single-frame latency of DisparityWLSFilter::filter (one call = one stereo pair), device-resident and through the
host facade, at cfg 1 (640x480 sample default, half-res), cfg 2 (1080p) and cfg 4 (1080p half-res), with the
CUDA-graph replay of the launch sequence off and on."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import addingdisparityfiltering_b200 as A
from addingdisparityfiltering_b200 import synth
from addingdisparityfiltering_b200.batch import BatchFilter

dev = torch.device("cuda", 0)
ctx = A.Context(0)
CFG = {   # name: (W, H, half_res, radius, lrc)
    "cfg1_640x480_half_res": (640, 480, True, 3, 12),
    "cfg2_1080p": (1920, 1080, False, 2, 24),
    "cfg4_1080p_half_res": (1920, 1080, True, 3, 12),
}


def measure(W, H, half, radius, lrc):
    view, dl, dr, roi = synth.make_artificial_example(W, H, 3, seed=11)
    if half:
        dl, dr = (np.ascontiguousarray(dl[::2, ::2] // 2), np.ascontiguousarray(dr[::2, ::2] // 2))
        roi = tuple(v // 2 for v in roi)
    bf = BatchFilter(True, 8000.0, 1.5, lrc, radius, ctx=ctx)
    tv, tl, tr = (torch.from_numpy(a[None]).to(dev) for a in (view, dl, dr))
    out = torch.empty((1, H, W), dtype=torch.int16, device=dev)
    conf = torch.empty((1, H, W), dtype=torch.float32, device=dev)
    for _ in range(5):
        bf.filter_device(tl, tv, out, tr, conf, roi)
    torch.cuda.synchronize()
    reps = 50
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        bf.filter_device(tl, tv, out, tr, conf, roi)
    e1.record(); torch.cuda.synchronize()
    dev_us = e0.elapsed_time(e1) * 1e3 / reps
    t0 = time.perf_counter()                      # one blocking call at a time: latency as a caller sees it
    for _ in range(reps):
        bf.filter_device(tl, tv, out, tr, conf, roi); torch.cuda.synchronize()
    lat_us = (time.perf_counter() - t0) * 1e6 / reps
    # host facade: numpy in, numpy out (pageable memory), the call a cv::Mat user makes
    f = A.createDisparityWLSFilterGeneric(True, ctx=ctx)
    f.setSigmaColor(1.5); f.setLRCthresh(lrc); f.setDepthDiscontinuityRadius(radius)
    for _ in range(3):
        f.filter(dl, view, None, dr, roi)
    t0 = time.perf_counter()
    for _ in range(20):
        f.filter(dl, view, None, dr, roi)
    host_us = (time.perf_counter() - t0) * 1e6 / 20
    return {"device_us_back_to_back": round(dev_us, 1), "device_us_blocking": round(lat_us, 1), "host_facade_us": round(host_us, 1),
            "Mpix_per_s_device": round(W * H / dev_us, 1), "Mpix_per_s_host_facade": round(W * H / host_us, 1)}


res = {}
for graphs in (False, True):
    ctx.set_cuda_graphs(graphs)
    for name, cfg in CFG.items():
        res[name + ("/graphs" if graphs else "/plain")] = measure(*cfg)
print(json.dumps(res, indent=1))
