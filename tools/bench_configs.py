#!/usr/bin/env python
"""Times the other BASELINE.json configs (not bench lines: recorded in profiles/README.md).
cfg3: fastGlobalSmootherFilter 3840x2160 3-channel guide, num_iter=3, lambda_attenuation=0.25 (device-resident, 32FC1 and 8UC3 src)
cfg4: half-resolution WLS filtering with upscaled output at 1080p (disparities 960x540)."""
import ctypes as C
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, cv2
import addingdisparityfiltering_b200 as A
from addingdisparityfiltering_b200 import synth, _lib
from addingdisparityfiltering_b200.batch import BatchFilter

dev = torch.device("cuda", 0)
ctx = A.Context(0)
res = {}

def timeit(fn, reps=10, warm=3):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps

# ---- cfg4: half-res, batch of 32 ----
W, H, B = 1920, 1080, 32
view, dl, dr = synth.make_textured_scene(W, H, 3, seed=42, max_disp=160)
dl2 = (cv2.resize(dl, None, fx=0.5, fy=0.5) // 2).astype(np.int16); dr2 = (cv2.resize(dr, None, fx=0.5, fy=0.5) // 2).astype(np.int16)
tv = torch.from_numpy(view).to(dev).expand(B, -1, -1, -1).contiguous()
tl = torch.from_numpy(dl2).to(dev).expand(B, -1, -1).contiguous(); tr = torch.from_numpy(dr2).to(dev).expand(B, -1, -1).contiguous()
out = torch.empty((B, H, W), dtype=torch.int16, device=dev); conf = torch.empty((B, H, W), dtype=torch.float32, device=dev)
bf = BatchFilter(True, 8000.0, 1.5, 24, 3, ctx=ctx)
roi = (83, 3, 874, 534)
ms = timeit(lambda: bf.filter_device(tl, tv, out, tr, conf, roi))
ctx.kernel_timing_report(); ctx.set_kernel_timing(True)
for _ in range(5): bf.filter_device(tl, tv, out, tr, conf, roi)
kt = ctx.kernel_timing_report(); ctx.set_kernel_timing(False)
res["cfg4_half_res_1080p"] = {"ms_per_step": ms, "frames": B, "Mpix_per_s": B * W * H / ms / 1e3,
                              "kernels_us_per_frame": {k: round(t / 5 / B * 1e3, 2) for k, (n, t) in kt.items()}}

# ---- cfg3: FGS standalone 4K through dwls_fgs_dev ----
W, H = 3840, 2160
rng = np.random.default_rng(7)
g = cv2.GaussianBlur(rng.integers(0, 256, (H, W, 3)).astype(np.uint8), (0, 0), 2.0)
lib = _lib.lib()
lib.dwls_fgs_dev.argtypes = [C.c_void_p, C.POINTER(_lib.Image), C.POINTER(_lib.Image), C.POINTER(_lib.Image), C.c_double, C.c_double, C.c_double, C.c_int, C.c_void_p]
def img(t, depth, cn):
    return _lib.Image(t.data_ptr(), t.stride(0) * t.element_size(), t.shape[0], t.shape[1], depth, cn)
tg = torch.from_numpy(g).to(dev)
for name, src, depth, cn in (("32FC1", torch.rand((H, W), device=dev) * 255, 5, 1), ("8UC3", tg.clone(), 0, 3)):
    dst = torch.empty_like(src)
    G, S_, D = img(tg, 0, 3), img(src, depth, cn), img(dst, depth, cn)
    stream = torch.cuda.current_stream().cuda_stream or 1
    def run():
        st = lib.dwls_fgs_dev(ctx.handle, C.byref(G), C.byref(S_), C.byref(D), 1000.0, 10.0, 0.25, 3, C.c_void_p(stream)); assert st == 0, st
    ms = timeit(run)
    ctx.kernel_timing_report(); ctx.set_kernel_timing(True)
    for _ in range(5): run()
    kt = ctx.kernel_timing_report(); ctx.set_kernel_timing(False)
    res["cfg3_fgs_4k_" + name] = {"ms_per_call": ms, "Mpix_per_s": W * H / ms / 1e3, "kernels_us_per_call": {k: round(t / 5 * 1e3, 1) for k, (n, t) in kt.items()}}
print(json.dumps(res, indent=1))
