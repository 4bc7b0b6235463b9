#!/usr/bin/env python
"""bench.py — 1080p DisparityWLSFilter::filter throughput on B200 (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--frames B]

One "step" = one pass of the hot path (confidence + LRC + weights + 6 FGS sweeps + normalise) over a batch of B = 512
independent 1080p stereo frames (BASELINE.json configs[4]; every frame has the configs[1] shape the metric is quoted on:
1920x1080, cv2.StereoSGBM(0,160,3) left + right disparities of a synthetic stereo pair, ROI (160,0,1760,1080),
discontinuity radius 2, LRC threshold 24, lambda 8000, sigma 1.5, 3-channel guide), resident in HBM.

N > 1 (torchrun, one process per GPU): the SAME 512 frames are sharded in contiguous blocks over the ranks
("strong" scaling; frames are independent, disparity_filters.cpp:138-139), each rank filters its block chunk by chunk and
the CV_16S results are gathered to rank 0 over NCCL INSIDE the timed region: the grouped send/recv of chunk k runs on
NCCL's stream beside the kernels of chunk k+1 (batch.ShardedBatch).  `value` = compute + gather; `compute_only` is the
same step without the gather.  Time = CUDA events, max over ranks.

`e2e` times the host-buffer C-ABI call (dwls_filter_batch_host: pinned host memory, H2D and D2H copies inside the timed
region) with and without the confidence map copied back.  At N = 1 the line also carries `single_call` (one 1080p frame:
device call and the cv::Mat-facade call on pageable memory) and `other_configs` (configs[2] FGS 3840x2160, configs[3]
half-resolution 1080p).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "Mpix/s for 1080p DisparityWLSFilter::filter at 1/2/4/8 B200; achieved HBM GB/s"
W, H, CN = 1920, 1080, 3
ROI = (160, 0, 1760, 1080)
LAMBDA, SIGMA, RADIUS, LRC = 8000.0, 1.5, 2, 24
WORKLOAD = "configs[1] frames (1920x1080 StereoSGBM left+right disparities, confidence-aware WLS filter with LRC check) in the " \
           "configs[4] batch of 512 frames sharded over the GPUs"
N_BASE = 2


def algorithmic_bytes(n_roi, f_disp, f_view, cn, S=2, P=6):
    """SURVEY section 8d B_alg, full-resolution confidence path, bytes per frame (whole pipeline and per kernel)."""
    per = {
        "conf_kernel<right>": 2 * n_roi + 4 * n_roi,
        "conf_kernel<left+lrc>": 2 * n_roi + 2 * n_roi + 4 * n_roi + 4 * f_disp + 4 * n_roi,
        "weights_kernel": (cn + 8) * n_roi,
        "hsolve_kernel": (4 + 8 * S) * n_roi,
        "vsolve_kernel": (4 + 8 * S) * n_roi,
        "vsolve_kernel<final>": (4 + 4 * S) * n_roi + 2 * n_roi,
        "border_kernel": 2 * (f_view - n_roi),
    }
    total = per["conf_kernel<right>"] + per["conf_kernel<left+lrc>"] + per["weights_kernel"] + 3 * per["hsolve_kernel"] \
        + 2 * per["vsolve_kernel"] + per["vsolve_kernel<final>"] + per["border_kernel"]
    return total, per


def measured_peak_gbs():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    FIELDS = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.lines, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.FIELDS, "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def mark(self):
        return len(self.lines)

    def stop(self, lo=0, hi=None):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.12)
        self.proc.terminate()
        rows = [l.split(", ") for l in self.lines[lo:hi] if l.count(",") >= 5]
        if not rows:
            rows = [l.split(", ") for l in self.lines if l.count(",") >= 5]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm = sorted(float(r[0]) for r in rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[2 + i].strip() == "Active" for r in rows)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(rows[0][1]), "reasons": reasons, "samples": len(rows)}


def make_base_frames(n_base):
    """The distinct stereo frames every step is built from (frame f of the batch = base frame f % n_base shifted
    horizontally).  Generator G2 of SURVEY section 8d: a synthetic colour stereo pair matched by the REAL
    cv2.StereoSGBM configured exactly as createDisparityWLSFilter / createRightMatcher configure it
    (disparity_filters.cpp:404-408, 430-441; sample lines 162-175).  Returns ([(view, disp_left, disp_right)], source)."""
    from addingdisparityfiltering_b200 import synth
    try:
        import cv2
        import addingdisparityfiltering_b200 as A
        frames = []
        for i in range(n_base):
            left, right, _ = synth.make_stereo_pair(W, H, seed=42 + i)
            lm = cv2.StereoSGBM_create(0, 160, 3)
            lm.setP1(24 * 9); lm.setP2(96 * 9); lm.setPreFilterCap(63); lm.setMode(cv2.STEREO_SGBM_MODE_SGBM_3WAY)
            f = A.createDisparityWLSFilter(lm)                   # host-only: mutates the matcher, derives the margins
            assert f._offsets == (ROI[0], 0, 0, 0) and f.getDepthDiscontinuityRadius() == RADIUS
            rm = A.createRightMatcher(lm)
            frames.append((left, lm.compute(left, right), rm.compute(right, left)))
        return frames, "cv2.StereoSGBM(0,160,3,P1=216,P2=864,cap 63,3WAY) on synth.make_stereo_pair (G2)"
    except ImportError:
        return [synth.make_textured_scene(W, H, CN, seed=42 + i, max_disp=160) for i in range(n_base)], "synth.make_textured_scene (cv2 not importable)"


def cpu_kind():
    """"reference": oracle/_ref, the reference's own sources compiled from /root/reference against the
    OpenCV-core stand-in (prebuilt, travels with the repo); else "port": the threaded CPU oracle."""
    from oracle import ref as R
    return "reference" if R.available() else "port"


def cpu_frames_per_sec(base, n_frames, threads):
    """Times the reference's CPU implementation of the path on n_frames of the workload, all host threads."""
    from oracle import oracle as O
    from oracle import ref as R
    use_ref = R.available()
    if use_ref:
        R.set_num_threads(threads)
    t0 = time.perf_counter()
    for i in range(n_frames):
        view, dl, dr = base[i % len(base)]
        if use_ref:
            r = R.wls_filter(dl, view, dr, ROI, True, LAMBDA, SIGMA, LRC, RADIUS)
        else:
            r = O.wls_filter(dl, view, dr, roi=ROI, lam=LAMBDA, sigma=SIGMA, lrc_thresh=LRC, disc_radius=RADIUS,
                             assoc=1, nstripes=threads, threaded=True)
        assert r["rc"] == 0
    return n_frames / (time.perf_counter() - t0)


def default_chunk(world):
    """Frames per launch / gather chunk: 32-frame launches are the most efficient ones (one rank, or two ranks with 256 frames
    each); with 128 or 64 frames per rank the chunks shrink to 16 so that the gather starts early and its exposed tail is short."""
    return 32 if world <= 2 else 16


def default_tail(world):
    return 0 if world <= 2 else 4          # N = 2, measured: 32-frame chunks without a halved tail 72.5 Gpix/s, with one 71.9


def base_config(frames_total, world, source, chunk=0):
    """`config` is the same dict in the GPU arm and in the reference arm (the driver compares them); what is specific to a
    GPU run (shards, chunk plan, gather) goes into the line's `run` block."""
    chunk = chunk or default_chunk(world)
    return {"workload": WORKLOAD, "frames_total": frames_total, "roi": list(ROI), "lambda": LAMBDA, "sigma_color": SIGMA,
            "disc_radius": RADIUS, "lrc_thresh": LRC, "guide_channels": CN, "confidence_export": True, "inputs": source,
            "l2": "inputs larger than L2: a launch chunk's working set (%.0f MB inputs+outputs for %d frames) exceeds the 126 MB L2, "
                  "no flush between iterations" % (chunk * (2 * 2 * W * H + CN * W * H + 6 * W * H) / 1e6, chunk)}


CPU_NOTE = "kind=reference: the reference's fgs_filter.cpp + disparity_filters.cpp compiled unmodified (-O3, SSE2 universal intrinsics, " \
           "parallel_for_ over all host threads) against the in-repo OpenCV-core stand-in; boxFilter/resize are the stand-in's"


def run_reference(args, out=sys.stdout):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import oracle as O
    threads = O.hardware_threads()
    base, source = make_base_frames(N_BASE)
    S = args.ref_frames
    for _ in range(args.warmup):
        cpu_frames_per_sec(base, S, threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_frames_per_sec(base, S, threads)
    dt = time.perf_counter() - t0
    v = args.steps * S * W * H / dt / 1e6
    sample = "%d frames per step of the same 1080p workload (a bounded sample of the %d-frame batch), %d steps" % (S, args.frames, args.steps)
    cfg = base_config(args.frames, args.gpus, source)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "Mpix/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": cfg,
        "cpu_baseline": {"value": v, "unit": "Mpix/s", "cores": threads, "kind": cpu_kind(), "sample": sample},
        "e2e": {"value": v, "unit": "Mpix/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "note": CPU_NOTE}), file=out, flush=True)


def cuda_time(torch, fn, reps, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def extras_single_gpu(torch, A, ctx, dev, base, peak):
    """N = 1 only: single-call latency (device and cv::Mat-facade path) and the other BASELINE configs, each device-timed."""
    import ctypes as C
    from addingdisparityfiltering_b200 import _lib, synth
    from addingdisparityfiltering_b200.batch import BatchFilter
    res = {}
    view, dl, dr = base[0]
    # ---- one 1080p frame per call ----
    bf = BatchFilter(True, LAMBDA, SIGMA, LRC, RADIUS, ctx=ctx)
    tv, tl, tr = (torch.from_numpy(a[None]).to(dev) for a in (view, dl, dr))
    out = torch.empty((1, H, W), dtype=torch.int16, device=dev)
    conf = torch.empty((1, H, W), dtype=torch.float32, device=dev)
    dev_us = cuda_time(torch, lambda: bf.filter_device(tl, tv, out, tr, conf, ROI), 50, 5) * 1e3
    n_roi = ROI[2] * ROI[3]
    alg, _ = algorithmic_bytes(n_roi, W * H, W * H, CN)
    f = A.createDisparityWLSFilterGeneric(True, ctx=ctx)
    f.setLambda(LAMBDA); f.setSigmaColor(SIGMA); f.setLRCthresh(LRC); f.setDepthDiscontinuityRadius(RADIUS)
    for _ in range(3):
        f.filter(dl, view, None, dr, ROI)
    t0 = time.perf_counter()
    for _ in range(20):
        f.filter(dl, view, None, dr, ROI)
    host_ms = (time.perf_counter() - t0) * 1e3 / 20
    res["single_call"] = {"device_us": round(dev_us, 1), "device_Mpix_per_s": round(W * H / dev_us, 1), "device_alg_roofline_frac": round(alg / (dev_us * 1e-6) / 1e9 / peak, 3),
                          "facade_pageable_ms": round(host_ms, 3), "facade_Mpix_per_s": round(W * H / host_ms / 1e3, 1),
                          "what": "one 1080p frame per call: dwls_filter_batch_dev with n_frames=1 on device buffers (graph replay), and "
                                  "dwls_filter_host on pageable numpy arrays incl. the confidence map (what wls->filter(cv::Mat...) calls)"}
    other = {}
    # ---- the device matcher feeding the filter (sample default path at 1080p: views halved, StereoBM(80, 7) left + right
    #      maps, filtered at full resolution; samples/disparity_filtering.cpp:135-190), everything born in HBM ----
    try:
        from addingdisparityfiltering_b200.stereo import halve_views_device
        B = 32
        pairs = [synth.make_stereo_pair(W, H, seed=60 + i) for i in range(2)]
        tl2 = torch.stack([torch.from_numpy(pairs[f % 2][0]).to(dev) for f in range(B)]).contiguous()
        tr2 = torch.stack([torch.from_numpy(pairs[f % 2][1]).to(dev) for f in range(B)]).contiguous()
        bm = A.StereoBM.create(80, 7, ctx=ctx)
        wls = A.createDisparityWLSFilter(bm, ctx=ctx)             # margins (83,3,3,3), radius 3; switches the matcher's post-filters off
        bfm = BatchFilter(True, LAMBDA, SIGMA, wls.getLRCthresh(), wls.getDepthDiscontinuityRadius(), offsets=wls._offsets, ctx=ctx)
        mdl = torch.empty((B, H // 2, W // 2), dtype=torch.int16, device=dev)
        mdr = torch.empty_like(mdl)
        mo = torch.empty((B, H, W), dtype=torch.int16, device=dev)
        mc = torch.empty((B, H, W), dtype=torch.float32, device=dev)

        def match_and_filter():
            hl, hr = halve_views_device(tl2, tr2, ctx=ctx)
            bm.compute_device(hl, hr, mdl, mdr)
            bfm.filter_device(mdl, tl2, mo, mdr, mc)
        ms_all = cuda_time(torch, match_and_filter, 5, 2)
        ms_match = cuda_time(torch, lambda: bm.compute_device(*halve_views_device(tl2, tr2, ctx=ctx), mdl, mdr), 5, 2)
        t0 = time.perf_counter()
        import cv2
        cm = cv2.StereoBM_create(80, 7)
        A.createDisparityWLSFilter(cm)
        g0, g1 = (cv2.cvtColor(cv2.resize(pairs[0][k], None, fx=0.5, fy=0.5), cv2.COLOR_BGR2GRAY) for k in range(2))
        t0 = time.perf_counter()
        want_l = cm.compute(g0, g1)
        want_r = A.createRightMatcher(cm).compute(g1, g0)
        cv_ms = (time.perf_counter() - t0) * 1e3
        assert (mdl[0].cpu().numpy() == want_l).all() and (mdr[0].cpu().numpy() == want_r).all(), "device matcher differs from cv2.StereoBM"
        res["match_and_filter"] = {
            "what": "sample default path at 1080p on device-resident views, 32 frames per call: halve both views, StereoBM(80,7) left + right "
                    "matcher (bit-equal to cv2.StereoBM, checked here on frame 0), confidence-aware filter at full resolution",
            "us_per_frame": round(ms_all * 1e3 / B, 2), "Mpix_per_s": round(B * W * H / ms_all / 1e3, 1),
            "matching_us_per_frame": round(ms_match * 1e3 / B, 2), "cv2_StereoBM_both_maps_ms_per_frame_host": round(cv_ms, 2)}
        del tl2, tr2, mdl, mdr, mo, mc
    except ImportError:
        pass
    # ---- configs[0]: 640x480 sample default (320x240 StereoBM maps, ROI (83,3,234,234), r=3), 128 frames ----
    try:
        import cv2
        from tests.stereo_inputs import sample_matching
        B = 128
        l0, r0, _ = synth.make_stereo_pair(640, 480, seed=42, max_disp=140)
        f0, d0l, d0r = sample_matching(l0, r0, "bm", downscale=True, max_disp=160, ctx=ctx)
        sv = torch.from_numpy(l0).to(dev).expand(B, -1, -1, -1).contiguous()
        sl = torch.from_numpy(d0l).to(dev).expand(B, -1, -1).contiguous()
        sr = torch.from_numpy(d0r).to(dev).expand(B, -1, -1).contiguous()
        so = torch.empty((B, 480, 640), dtype=torch.int16, device=dev)
        sc = torch.empty((B, 480, 640), dtype=torch.float32, device=dev)
        bf0 = BatchFilter(True, LAMBDA, SIGMA, f0.getLRCthresh(), f0.getDepthDiscontinuityRadius(), offsets=f0._offsets, ctx=ctx)
        ms = cuda_time(torch, lambda: bf0.filter_device(sl, sv, so, sr, sc), 5, 2)
        N0, n0, fd0, F0 = 468 * 468, 234 * 234, 320 * 240, 640 * 480
        alg0 = 14 * n0 + 4 * fd0 + 6 * fd0 + 6 * F0 + 14 * N0 + (CN + 8) * N0 + 6 * (4 * N0 + 16 * N0) - 8 * N0 + 2 * F0
        other["configs[0] 640x480 StereoBM sample default (320x240 maps, ROI (83,3,234,234), r=3), 128 frames"] = {
            "Mpix_per_s": round(B * F0 / ms / 1e3, 1), "us_per_frame": round(ms * 1e3 / B, 2), "alg_roofline_frac": round(alg0 * B / (ms * 1e-3) / 1e9 / peak, 3)}
        del sv, sl, sr, so, sc
    except ImportError:
        pass
    # ---- configs[3]: half resolution, 32 frames ----
    try:
        import cv2
        B = 32
        dl2 = (cv2.resize(dl, None, fx=0.5, fy=0.5) // 2).astype(np.int16)
        dr2 = (cv2.resize(dr, None, fx=0.5, fy=0.5) // 2).astype(np.int16)
        hv = torch.from_numpy(view).to(dev).expand(B, -1, -1, -1).contiguous()
        hl = torch.from_numpy(dl2).to(dev).expand(B, -1, -1).contiguous()
        hr = torch.from_numpy(dr2).to(dev).expand(B, -1, -1).contiguous()
        o2 = torch.empty((B, H, W), dtype=torch.int16, device=dev)
        c2 = torch.empty((B, H, W), dtype=torch.float32, device=dev)
        bf2 = BatchFilter(True, LAMBDA, SIGMA, 24, 3, ctx=ctx)
        roi2 = (83, 3, 874, 534)
        ms = cuda_time(torch, lambda: bf2.filter_device(hl, hv, o2, hr, c2, roi2), 5, 2)
        N, n, fd = 1748 * 1068, 874 * 534, 960 * 540
        alg4 = 14 * n + 4 * fd + 6 * fd + 6 * W * H + 14 * N + (CN + 8) * N + 6 * (4 * N + 16 * N) - 8 * N + 2 * W * H
        other["configs[3] half-res 1080p (960x540 maps, ROI (83,3,874,534), r=3), 32 frames"] = {
            "Mpix_per_s": round(B * W * H / ms / 1e3, 1), "us_per_frame": round(ms * 1e3 / B, 2), "alg_roofline_frac": round(alg4 * B / (ms * 1e-3) / 1e9 / peak, 3)}
        del hv, hl, hr, o2, c2
    except ImportError:
        pass
    # ---- configs[2]: fastGlobalSmootherFilter 3840x2160, 3-channel guide ----
    W4, H4 = 3840, 2160
    rng = np.random.default_rng(7)
    g = synth._gauss_blur(rng.integers(0, 256, (H4, W4, 3)).astype(np.float32), 2.0)
    g = np.clip(np.rint(128 + (g - 128) * 3.0), 0, 255).astype(np.uint8)
    lib = _lib.lib()
    lib.dwls_fgs_dev.argtypes = [C.c_void_p, C.POINTER(_lib.Image), C.POINTER(_lib.Image), C.POINTER(_lib.Image), C.c_double, C.c_double, C.c_double, C.c_int, C.c_void_p]

    def img(t, depth, cn):
        return _lib.Image(t.data_ptr(), t.stride(0) * t.element_size(), t.shape[0], t.shape[1], depth, cn)
    tg = torch.from_numpy(g).to(dev)
    N4 = W4 * H4
    for name, src, depth, cn, b in (("32FC1", torch.rand((H4, W4), device=dev) * 255, 5, 1, 4), ("8UC3", tg.clone(), 0, 3, 1)):
        dst = torch.empty_like(src)
        G_, S_, D_ = img(tg, 0, 3), img(src, depth, cn), img(dst, depth, cn)
        stream = torch.cuda.current_stream().cuda_stream or 1

        def run():
            st = lib.dwls_fgs_dev(ctx.handle, C.byref(G_), C.byref(S_), C.byref(D_), 1000.0, 10.0, 0.25, 3, C.c_void_p(stream))
            assert st == 0, st
        ms = cuda_time(torch, run, 10, 3)
        alg3 = (3 + 8) * N4 + 6 * (4 * N4 + 8 * cn * N4) - 2 * (4 - b) * cn * N4
        other["configs[2] fastGlobalSmootherFilter 3840x2160 guide 8UC3, src %s, one frame" % name] = {
            "ms_per_call": round(ms, 4), "Mpix_per_s": round(N4 / ms / 1e3, 1), "alg_roofline_frac": round(alg3 / (ms * 1e-3) / 1e9 / peak, 3)}
    res["other_configs"] = other
    return res


def main(out=sys.stdout):
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--frames", type=int, default=512, help="frames of the whole batch per step (sharded over the GPUs)")
    ap.add_argument("--chunk", type=int, default=0, help="frames per launch / gather chunk (0 = 32 on one or two GPUs, 16 on more)")
    ap.add_argument("--tail-min", type=int, default=-1, help="smallest halved tail chunk (-1 = none on one or two GPUs, 4 on more)")
    ap.add_argument("--gather-mode", default="auto", choices=["auto", "root", "all", "peer"],
                    help="auto = peer-memory pushes (CUDA IPC + copy engines), NCCL send/recv to the root if that cannot be set up")
    ap.add_argument("--root-extra", type=int, default=0, help="frames the gather root owns beyond an even share (its results need no transfer)")
    ap.add_argument("--group", type=int, default=0, help="frames per kernel launch group inside a chunk (0 = library default)")
    ap.add_argument("--lanes", type=int, default=0, help="internal streams the launch groups of a chunk are dealt to (0 = library default)")
    ap.add_argument("--e2e-frames", type=int, default=64, help="frames per rank and step of the end-to-end (host buffer) leg")
    ap.add_argument("--ref-frames", type=int, default=4)
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--interleave", action="store_true", help="A/B switch: sector-interleaved solver planes (dwls_set_plane_interleave)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline leg (tuning runs only)")
    ap.add_argument("--no-extras", action="store_true", help="skip single_call / other_configs (tuning runs only)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args, out)

    import torch
    import addingdisparityfiltering_b200 as A
    from addingdisparityfiltering_b200.batch import BatchFilter, ShardedBatch

    rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    ctx = A.Context(local)
    if args.group:
        ctx.set_group_frames(args.group)
    if args.interleave:
        ctx.set_plane_interleave(True)
    if args.lanes:
        ctx.set_lanes(args.lanes)
    chunk = args.chunk or default_chunk(world)
    tail_min = args.tail_min if args.tail_min >= 0 else default_tail(world)

    # ---- inputs: rank 0 runs the stereo matcher, the other ranks receive the base frames over NCCL ----
    if rank == 0:
        base, source = make_base_frames(N_BASE)
        tb = [[torch.from_numpy(a).to(dev) for a in fr] for fr in base]
    else:
        base, source = None, None
        tb = [[torch.empty((H, W, CN), dtype=torch.uint8, device=dev), torch.empty((H, W), dtype=torch.int16, device=dev),
               torch.empty((H, W), dtype=torch.int16, device=dev)] for _ in range(N_BASE)]
    if dist is not None:
        for fr in tb:
            for t in fr:
                dist.broadcast(t.view(torch.uint8), src=0)
    gather_mode = "peer" if args.gather_mode == "auto" else args.gather_mode
    fallback = None
    try:
        sb = ShardedBatch(args.frames, (H, W), chunk=chunk, tail_min=tail_min, mode=gather_mode, dst=0, device=dev, root_extra=args.root_extra, ctx=ctx)
    except A.DwlsError as e:                 # raised on every rank together (ShardedBatch._peer_setup)
        if args.gather_mode != "auto":
            raise
        fallback, gather_mode = str(e), "root"
        sb = ShardedBatch(args.frames, (H, W), chunk=chunk, tail_min=tail_min, mode=gather_mode, dst=0, device=dev, root_extra=args.root_extra, ctx=ctx)
    nloc = sb.end - sb.begin

    def build(k):          # frame f of the batch = base frame f % N_BASE rolled by 3 f columns (left / right / view alike)
        return torch.stack([torch.roll(tb[f % N_BASE][k], shifts=3 * f, dims=1) for f in range(sb.begin, sb.end)]).contiguous()
    view, dl, dr = build(0), build(1), build(2)
    conf = torch.empty((nloc, H, W), dtype=torch.float32, device=dev)
    bf = BatchFilter(True, LAMBDA, SIGMA, LRC, RADIUS, ctx=ctx)

    def filt(f0, nf, out):
        bf.filter_device(dl[f0:f0 + nf], view[f0:f0 + nf], out, dr[f0:f0 + nf], conf[f0:f0 + nf], ROI)

    def sync_all():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(steps, gather):
        sync_all()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            sb.run(filt, gather=gather)
        e1.record()
        sync_all()
        ms = e0.elapsed_time(e1)
        if dist is not None:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    for _ in range(args.warmup):
        sb.run(filt, gather=True)
    sync_all()
    sampler = ClockSampler(local) if rank == 0 else None
    time.sleep(0.15 if rank == 0 else 0.0)
    lo = sampler.mark() if sampler else 0
    launches0 = ctx.kernel_launches()
    sent0 = sb.bytes_sent
    ms = timed(args.steps, True)                                   # THE timed region: compute + gather
    launches = ctx.kernel_launches() - launches0
    sent = torch.tensor([float(sb.bytes_sent - sent0)], device=dev)
    if dist is not None:
        dist.all_reduce(sent)
    ms_compute = timed(args.steps, False) if world > 1 else ms     # same step without the collective
    # per-kernel durations: the same K steps again with a CUDA-event bracket around every launch (on the launching
    # stream).  The brackets cost a few % of the step, so they stay out of the pass `value` is taken from.
    ctx.kernel_timing_report()
    ctx.set_kernel_timing(True)
    ms_bracketed = timed(args.steps, False)
    hi = sampler.mark() if sampler else 0
    ktimes = ctx.kernel_timing_report()
    ctx.set_kernel_timing(False)
    clocks = sampler.stop(lo, hi + 1) if sampler else None
    value = args.frames * args.steps * W * H / (ms * 1e-3) / 1e6
    # gathered batch: spot-check on the root that what arrived is what a rank computed (frame 0 of the last rank's shard)
    if world > 1:
        probe = sb.local[0].clone() if rank == world - 1 else torch.empty((H, W), dtype=torch.int16, device=dev)
        dist.broadcast(probe.view(torch.uint8), src=world - 1)
        if rank == 0:
            assert torch.equal(sb.full[sb.spans[world - 1][0]], probe), "gathered frame differs from the frame its rank computed"
    zero_conf = float((conf[0, ROI[1]:ROI[1] + ROI[3], ROI[0]:ROI[0] + ROI[2]] == 0).float().mean().item())

    # ---- end to end through the host-buffer entry point (pinned memory), with and without the confidence map ----
    e2e = None
    if not args.no_e2e:
        B = min(args.e2e_frames, nloc)
        hv, hl, hr = view[:B].cpu().pin_memory(), dl[:B].cpu().pin_memory(), dr[:B].cpu().pin_memory()
        ho = torch.empty((B, H, W), dtype=torch.int16).pin_memory()
        hc = torch.empty((B, H, W), dtype=torch.float32).pin_memory()
        esteps = max(2, min(args.steps, 5))

        def e2e_leg(hconf):
            for _ in range(2):
                bf.filter_host(hl, hv, ho, hr, hconf, ROI)
            sync_all()
            x0 = ctx.transfer_bytes()
            t0 = time.perf_counter()
            for _ in range(esteps):
                bf.filter_host(hl, hv, ho, hr, hconf, ROI)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            x1 = ctx.transfer_bytes()
            if dist is not None:
                t = torch.tensor([dt], device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                dt = float(t.item())
            return world * B * esteps * W * H / dt / 1e6, (x1[0] - x0[0]) // esteps, (x1[1] - x0[1]) // esteps
        v0, h2d0, d2h0 = e2e_leg(None)
        assert torch.equal(ho.to(dev), sb.local[:B]), "host-path result differs from the device-path result"
        v1, h2d1, d2h1 = e2e_leg(hc)
        assert torch.equal(hc.to(dev), conf[:B]), "host-path confidence map differs from the device-path one"
        e2e = {"value": v0, "unit": "Mpix/s", "h2d_bytes_per_step": h2d0, "d2h_bytes_per_step": d2h0,
               "with_confidence_map": {"value": v1, "unit": "Mpix/s", "h2d_bytes_per_step": h2d1, "d2h_bytes_per_step": d2h1},
               "frames_per_rank_per_step": B, "steps": esteps,
               "host_input_bytes_per_step": int(hv.numel() + 2 * hl.numel() + 2 * hr.numel()),
               "api": "dwls_filter_batch_host (pinned host buffers, 2-lane copy/compute pipeline; uploads the ROI windows the filter "
                      "reads, bytes counted by the library from the copies it issues); `value` returns the CV_16S maps only, "
                      "`with_confidence_map` also copies back what getConfidenceMap() holds"}
        del hv, hl, hr, ho, hc

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    n_roi = ROI[2] * ROI[3]
    total_alg, per_alg = algorithmic_bytes(n_roi, W * H, W * H, CN)
    peak, peak_src = measured_peak_gbs()
    dom = max(ktimes.items(), key=lambda kv: kv[1][1])
    dom_name, (dom_n, dom_ms) = dom
    frames_per_launch = nloc * args.steps * {"hsolve_kernel": 3, "vsolve_kernel": 2}.get(dom_name, 1) / dom_n
    achieved = per_alg.get(dom_name, 0) * frames_per_launch / (dom_ms / dom_n * 1e-3) / 1e9
    traffic = None           # DRAM bytes of the dominant kernel per launch from the committed ncu --set full capture
    for prof in ("ncu_traffic_r02.json", "ncu_traffic_r01.json"):
        try:
            tk = json.load(open(os.path.join(ROOT, "profiles", prof)))["kernels"].get(dom_name)
            if tk:
                traffic = tk["dram_bytes_per_frame"] * frames_per_launch
                break
        except Exception:
            pass
    ksum = sum(v[1] for v in ktimes.values())
    per_frame_launches = {"hsolve_kernel": 3, "vsolve_kernel": 2}
    roofline = {"bound": "hbm", "kernel": dom_name, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "peak_source": peak_src, "alg_bytes_per_launch": per_alg.get(dom_name, 0) * frames_per_launch,
                "avg_launch_ms": dom_ms / dom_n,
                "timing": "per-kernel CUDA events on the launching stream over a further pass of the same %d steps without the gather "
                          "(%.4f ms/step with the brackets, %.4f without)" % (args.steps, ms_bracketed / args.steps, ms_compute / args.steps),
                "pipeline": {"alg_bytes_per_frame": total_alg, "achieved": total_alg * nloc * args.steps / (ms_compute * 1e-3) / 1e9,
                             "frac": total_alg * nloc * args.steps / (ms_compute * 1e-3) / 1e9 / peak,
                             "what": "whole step of this rank (all kernels, no gather): algorithmic bytes of its frames / its time",
                             "bytes_not_moved_per_frame": 10 * n_roi,
                             "note": "B_alg is SURVEY 8d's fixed count.  The library moves 10 B/ROI pixel less than it: the exported "
                                     "confidence map doubles as plane c and the first row pass forms d*c from the CV_16S disparity, so "
                                     "conf_kernel<left+lrc> does not write the two planes (-8 B/px) and the first hsolve_kernel reads "
                                     "2 instead of 4 B/px of plane 0; time / peak against the bytes actually moved is frac * "
                                     "(1 - bytes_not_moved / alg_bytes)"},
                "kernels": {k: {"launches": n, "total_ms": round(t, 4), "share": round(t / ksum, 4),
                                "us_per_frame": round(t * 1e3 / (nloc * args.steps), 3),
                                "frac": round(per_alg[k] * per_frame_launches.get(k, 1) * nloc * args.steps / (t * 1e-3) / 1e9 / peak, 3) if k in per_alg else None}
                            for k, (n, t) in ktimes.items()}}
    cpu = None               # ---- CPU baseline: the reference's CPU code on a bounded sample, all host threads ----
    if not args.no_cpu:
        from oracle import oracle as O
        threads = O.hardware_threads()
        fps1 = cpu_frames_per_sec(base, 1, threads)
        n_cpu = int(max(2, min(64, args.cpu_seconds * fps1)))
        fps = cpu_frames_per_sec(base, n_cpu, threads)
        cpu = {"value": fps * W * H / 1e6, "unit": "Mpix/s", "cores": threads, "kind": cpu_kind(),
               "sample": "%d frames of the same 1080p workload, reference CPU code on all host threads" % n_cpu,
               "zero_confidence_share_of_roi": round(zero_conf, 4), "note": CPU_NOTE}
    cfg = base_config(args.frames, world, source, chunk)
    run = {"frames_per_gpu": [e - b for b, e in sb.spans], "chunk_plan_rank0": [c for _, c in sb.plan],
           "gather": "none (one rank holds the whole batch)" if world == 1 else
                     {"peer": "CV_16S results of every chunk pushed into rank 0's buffer (CUDA IPC mapping) with peer-to-peer copies on a copy "
                              "stream - copy engines over NVLink, no SMs - while the next chunk's kernels run; one 4-byte NCCL all-reduce per "
                              "step orders the root's reads after the pushes",
                      "root": "CV_16S results of every chunk sent to rank 0 with grouped ncclSend/ncclRecv on NCCL's stream while the next "
                              "chunk's kernels run",
                      "all": "every rank receives every chunk (grouped ncclSend/ncclRecv between all pairs)"}[gather_mode],
           "gather_mode": gather_mode, "gather_fallback": fallback,
           "gather_in_timed_region": world > 1, "zero_confidence_share_of_roi": round(zero_conf, 4)}
    line = {
        "metric": METRIC, "value": value, "unit": "Mpix/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": cfg, "run": run, "gather_in_timed_region": world > 1,
        "compute_only": {"value": args.frames * args.steps * W * H / (ms_compute * 1e-3) / 1e6, "unit": "Mpix/s", "ms_per_step": ms_compute / args.steps,
                         "what": "the same step without the NCCL gather (separate timed pass)"},
        "nvlink_bytes_per_step": int(sent.item()) // args.steps,
        "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks}
    if world == 1 and not args.no_extras:
        line.update(extras_single_gpu(torch, A, ctx, dev, base, peak))
    print(json.dumps(line), file=out, flush=True)
    if dist is not None:
        dist.destroy_process_group()


def _only_the_json_line_on_stdout(fn):
    """The contract is ONE line on stdout.  Libraries write there too (NCCL prints its version banner on stdout when
    NCCL_DEBUG is set on the box), so file descriptor 1 points at stderr while the bench runs and is restored for the line."""
    sys.stdout.flush()
    saved = os.dup(1)
    os.dup2(2, 1)
    real = os.fdopen(saved, "w")
    py_stdout, sys.stdout = sys.stdout, sys.stderr
    try:
        return fn(real)
    finally:
        sys.stdout = py_stdout
        real.flush()
        os.dup2(real.fileno(), 1)


if __name__ == "__main__":
    _only_the_json_line_on_stdout(main)
