#!/usr/bin/env python
"""bench.py — 1080p DisparityWLSFilter::filter throughput on B200 (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--frames B]

One "step" = one pass of the hot path (confidence + LRC + weights + 6 FGS sweeps + normalise) over
a batch of B independent synthetic 1080p stereo frames (BASELINE.json configs[1]: 1920x1080,
SGBM-style geometry: ROI (160,0,1760,1080), discontinuity radius 2, LRC threshold 24, lambda 8000,
sigma 1.5, 3-channel guide).  `value` times the device-resident path (inputs already in HBM, batch
larger than L2); `e2e` times the same batch through the host-buffer C-ABI call with pinned host
memory, H2D and D2H copies inside the timed region.  N > 1: one process per GPU (torchrun), every
rank filters its own B frames, no collective on the data path ("weak" scaling); time = max over ranks.
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
WORKLOAD = "configs[1]: 1920x1080 SGBM-geometry disparities, confidence-aware WLS filter with LRC check"


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
    from addingdisparityfiltering_b200 import synth
    return [synth.make_textured_scene(W, H, CN, seed=42 + i, max_disp=160) for i in range(n_base)]


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


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import oracle as O
    threads = O.hardware_threads()
    base = make_base_frames(2)
    S = args.ref_frames
    for _ in range(args.warmup):
        cpu_frames_per_sec(base, S, threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_frames_per_sec(base, S, threads)
    dt = time.perf_counter() - t0
    v = args.steps * S * W * H / dt / 1e6
    sample = "%d frames per step of the same 1080p workload, %d steps" % (S, args.steps)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "Mpix/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": {"workload": WORKLOAD, "frames_per_step": S, "roi": list(ROI), "lambda": LAMBDA, "sigma_color": SIGMA},
        "cpu_baseline": {"value": v, "unit": "Mpix/s", "cores": threads, "kind": cpu_kind(), "sample": sample},
        "e2e": {"value": v, "unit": "Mpix/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "kind=reference: the reference's fgs_filter.cpp + disparity_filters.cpp compiled unmodified (SSE2 universal intrinsics, "
                "parallel_for_ over all host threads) against the in-repo OpenCV-core stand-in; boxFilter/resize are the stand-in's"}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--frames", type=int, default=32, help="frames per step per GPU")
    ap.add_argument("--group", type=int, default=0, help="frames per kernel launch group (0 = library default)")
    ap.add_argument("--ref-frames", type=int, default=4)
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--interleave", action="store_true", help="A/B switch: sector-interleaved solver planes (dwls_set_plane_interleave)")
    ap.add_argument("--gather", action="store_true", help="all_gather the outputs over NCCL inside the timed region")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline leg (tuning runs only)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import addingdisparityfiltering_b200 as A
    from addingdisparityfiltering_b200.batch import BatchFilter, gather_results

    rank, world, local = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    ctx = A.Context(local)
    if args.group:
        ctx.set_group_frames(args.group)
    if args.interleave:
        ctx.set_plane_interleave(True)
    B = args.frames

    base = make_base_frames(2)
    tb = [[torch.from_numpy(a).to(dev) for a in fr] for fr in base]
    view = torch.stack([torch.roll(tb[(f + rank) % 2][0], shifts=3 * f + rank, dims=1) for f in range(B)]).contiguous()
    dl = torch.stack([torch.roll(tb[(f + rank) % 2][1], shifts=3 * f + rank, dims=1) for f in range(B)]).contiguous()
    dr = torch.stack([torch.roll(tb[(f + rank) % 2][2], shifts=3 * f + rank, dims=1) for f in range(B)]).contiguous()
    out = torch.empty((B, H, W), dtype=torch.int16, device=dev)
    conf = torch.empty((B, H, W), dtype=torch.float32, device=dev)
    bf = BatchFilter(True, LAMBDA, SIGMA, LRC, RADIUS, ctx=ctx)

    def step():
        bf.filter_device(dl, view, out, dr, conf, ROI)
        if args.gather and world > 1:
            gather_results(out)

    def sync_all():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    sync_all()
    sampler = ClockSampler(local) if rank == 0 else None
    time.sleep(0.15 if rank == 0 else 0.0)
    sync_all()
    lo = sampler.mark() if sampler else 0
    launches0 = ctx.kernel_launches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    sync_all()
    ms = e0.elapsed_time(e1)
    launches = ctx.kernel_launches() - launches0
    # per-kernel durations: the same K steps again with a CUDA-event bracket around every launch (on the launching
    # stream).  The brackets cost about 3 % of the step, so they stay out of the pass `value` is taken from.
    ctx.kernel_timing_report()
    ctx.set_kernel_timing(True)
    b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    b0.record()
    for _ in range(args.steps):
        step()
    b1.record()
    sync_all()
    hi = sampler.mark() if sampler else 0
    ms_bracketed = b0.elapsed_time(b1)
    ktimes = ctx.kernel_timing_report()
    ctx.set_kernel_timing(False)
    if dist is not None:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    clocks = sampler.stop(lo, hi + 1) if sampler else None
    value = world * B * args.steps * W * H / (ms * 1e-3) / 1e6

    # ---- end to end through the host-buffer entry point (pinned memory) ----
    e2e = None
    if not args.no_e2e:
        hv, hl, hr = view.cpu().pin_memory(), dl.cpu().pin_memory(), dr.cpu().pin_memory()
        ho = torch.empty((B, H, W), dtype=torch.int16).pin_memory()
        for _ in range(2):
            bf.filter_host(hl, hv, ho, hr, None, ROI)
        sync_all()
        x0 = ctx.transfer_bytes()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            bf.filter_host(hl, hv, ho, hr, None, ROI)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        x1 = ctx.transfer_bytes()
        assert torch.equal(ho.to(dev), out), "host-path result differs from the device-path result"
        if dist is not None:
            t = torch.tensor([dt], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        e2e = {"value": world * B * args.steps * W * H / dt / 1e6, "unit": "Mpix/s",
               "h2d_bytes_per_step": (x1[0] - x0[0]) // args.steps, "d2h_bytes_per_step": (x1[1] - x0[1]) // args.steps,
               "host_input_bytes_per_step": int(hv.numel() + 2 * hl.numel() + 2 * hr.numel()),
               "api": "dwls_filter_batch_host (pinned host buffers, 2-lane copy/compute pipeline; uploads the ROI windows the filter "
                      "reads, bytes counted by the library from the copies it issues)"}

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return
    n_roi = ROI[2] * ROI[3]
    total_alg, per_alg = algorithmic_bytes(n_roi, W * H, W * H, CN)
    peak, peak_src = measured_peak_gbs()
    dom = max(ktimes.items(), key=lambda kv: kv[1][1])
    dom_name, (dom_n, dom_ms) = dom
    frames_per_launch = B * args.steps * {"hsolve_kernel": 3, "vsolve_kernel": 2}.get(dom_name, 1) / dom_n
    achieved = per_alg.get(dom_name, 0) * frames_per_launch / (dom_ms / dom_n * 1e-3) / 1e9
    # DRAM bytes of the dominant kernel per launch from the committed ncu --set full capture (profiles/ncu_traffic_r01.json)
    traffic = None
    try:
        tk = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic_r01.json")))["kernels"].get(dom_name)
        if tk:
            traffic = tk["dram_bytes_per_frame"] * frames_per_launch
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": dom_name, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "peak_source": peak_src, "alg_bytes_per_launch": per_alg.get(dom_name, 0) * frames_per_launch,
                "avg_launch_ms": dom_ms / dom_n,
                "timing": "per-kernel CUDA events on the launching stream over a second pass of the same %d steps (%.4f ms/step with the "
                          "brackets, %.4f without)" % (args.steps, ms_bracketed / args.steps, ms / args.steps),
                "pipeline": {"alg_bytes_per_frame": total_alg, "achieved": total_alg * B * args.steps / (ms * 1e-3) / 1e9,
                             "frac": total_alg * B * args.steps / (ms * 1e-3) / 1e9 / peak},
                "kernels": {k: {"launches": n, "total_ms": round(t, 4), "share": round(t / sum(v[1] for v in ktimes.values()), 4)}
                            for k, (n, t) in ktimes.items()}}
    # ---- CPU baseline: the oracle on a bounded sample, all host threads ----
    cpu = None
    if not args.no_cpu:
        from oracle import oracle as O
        threads = O.hardware_threads()
        fps1 = cpu_frames_per_sec(base, 1, threads)
        n_cpu = int(max(2, min(64, args.cpu_seconds * fps1)))
        fps = cpu_frames_per_sec(base, n_cpu, threads)
        cpu = {"value": fps * W * H / 1e6, "unit": "Mpix/s", "cores": threads, "kind": cpu_kind(),
               "sample": "%d frames of the same 1080p workload, reference CPU code on all host threads" % n_cpu}
    print(json.dumps({
        "metric": METRIC, "value": value, "unit": "Mpix/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "frames_per_step_per_gpu": B, "roi": list(ROI), "lambda": LAMBDA, "sigma_color": SIGMA,
                   "disc_radius": RADIUS, "lrc_thresh": LRC, "guide_channels": CN, "confidence_export": True,
                   "l2": "batch working set (%.0f MB inputs+outputs per step) larger than the 126 MB L2; no flush needed" % (B * (2 * 2 * W * H + CN * W * H + 6 * W * H) / 1e6),
                   "gather_in_timed_region": bool(args.gather and world > 1)},
        "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches, "clocks": clocks}))
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
