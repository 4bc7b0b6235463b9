#!/usr/bin/env python
"""CPU baseline for every BASELINE.json config: the reference's own CPU code (oracle/_ref, all host threads) timed on
this box's host cores - a reported baseline for BASELINE.md section 3, not a target.  Prints one JSON object."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from addingdisparityfiltering_b200 import synth
from oracle import oracle as O
from oracle import ref as R
from tests.stereo_inputs import sample_matching

assert R.available()
threads = O.hardware_threads()
R.set_num_threads(threads)
res = {"host_threads": threads, "kind": "reference (oracle/_ref: the reference's two source files compiled unmodified, -O3)"}


def timeit(fn, seconds=4.0):
    fn()
    n, t0 = 0, time.perf_counter()
    while time.perf_counter() - t0 < seconds or n < 2:
        fn(); n += 1
    return (time.perf_counter() - t0) / n * 1e3


def wls_case(name, left, right, algo, downscale):
    f, dl, dr = sample_matching(left, right, algo, downscale=downscale, max_disp=160)
    lo, ro, to, bo = f._offsets
    roi = (lo, to, dl.shape[1] - lo - ro, dl.shape[0] - to - bo)
    ms = timeit(lambda: R.wls_filter(dl, left, dr, roi, True, 8000.0, 1.5, f.getLRCthresh(), f.getDepthDiscontinuityRadius()))
    res[name] = {"cpu_ms_per_frame": round(ms, 2), "cpu_Mpix_per_s": round(left.shape[0] * left.shape[1] / ms / 1e3, 1), "roi": list(roi)}


l, r, _ = synth.make_stereo_pair(640, 480, seed=42, max_disp=140)
wls_case("configs[0] 640x480 StereoBM, sample default (half-res matching)", l, r, "bm", True)
wls_case("configs[0] 640x480 StereoBM, --no-downscale", l, r, "bm", False)
l, r, _ = synth.make_stereo_pair(1920, 1080, seed=42)
wls_case("configs[1] 1920x1080 StereoSGBM", l, r, "sgbm", False)
wls_case("configs[3] 1920x1080 view, 960x540 StereoBM maps", l, r, "bm", True)
rng = np.random.default_rng(7)
g = synth._gauss_blur(rng.integers(0, 256, (2160, 3840, 3)).astype(np.float32), 2.0)
g = np.clip(np.rint(128 + (g - 128) * 3.0), 0, 255).astype(np.uint8)
src = (rng.random((2160, 3840)) * 255).astype(np.float32)
for name, s in (("32FC1", src), ("8UC3", g)):
    ms = timeit(lambda: R.fgs_filter(g, s, 1000.0, 10.0))
    res["configs[2] FGS 3840x2160 src " + name] = {"cpu_ms_per_frame": round(ms, 2), "cpu_Mpix_per_s": round(3840 * 2160 / ms / 1e3, 1)}
print(json.dumps(res, indent=1))
