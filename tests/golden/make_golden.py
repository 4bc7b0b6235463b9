"""Generates tests/golden/*.npz from the REFERENCE itself: oracle/_ref/libref_ximgproc.so is the
reference's fgs_filter.cpp + disparity_filters.cpp compiled unmodified from /root/reference
(`make -C oracle ref`).  Run in the build container (needs /root/reference to have built _ref):

    python tests/golden/make_golden.py

Inputs are regenerated from seeds by the tests (addingdisparityfiltering_b200.synth); only the
reference's outputs are stored, as compressed arrays.  num_threads is fixed to 4 (the reference's
SIMD/scalar association depends on its stripe count, test_fgs_filter.cpp:109-151)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from addingdisparityfiltering_b200 import synth  # noqa: E402
from oracle import ref as R  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
THREADS = 4

FGS_CASES = [  # (name, w, h, guide channels, src kind, lambda, sigma)
    ("fgs_127x61_c1_f32", 127, 61, 1, "f32", 8000.0, 1.5),
    ("fgs_160x120_c3_f32", 160, 120, 3, "f32", 1000.0, 10.0),
    ("fgs_96x64_c3_s16c3", 96, 64, 3, "s16c3", 5000.0, 20.0),
    ("fgs_96x64_c1_u8c4", 96, 64, 1, "u8c4", 500.0, 1.5),
]
WLS_CASES = [  # (name, w, h, channels, use_conf, half_res, radius, seed)
    ("wls_160x120_conf", 160, 120, 3, True, False, 3, 11),
    ("wls_160x120_noconf", 160, 120, 3, False, False, 3, 12),
    ("wls_127x61_conf_c1", 127, 61, 1, True, False, 2, 13),
    ("wls_160x120_conf_half", 160, 120, 3, True, True, 3, 14),
    ("wls_160x120_noconf_half", 160, 120, 3, False, True, 3, 15),
]


def fgs_inputs(w, h, cn, kind, seed):
    rng = np.random.default_rng(seed)
    g = rng.integers(0, 256, (h, w, cn)).astype(np.float32)
    g = ((g + np.roll(g, 1, 0) + np.roll(g, 1, 1)) / 3).astype(np.uint8)
    g = g[:, :, 0] if cn == 1 else g
    if kind == "f32":
        src = (rng.random((h, w)) * 255).astype(np.float32)
    elif kind == "s16c3":
        src = rng.integers(-32767, 32767, (h, w, 3)).astype(np.int16)
    else:
        src = rng.integers(0, 256, (h, w, 4)).astype(np.uint8)
    return g, src


def wls_inputs(w, h, cn, half, seed):
    view, dl, dr, roi = synth.make_artificial_example(w, h, cn, seed=seed)
    if half:   # decimate + halve, like the reference sample's downscale mode (values /2, ROI /2)
        dl = (dl[::2, ::2] // 2).astype(np.int16)
        dr = (dr[::2, ::2] // 2).astype(np.int16)
        roi = tuple(v // 2 for v in roi)
    return view, np.ascontiguousarray(dl), np.ascontiguousarray(dr), roi


def main():
    assert R.available(), "build oracle/_ref first: make -C oracle ref"
    R.set_num_threads(THREADS)
    for i, (name, w, h, cn, kind, lam, sigma) in enumerate(FGS_CASES):
        g, src = fgs_inputs(w, h, cn, kind, 100 + i)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), out=R.fgs_filter(g, src, lam, sigma))
    for name, w, h, cn, conf, half, radius, seed in WLS_CASES:
        view, dl, dr, roi = wls_inputs(w, h, cn, half, seed)
        r = R.wls_filter(dl, view, dr, roi, conf, 8000.0, 1.5, 24, radius)
        assert r["rc"] == 0
        np.savez_compressed(os.path.join(HERE, name + ".npz"), out=r["out"], conf=r["conf"] if conf else np.zeros(0, np.float32), roi=np.array(r["roi"]))
    print("golden vectors written to", HERE)


if __name__ == "__main__":
    main()
