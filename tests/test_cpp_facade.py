"""The C++ facade (cv::ximgproc names on cv::Mat) compiled with g++ against the C-ABI library:
compile + link on CPU; on the GPU box the binary runs and its outputs are checked against the oracle."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "addingdisparityfiltering_b200")


def _build(tmp_path):
    from addingdisparityfiltering_b200 import _lib
    _lib.build()
    exe = str(tmp_path / "facade_test")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-DDWLS_USE_MINI_CV", os.path.join(ROOT, "tests", "cpp", "facade_test.cpp"), "-o", exe,
                           "-L" + PKG, "-ldwls_b200", "-Wl,-rpath," + PKG, "-L/usr/local/cuda/lib64", "-Wl,-rpath,/usr/local/cuda/lib64", "-lcudart"])
    return exe


def test_facade_compiles_and_links(tmp_path):
    assert os.path.exists(_build(tmp_path))


@pytest.mark.gpu
def test_facade_matches_oracle(tmp_path):
    from addingdisparityfiltering_b200 import synth
    from oracle import oracle as O
    from tests.parity import assert_s16_parity
    exe = _build(tmp_path)
    w, h, cn = 320, 240, 3
    view, dl, dr, roi = synth.make_artificial_example(w, h, cn, seed=21)
    src = (np.random.default_rng(1).random((h, w)) * 255).astype(np.float32)
    import cv2
    left, right, _ = synth.make_stereo_pair(w, h, seed=5, bg_disp=6, min_disp=10, max_disp=40)
    gl, gr = cv2.cvtColor(left, cv2.COLOR_BGR2GRAY), cv2.cvtColor(right, cv2.COLOR_BGR2GRAY)
    for name, a in (("dl", dl), ("dr", dr), ("view", view), ("src", src), ("gl", gl), ("gr", gr)):
        a.tofile(str(tmp_path / (name + ".bin")))
    lam, sigma, radius = 8000.0, 1.5, 3
    res = subprocess.run([exe, str(tmp_path), str(w), str(h), str(w), str(h), str(cn)] + [str(v) for v in roi] + [str(radius), str(lam), str(sigma)],
                         capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout + res.stderr
    lines = dict((l.split()[0], l.split()[1:]) for l in res.stdout.strip().splitlines())
    ref = O.wls_filter(dl, view, dr, roi=roi, lam=lam, sigma=sigma, disc_radius=radius)
    assert tuple(int(v) for v in lines["ROI"]) == ref["roi"]
    assert lines["FACTORY"] == ["radius", "3", "right_min", "-79", "disp12", "1000000"]
    assert tuple(int(v) for v in lines["ROI2"]) == (83, 3, w - 83 - 3, h - 6)
    assert lines["CAUGHT"] == ["3"]
    out = np.fromfile(str(tmp_path / "out.bin"), np.int16).reshape(h, w)
    conf = np.fromfile(str(tmp_path / "conf.bin"), np.float32).reshape(h, w)
    fgs = np.fromfile(str(tmp_path / "fgs.bin"), np.float32).reshape(h, w)
    assert_s16_parity(out, ref["out"], roi)
    assert np.abs(conf - ref["conf"]).max() <= 1e-4
    assert np.abs(fgs - O.fgs_filter(view, src, lam, sigma)).max() <= 255 * 1e-3
    # the device matcher through the C++ facade == cv2.StereoBM, and its filtered map == the oracle on cv2's maps
    import addingdisparityfiltering_b200 as A
    lm = cv2.StereoBM_create(48, 7)
    f = A.createDisparityWLSFilter(lm)
    rm = A.createRightMatcher(lm)
    want_l, want_r = lm.compute(gl, gr), rm.compute(gr, gl)
    bm_l = np.fromfile(str(tmp_path / "bm_l.bin"), np.int16).reshape(h, w)
    bm_r = np.fromfile(str(tmp_path / "bm_r.bin"), np.int16).reshape(h, w)
    assert (bm_l == want_l).all() and (bm_r == want_r).all()
    ref3 = O.wls_filter(want_l, view, want_r, lam=lam, sigma=sigma, disc_radius=f.getDepthDiscontinuityRadius(), offsets=f._offsets)
    assert tuple(int(v) for v in lines["ROI3"]) == ref3["roi"]
    assert_s16_parity(np.fromfile(str(tmp_path / "bm_out.bin"), np.int16).reshape(h, w), ref3["out"], ref3["roi"])


def test_quality_helpers_match_the_reference_bit_for_bit(tmp_path):
    """readGT / computeMSE / computeBadPixelPercent / getDisparityVis (disparity_filter.hpp:163-204) of the C++ facade and of
    the Python mirror against the REFERENCE's own implementations (disparity_filters.cpp:460-556, compiled into
    oracle/_ref).  Host code on both sides: runs without a GPU."""
    cv2 = pytest.importorskip("cv2")
    import addingdisparityfiltering_b200 as A
    from addingdisparityfiltering_b200 import _lib
    from oracle import ref as R
    if not R.available():
        pytest.skip("oracle/_ref is not built")
    _lib.build()
    exe = str(tmp_path / "helpers_test")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-DDWLS_USE_MINI_CV", os.path.join(ROOT, "tests", "cpp", "helpers_test.cpp"), "-o", exe,
                           "-L" + PKG, "-ldwls_b200", "-Wl,-rpath," + PKG, "-L/usr/local/cuda/lib64", "-Wl,-rpath,/usr/local/cuda/lib64", "-lcudart"])
    rng = np.random.default_rng(5)
    w, h = 70, 50
    gt = rng.integers(-300, 3000, (h, w)).astype(np.int16)
    src = (gt + rng.integers(-60, 60, gt.shape)).astype(np.int16)
    src[::7, ::5] = rng.integers(-32768, 32768, src[::7, ::5].shape)          # saturating visualisation values
    gt[5:12, 8:20] = 16320                                                     # UNKNOWN_DISPARITY (:460)
    src[3, 3] = 16320
    roi, thresh, scale = (6, 4, 55, 40), 17, 2.37
    g = rng.integers(0, 256, (31, 43)).astype(np.uint8)
    g[4, 7] = 0
    c = rng.integers(0, 256, (29, 37, 3)).astype(np.uint8)
    gt.tofile(str(tmp_path / "gt.bin")); src.tofile(str(tmp_path / "src.bin"))
    for name, img in (("g.pgm", g), ("c.ppm", c), ("g.png", g), ("c.png", c)):
        assert cv2.imwrite(str(tmp_path / name), img)
    res = subprocess.run([exe, str(tmp_path), str(w), str(h)] + [str(v) for v in roi] + [str(thresh), repr(scale)], capture_output=True, text=True, timeout=120)
    assert res.returncode == 0, res.stdout + res.stderr
    lines = dict((l.split()[0], l.split()[1:]) for l in res.stdout.strip().splitlines())
    # the reference's numbers
    mse, bad, bad24 = R.compute_mse(gt, src, roi), R.bad_pixel_percent(gt, src, roi, thresh), R.bad_pixel_percent(gt, src, roi)
    vis = R.disparity_vis(src, scale)
    rc1, gt1 = R.read_gt(str(tmp_path / "g.pgm"))
    rc3, gt3 = R.read_gt(str(tmp_path / "c.ppm"))
    assert rc1 == 0 and rc3 == 0 and R.read_gt(str(tmp_path / "missing.pgm"))[0] == 1
    # C++ facade == reference
    assert float(lines["MSE"][0]) == mse and float(lines["BAD"][0]) == bad and float(lines["BAD_DEFAULT"][0]) == bad24
    assert lines["READ"] == ["0", "0", "1"] and lines["SIZE"] == ["43", "31", "37", "29"] and lines["CAUGHT"] == ["7"]
    assert (np.fromfile(str(tmp_path / "vis.bin"), np.uint8).reshape(h, w) == vis).all()
    assert (np.fromfile(str(tmp_path / "gt_pgm.bin"), np.int16).reshape(31, 43) == gt1).all()
    assert (np.fromfile(str(tmp_path / "gt_ppm.bin"), np.int16).reshape(29, 37) == gt3).all()
    # Python mirror == reference; readGT on PNGs written by cv2.imwrite, both the 1- and the 3-channel branch (:466-493)
    assert A.computeMSE(gt, src, roi) == mse and A.computeBadPixelPercent(gt, src, roi, thresh) == bad
    assert A.computeBadPixelPercent(gt, src, roi) == bad24
    assert (A.getDisparityVis(src, scale) == vis).all() and (A.getDisparityVis(src) == R.disparity_vis(src, 1.0)).all()
    s1, p1 = A.readGT(str(tmp_path / "g.png"))
    s3, p3 = A.readGT(str(tmp_path / "c.png"))
    assert s1 == 0 and s3 == 0 and p1.dtype == np.int16 and (p1 == gt1).all() and (p3 == gt3).all()
    assert (p1[4, 7], p1[0, 0]) == (16320, 16 * int(g[0, 0]))
    assert A.readGT(str(tmp_path / "missing.png"))[0] == 1
