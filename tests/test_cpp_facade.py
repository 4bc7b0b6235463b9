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
    for name, a in (("dl", dl), ("dr", dr), ("view", view), ("src", src)):
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
