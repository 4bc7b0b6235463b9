"""The numpy restatement of cv::StereoBM (oracle/stereobm_model.py) pinned against the real thing, cv2.StereoBM.compute:
bit-exact CV_16S for the matcher pair the WLS factories configure and for OpenCV's default texture / uniqueness tests."""
import numpy as np
import pytest

from addingdisparityfiltering_b200 import synth
from oracle import stereobm_model as M

BM_CASES = [  # (numDisparities, blockSize, minDisparity, textureThreshold, uniquenessRatio)
    (80, 7, 0, 0, 0), (80, 7, -79, 0, 0),          # sample default: left matcher / createRightMatcher (disparity_filters.cpp:424)
    (64, 15, 0, 0, 0), (64, 15, -63, 0, 0),        # --no-downscale window
    (48, 9, 0, 10, 15), (32, 5, -16, 10, 15), (80, 21, 0, 0, 0), (16, 5, -3, 0, 7),
]


def gray_pair(w, h, seed, max_disp):
    cv2 = pytest.importorskip("cv2")
    left, right, _ = synth.make_stereo_pair(w, h, seed=seed, bg_disp=min(16, max(1, max_disp // 3)), min_disp=min(24, max(2, max_disp // 2)), max_disp=max_disp)
    return cv2.cvtColor(left, cv2.COLOR_BGR2GRAY), cv2.cvtColor(right, cv2.COLOR_BGR2GRAY), left, right


def cv2_bm(nd, ws, md, tex, un):
    import cv2
    m = cv2.StereoBM_create(nd, ws)
    m.setMinDisparity(md); m.setTextureThreshold(tex); m.setUniquenessRatio(un); m.setDisp12MaxDiff(1000000); m.setSpeckleWindowSize(0)
    return m


@pytest.mark.parametrize("case", BM_CASES)
@pytest.mark.parametrize("size", [(320, 240), (203, 97)])
def test_numpy_model_equals_cv2_stereobm(case, size):
    nd, ws, md, tex, un = case
    gl, gr, _, _ = gray_pair(size[0], size[1], seed=1 + ws, max_disp=min(60, nd - 4))
    a, b = (gl, gr) if md >= 0 else (gr, gl)
    m = cv2_bm(nd, ws, md, tex, un)
    want = m.compute(a, b)
    got = M.stereo_bm(a, b, nd, ws, md, m.getPreFilterCap(), tex, un)
    assert (got == want).all(), int((got != want).sum())


def test_gray_conversion_and_half_resize_formulas():
    """The two host-side steps of the sample the device path also covers: cvtColor BGR2GRAY and resize by 0.5."""
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(0)
    a = rng.integers(0, 256, (48, 64, 3)).astype(np.uint8)
    b, g, r = [a[..., i].astype(np.int64) for i in range(3)]
    assert ((3735 * b + 19235 * g + 9798 * r + 16384) >> 15 == cv2.cvtColor(a, cv2.COLOR_BGR2GRAY)).all()
    s = a.astype(np.int64)
    half = (s[0::2, 0::2] + s[1::2, 0::2] + s[0::2, 1::2] + s[1::2, 1::2] + 2) >> 2
    assert (half == cv2.resize(a, None, fx=0.5, fy=0.5)).all()
