"""Inputs produced by REAL stereo matchers, the way the reference's sample builds them
(modules/ximgproc/samples/disparity_filtering.cpp:128-190): cv2.StereoBM / cv2.StereoSGBM configured by
createDisparityWLSFilter / createRightMatcher (disparity_filters.cpp:385-449).  Test infrastructure."""
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def sample_matching(left, right, algo="bm", downscale=True, max_disp=160, wsize=None, ctx=None):
    """The `wls_conf` branch of the sample up to (not including) the filter call.
    Returns (wls_filter, left_disp, right_disp): the filter object comes from createDisparityWLSFilter(left_matcher)."""
    import cv2
    import addingdisparityfiltering_b200 as A
    if wsize is None:                                                   # sample lines 62-72
        wsize = 3 if algo == "sgbm" else (7 if downscale else 15)
    lm, rm = left, right
    if downscale:                                                       # 135-139
        max_disp //= 2
        if max_disp % 16:
            max_disp += 16 - (max_disp % 16)
        lm = cv2.resize(left, None, fx=0.5, fy=0.5)
        rm = cv2.resize(right, None, fx=0.5, fy=0.5)
    if algo == "bm":                                                    # 148-160
        left_matcher = cv2.StereoBM_create(max_disp, wsize)
        wls = A.createDisparityWLSFilter(left_matcher, ctx=ctx)
        right_matcher = A.createRightMatcher(left_matcher)
        lm, rm = cv2.cvtColor(lm, cv2.COLOR_BGR2GRAY), cv2.cvtColor(rm, cv2.COLOR_BGR2GRAY)
    else:                                                               # 162-175
        left_matcher = cv2.StereoSGBM_create(0, max_disp, wsize)
        left_matcher.setP1(24 * wsize * wsize)
        left_matcher.setP2(96 * wsize * wsize)
        left_matcher.setPreFilterCap(63)
        left_matcher.setMode(cv2.STEREO_SGBM_MODE_SGBM_3WAY)
        wls = A.createDisparityWLSFilter(left_matcher, ctx=ctx)
        right_matcher = A.createRightMatcher(left_matcher)
    left_disp = left_matcher.compute(lm, rm)
    right_disp = right_matcher.compute(rm, lm)
    return wls, left_disp, right_disp


def ambush_fixture():
    """tests/golden/ambush_5_bm.npz (made by tests/golden/make_golden_real.py from the reference's tutorial pair)."""
    import cv2
    z = np.load(os.path.join(HERE, "golden", "ambush_5_bm.npz"))
    left = cv2.imdecode(z["left_jpg"], cv2.IMREAD_COLOR)
    right = cv2.imdecode(z["right_jpg"], cv2.IMREAD_COLOR)
    return z, left, right
