"""numpy model of cv::StereoBM::compute — TEST INFRASTRUCTURE (CPU restatement for the stereo-matching row).

StereoBM lives in OpenCV's calib3d, which is NOT part of /root/reference (the reference only calls it:
samples/disparity_filtering.cpp:148-160, parameter contract src/disparity_filters.cpp:389-403, 419-428).  This restates
the published algorithm - PREFILTER_XSOBEL with cap, SAD over a square window, first-minimum winner-takes-all, optional
texture / uniqueness tests, parabola sub-pixel fit, valid-disparity rectangle - and is pinned bit for bit against
cv2.StereoBM.compute (OpenCV 4.13 on the box) in tests/test_stereo_model.py; the CUDA matcher is compared with cv2 itself.
minDisparity <= 0 only: for minDisparity > 0 OpenCV writes its last `minDisparity` computed columns past the end of each row
(they land in the first columns of the next row), which is not something to reproduce.
"""
import numpy as np

def prefilter_xsobel(src, ftzero):
    h, w = src.shape
    s = src.astype(np.int32)
    dst = np.full((h, w), ftzero, np.uint8)
    d = np.zeros((h, w), np.int32)
    d[:, 1:w-1] = s[:, 2:] - s[:, :-2]
    def row(i): return d[i]
    for y in range(0, h - 1, 2):
        r1 = y
        r0 = y - 1 if y > 0 else (y + 1 if h > 1 else y)
        r2 = y + 1 if y < h - 1 else (y - 1 if h > 1 else y)
        r3 = y + 2 if y < h - 2 else y
        v0 = d[r0] + 2 * d[r1] + d[r2]
        v1 = d[r1] + 2 * d[r2] + d[r3]
        dst[y, 1:w-1] = (np.clip(v0, -ftzero, ftzero) + ftzero)[1:w-1]
        dst[y+1, 1:w-1] = (np.clip(v1, -ftzero, ftzero) + ftzero)[1:w-1]
    return dst

def stereo_bm(left, right, ndisp, wsz, mindisp=0, cap=31, texture=0, uniq=0):
    h, w = left.shape
    L = prefilter_xsobel(left, cap).astype(np.int32); R = prefilter_xsobel(right, cap).astype(np.int32)
    w2 = wsz // 2
    n = ndisp - 1 + mindisp
    lofs, rofs = max(n, 0), -min(n, 0)
    width1 = w - rofs - ndisp + 1
    FILTERED = (mindisp - 1) * 16
    out = np.full((h, w), FILTERED, np.int16)
    if lofs >= w or rofs >= w or width1 < 1: return out
    xr = np.arange(-w2, width1 + w2)             # relative window columns
    colL = np.clip(xr, -lofs, w - lofs - 1) + lofs
    colR0 = np.clip(xr, -rofs, w - rofs - ndisp) + rofs
    y0, y1 = w2, h - w2                          # valid rows
    if y1 <= y0: return out
    sad = np.zeros((ndisp, y1 - y0, width1), np.int64)
    def box(a):    # a: [h, len(xr)] -> window sums for valid rows / rel x in [0,width1)
        c = np.cumsum(np.pad(a, ((1, 0), (1, 0))), 0).cumsum(1)
        ys = np.arange(y0, y1); xs = np.arange(0, width1)
        A = c[np.ix_(ys + w2 + 1, xs + 2 * w2 + 1)]; B = c[np.ix_(ys - w2, xs + 2 * w2 + 1)]
        Cc = c[np.ix_(ys + w2 + 1, xs)]; D = c[np.ix_(ys - w2, xs)]
        return A - B - Cc + D
    for d in range(ndisp):
        diff = np.abs(L[:, colL] - R[:, colR0 + d])
        sad[d] = box(diff)
    mind = sad.argmin(0)                         # first minimum
    ms = np.take_along_axis(sad, mind[None], 0)[0]
    sp = np.concatenate([sad[1:2], sad, sad[ndisp-2:ndisp-1]], 0)   # sad[-1]=sad[1], sad[ndisp]=sad[ndisp-2]
    p = np.take_along_axis(sp, (mind + 2)[None], 0)[0]; nn = np.take_along_axis(sp, mind[None], 0)[0]
    dd = p + nn - 2 * ms + np.abs(p - nn)
    num = (p - nn) * 256
    q = np.where(dd != 0, np.trunc(num / np.where(dd == 0, 1, dd)).astype(np.int64), 0)   # C integer division truncates toward zero
    val = ((ndisp - mind - 1 + mindisp) * 256 + q + 15) >> 4
    res = val.astype(np.int16)
    if texture > 0:
        t = box(np.abs(L[:, colL] - cap))
        res[t < texture] = FILTERED
    if uniq > 0:
        thresh = ms + ms * uniq // 100
        dgrid = np.arange(ndisp)[:, None, None]
        bad = ((dgrid < mind[None] - 1) | (dgrid > mind[None] + 1)) & (sad <= thresh[None])
        res[bad.any(0)] = FILTERED
    out[y0:y1, lofs:min(lofs + width1, w)] = res[:, :max(0, min(lofs + width1, w) - lofs)]   # minDisparity > 0: the last columns fall off the row
    maxd = mindisp + ndisp - 1
    xmin, xmax = max(0, maxd) + w2, min(w, w - mindisp) - w2      # getValidDisparityROI
    out[:, :max(xmin, 0)] = FILTERED
    out[:, max(xmax, 0):] = FILTERED
    return out

