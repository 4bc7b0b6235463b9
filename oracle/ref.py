"""ctypes front-end of oracle/_ref/libref_ximgproc.so — the reference's own fgs_filter.cpp and
disparity_filters.cpp compiled unmodified from /root/reference against the OpenCV-core stand-in in
oracle/shim/ (recipe: `make -C oracle ref`).  TEST INFRASTRUCTURE: used to pin the CPU oracle, to
generate tests/golden/, and as the CPU baseline ("kind": "reference") of bench.py.  The prebuilt
.so travels to the GPU box; /root/reference itself is never read at run time."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(_HERE, "_ref", "libref_ximgproc.so")
DEPTH = {np.dtype(np.uint8): 0, np.dtype(np.int16): 3, np.dtype(np.float32): 5}
_LIB = None


def available():
    return os.path.exists(SO)


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(SO)
    return _LIB


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def set_num_threads(n):
    return int(lib().ref_set_num_threads(int(n)))


def fgs_filter(guide, src, lam, sigma, lambda_att=0.25, num_iter=3):
    guide, src = np.ascontiguousarray(guide), np.ascontiguousarray(src)
    h, w = guide.shape[:2]
    gcn = 1 if guide.ndim == 2 else guide.shape[2]
    scn = 1 if src.ndim == 2 else src.shape[2]
    dst = np.empty_like(src)
    rc = lib().ref_fgs_filter(_p(guide), C.c_longlong(guide.strides[0]), w, h, gcn, _p(src), C.c_longlong(src.strides[0]), DEPTH[src.dtype], scn,
                              _p(dst), C.c_longlong(dst.strides[0]), C.c_double(lam), C.c_double(sigma), C.c_double(lambda_att), int(num_iter))
    if rc:
        raise ValueError("reference fastGlobalSmootherFilter raised (code %d)" % rc)
    return dst


def wls_filter(disp_left, view, disp_right, roi, use_confidence=True, lam=8000.0, sigma=1.0, lrc_thresh=24, disc_radius=5):
    dl = np.ascontiguousarray(disp_left, np.int16)
    dr = None if disp_right is None else np.ascontiguousarray(disp_right, np.int16)
    view = np.ascontiguousarray(view, np.uint8)
    dh, dw = dl.shape
    H, W = view.shape[:2]
    vcn = 1 if view.ndim == 2 else view.shape[2]
    out = np.empty((H, W), np.int16)
    conf = np.zeros((H, W), np.float32) if use_confidence else None
    roi_in = (C.c_int * 4)(*[int(v) for v in roi])
    roi_out = (C.c_int * 4)()
    rc = lib().ref_wls_filter(_p(dl), C.c_longlong(dl.strides[0]), _p(dr) if dr is not None else None, C.c_longlong(dr.strides[0] if dr is not None else 0),
                              dw, dh, _p(view), C.c_longlong(view.strides[0]), W, H, vcn, int(bool(use_confidence)), C.c_double(lam), C.c_double(sigma),
                              int(lrc_thresh), int(disc_radius), roi_in, _p(out), C.c_longlong(out.strides[0]),
                              _p(conf) if conf is not None else None, C.c_longlong(W * 4), roi_out)
    return dict(out=out, conf=conf, roi=tuple(roi_out), rc=rc)


def factory_geometry(kind, min_disp, num_disp, block, dw, dh):
    """(ROI, discontinuity radius, right matcher's minDisparity) from createDisparityWLSFilter / createRightMatcher."""
    roi = (C.c_int * 4)()
    rad, rmin = C.c_int(), C.c_int()
    rc = lib().ref_factory_roi(0 if kind == "bm" else 1, min_disp, num_disp, block, dw, dh, roi, C.byref(rad), C.byref(rmin))
    assert rc == 0
    return tuple(roi), rad.value, rmin.value


def read_gt(path, cap=1 << 24):
    """readGT of the reference (disparity_filters.cpp:462-495) through the stand-in imread (binary PGM / PPM files)."""
    dst = np.empty(cap, np.int16)
    rows, cols = C.c_int(), C.c_int()
    rc = lib().ref_read_gt(path.encode(), _p(dst), int(cap), C.byref(rows), C.byref(cols))
    return rc, (dst[:rows.value * cols.value].reshape(rows.value, cols.value).copy() if rc == 0 else None)


def compute_mse(gt, src, roi):
    gt, src = np.ascontiguousarray(gt, np.int16), np.ascontiguousarray(src, np.int16)
    out = C.c_double()
    rc = lib().ref_compute_mse(_p(gt), _p(src), gt.shape[1], gt.shape[0], (C.c_int * 4)(*[int(v) for v in roi]), C.byref(out))
    assert rc == 0
    return out.value


def bad_pixel_percent(gt, src, roi, thresh=24):
    gt, src = np.ascontiguousarray(gt, np.int16), np.ascontiguousarray(src, np.int16)
    out = C.c_double()
    rc = lib().ref_bad_pixel_percent(_p(gt), _p(src), gt.shape[1], gt.shape[0], (C.c_int * 4)(*[int(v) for v in roi]), int(thresh), C.byref(out))
    assert rc == 0
    return out.value


def disparity_vis(src, scale=1.0):
    src = np.ascontiguousarray(src, np.int16)
    dst = np.empty(src.shape, np.uint8)
    rc = lib().ref_disparity_vis(_p(src), src.shape[1], src.shape[0], C.c_double(scale), _p(dst))
    assert rc == 0
    return dst
