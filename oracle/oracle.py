"""ctypes front-end of the CPU oracle (oracle/liboracle.so) — TEST INFRASTRUCTURE ONLY.

Importable from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs.  The product package never imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

DEPTH = {np.dtype(np.uint8): 0, np.dtype(np.int16): 3, np.dtype(np.float32): 5}


class OrcWlsParams(C.Structure):
    _fields_ = [("lambda_", C.c_double), ("sigma_color", C.c_double),
                ("lrc_thresh", C.c_int), ("disc_radius", C.c_int),
                ("left_offset", C.c_int), ("right_offset", C.c_int), ("top_offset", C.c_int), ("bottom_offset", C.c_int),
                ("use_confidence", C.c_int),
                ("roi_x", C.c_int), ("roi_y", C.c_int), ("roi_w", C.c_int), ("roi_h", C.c_int),
                ("assoc", C.c_int), ("nstripes", C.c_int), ("threaded", C.c_int)]


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    src = os.path.join(_HERE, "dwls_oracle.cpp")
    if force or not os.path.exists(so) or (os.path.exists(src) and os.path.getmtime(src) > os.path.getmtime(so)):
        subprocess.check_call(["make", "-s", "-C", _HERE, "liboracle.so"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        _LIB.orc_hardware_threads.restype = C.c_int
    return _LIB


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def hardware_threads():
    return int(lib().orc_hardware_threads())


def fgs_lut(sigma, n=3 * 256 * 256):
    out = np.empty(n, np.float32)
    lib().orc_fgs_lut(C.c_float(sigma), _p(out), C.c_int(n))
    return out


def fgs_weights(guide, sigma):
    guide = np.ascontiguousarray(guide)
    h, w = guide.shape[:2]
    cn = 1 if guide.ndim == 2 else guide.shape[2]
    chor = np.empty((h, w), np.float32)
    cvert = np.empty((h, w), np.float32)
    rc = lib().orc_fgs_weights(_p(guide), C.c_int64(guide.strides[0]), w, h, cn, C.c_double(sigma), _p(chor), _p(cvert))
    assert rc == 0
    return chor, cvert


def fgs_filter(guide, src, lam, sigma, lambda_att=0.25, num_iter=3, assoc=0, nstripes=1, threaded=False):
    """fastGlobalSmootherFilter(guide, src, dst, lambda, sigma_color, lambda_attenuation, num_iter)."""
    guide = np.ascontiguousarray(guide)
    src = np.ascontiguousarray(src)
    h, w = guide.shape[:2]
    gcn = 1 if guide.ndim == 2 else guide.shape[2]
    scn = 1 if src.ndim == 2 else src.shape[2]
    if src.shape[:2] != (h, w):
        raise ValueError("Size of the filtered image must be equal to the size of the guide image")
    dst = np.empty_like(src)
    rc = lib().orc_fgs_filter(_p(guide), C.c_int64(guide.strides[0]), w, h, gcn,
                              _p(src), C.c_int64(src.strides[0]), DEPTH[src.dtype], scn,
                              _p(dst), C.c_int64(dst.strides[0]),
                              C.c_double(lam), C.c_double(sigma), C.c_double(lambda_att), int(num_iter),
                              int(assoc), int(nstripes), int(bool(threaded)))
    if rc != 0:
        raise ValueError("oracle fgs_filter: bad argument")
    return dst


def box_mean_16s(src, r):
    src = np.ascontiguousarray(src, np.int16)
    h, w = src.shape
    out = np.empty((h, w), np.float32)
    lib().orc_box_mean_16s(_p(src), C.c_int64(w), w, h, int(r), _p(out))
    return out


def box_meansq_16s(src, r):
    src = np.ascontiguousarray(src, np.int16)
    h, w = src.shape
    out = np.empty((h, w), np.float32)
    lib().orc_box_meansq_16s(_p(src), C.c_int64(w), w, h, int(r), _p(out))
    return out


def resize_linear(src, dsize_wh):
    src = np.ascontiguousarray(src)
    sh, sw = src.shape
    dw, dh = dsize_wh
    out = np.empty((dh, dw), src.dtype)
    fn = lib().orc_resize_linear_16s if src.dtype == np.int16 else lib().orc_resize_linear_32f
    fn(_p(src), C.c_int64(sw), sw, sh, _p(out), C.c_int64(dw), dw, dh)
    return out


def wls_filter(disp_left, view, disp_right=None, roi=None, use_confidence=True, lam=8000.0, sigma=1.0,
               lrc_thresh=24, disc_radius=5, offsets=(0, 0, 0, 0), assoc=0, nstripes=1, threaded=False,
               want_mask=False):
    """DisparityWLSFilter::filter.  Returns dict(out, conf, roi, mask, rc)."""
    dl = np.ascontiguousarray(disp_left, np.int16)
    dr = None if disp_right is None else np.ascontiguousarray(disp_right, np.int16)
    view = np.ascontiguousarray(view, np.uint8)
    dh, dw = dl.shape
    H, W = view.shape[:2]
    vcn = 1 if view.ndim == 2 else view.shape[2]
    p = OrcWlsParams(lam, sigma, lrc_thresh, disc_radius, offsets[0], offsets[1], offsets[2], offsets[3],
                     int(bool(use_confidence)), 0, 0, 0, 0, int(assoc), int(nstripes), int(bool(threaded)))
    if roi is not None:
        p.roi_x, p.roi_y, p.roi_w, p.roi_h = [int(v) for v in roi]
    out = np.empty((H, W), np.int16)
    conf = np.zeros((H, W), np.float32) if use_confidence else None
    roi_out = (C.c_int * 4)()
    mask = np.zeros((dh, dw), np.uint8) if (want_mask and use_confidence) else None
    rc = lib().orc_wls_filter(_p(dl), C.c_int64(dl.strides[0] // 2),
                              _p(dr) if dr is not None else None, C.c_int64(dr.strides[0] // 2 if dr is not None else 0),
                              dw, dh, _p(view), C.c_int64(view.strides[0]), W, H, vcn, C.byref(p),
                              _p(out), C.c_int64(W), _p(conf) if conf is not None else None, C.c_int64(W),
                              roi_out, _p(mask) if mask is not None else None)
    return dict(out=out, conf=conf, roi=tuple(roi_out), mask=mask, rc=rc)
