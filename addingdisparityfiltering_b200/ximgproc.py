"""Host-side mirror of the reference's `cv::ximgproc` disparity-filtering API on numpy arrays
(cv2 Mats are numpy arrays in Python), calling the CUDA library through its C-ABI.

Same names, argument meaning and error behaviour as
  modules/ximgproc/include/opencv2/ximgproc/disparity_filter.hpp:52-204
  modules/ximgproc/include/opencv2/ximgproc/edge_filter.hpp:361-413
(the reference wraps them for Python with CV_WRAP / `WRAP python`, modules/ximgproc/CMakeLists.txt:3).
"""
import ctypes as C
import math

import numpy as np

from . import _lib
from ._lib import DwlsError, Image, Params, image_of

UNKNOWN_DISPARITY = 16320          # disparity_filters.cpp:460


def _params(lam, sigma, lrc, radius, offsets, use_conf, att=0.25, num_iter=3):
    return Params(float(lam), float(sigma), int(lrc), int(radius), int(offsets[0]), int(offsets[1]), int(offsets[2]),
                  int(offsets[3]), int(bool(use_conf)), float(att), int(num_iter))


class DisparityWLSFilter:
    """DisparityWLSFilterImpl (disparity_filters.cpp:49-140): parameters + last ROI / confidence map."""

    def __init__(self, use_confidence, l_offs=0, r_offs=0, t_offs=0, b_offs=0, ctx=None):
        # DisparityWLSFilterImpl::init (142-159) with create()'s defaults (215)
        self._offsets = (l_offs, r_offs, t_offs, b_offs)
        self._use_confidence = bool(use_confidence)
        self._lambda, self._sigma_color = 8000.0, 1.0
        self._lrc_thresh, self._disc_radius = 24, 5
        self._roi = (0, 0, 0, 0)
        self._confidence_map = None
        self._ctx = ctx

    # -- parameter accessors (disparity_filter.hpp:90-122) --
    def getLambda(self): return self._lambda
    def setLambda(self, _lambda): self._lambda = float(_lambda)
    def getSigmaColor(self): return self._sigma_color
    def setSigmaColor(self, _sigma_color): self._sigma_color = float(_sigma_color)
    def getLRCthresh(self): return self._lrc_thresh
    def setLRCthresh(self, _LRC_thresh): self._lrc_thresh = int(_LRC_thresh)
    def getDepthDiscontinuityRadius(self): return self._disc_radius
    def setDepthDiscontinuityRadius(self, _disc_radius): self._disc_radius = int(_disc_radius)
    def getConfidenceMap(self): return self._confidence_map
    def getROI(self): return self._roi

    def _c_params(self):
        return _params(self._lambda, self._sigma_color, self._lrc_thresh, self._disc_radius, self._offsets, self._use_confidence)

    def filter(self, disparity_map_left, left_view, filtered_disparity_map=None, disparity_map_right=None, ROI=None, right_view=None):
        """DisparityFilter::filter (disparity_filter.hpp:75).  Returns the filtered CV_16S map (view size)."""
        ctx = self._ctx or _lib.default_context()
        dl = np.asarray(disparity_map_left)
        view = np.asarray(left_view)
        if dl.size == 0 or dl.dtype != np.int16 or dl.ndim != 2:
            raise DwlsError(_lib.DWLS_BAD_ARG, "disparity_map_left must be a non-empty CV_16SC1 image")
        if view.size == 0 or view.dtype != np.uint8 or not (view.ndim == 2 or (view.ndim == 3 and view.shape[2] in (1, 3))):
            raise DwlsError(_lib.DWLS_BAD_ARG, "left_view must be a non-empty CV_8UC1 or CV_8UC3 image")
        dl = np.ascontiguousarray(dl)
        view = np.ascontiguousarray(view)
        dr = None
        if disparity_map_right is not None and np.asarray(disparity_map_right).size:
            dr = np.ascontiguousarray(disparity_map_right)
            if dr.dtype != np.int16 or dr.ndim != 2:
                raise DwlsError(_lib.DWLS_BAD_ARG, "disparity_map_right must be a CV_16SC1 image")
        H, W = view.shape[:2]
        out = filtered_disparity_map if (isinstance(filtered_disparity_map, np.ndarray) and filtered_disparity_map.shape == (H, W)
                                         and filtered_disparity_map.dtype == np.int16) else np.empty((H, W), np.int16)
        conf = np.empty((H, W), np.float32) if self._use_confidence else None
        roi_in = (C.c_int * 4)(*[int(v) for v in ROI]) if ROI is not None else None
        roi_out = (C.c_int * 4)()
        p = self._c_params()
        i_dl, i_view, i_out = image_of(dl), image_of(view), image_of(out)
        i_dr = image_of(dr) if dr is not None else None
        i_conf = image_of(conf) if conf is not None else None
        st = _lib.lib().dwls_filter_host(ctx.handle, C.byref(i_dl), C.byref(i_dr) if i_dr is not None else None, C.byref(i_view),
                                         roi_in, C.byref(p), C.byref(i_out), C.byref(i_conf) if i_conf is not None else None, roi_out)
        self._roi = tuple(roi_out)
        ctx.check(st)
        if conf is not None:
            self._confidence_map = conf
        return out

    def lrc_mask(self, disparity_map_left, disparity_map_right, view_cols=None, ROI=None):
        """Parity export: the per-pixel decision of ComputeDiscontinuityAwareLRC_ParBody (306-341)."""
        ctx = self._ctx or _lib.default_context()
        dl = np.ascontiguousarray(disparity_map_left, np.int16)
        dr = np.ascontiguousarray(disparity_map_right, np.int16)
        mask = np.empty(dl.shape, np.uint8)
        roi_in = (C.c_int * 4)(*[int(v) for v in ROI]) if ROI is not None else None
        p = self._c_params()
        i_dl, i_dr = image_of(dl), image_of(dr)
        ctx.check(_lib.lib().dwls_lrc_mask_host(ctx.handle, C.byref(i_dl), C.byref(i_dr), int(view_cols or dl.shape[1]), roi_in,
                                                C.byref(p), mask.ctypes.data_as(C.c_void_p), C.c_int64(mask.strides[0])))
        return mask


def createDisparityWLSFilterGeneric(use_confidence, ctx=None):
    """disparity_filters.cpp:451-455."""
    return DisparityWLSFilter(use_confidence, ctx=ctx)


def _matcher_kind(m):
    name = type(m).__name__
    if "StereoSGBM" in name:
        return "sgbm"
    if "StereoBM" in name:
        return "bm"
    return getattr(m, "kind", None)        # duck-typed matchers used by the tests


def createDisparityWLSFilter(matcher_left, ctx=None):
    """disparity_filters.cpp:385-414: mutates the matcher and derives ROI margins / discontinuity radius."""
    matcher_left.setDisp12MaxDiff(1000000)
    matcher_left.setSpeckleWindowSize(0)
    min_disp = matcher_left.getMinDisparity()
    num_disp = matcher_left.getNumDisparities()
    wsize = matcher_left.getBlockSize()
    wsize2 = wsize // 2
    kind = _matcher_kind(matcher_left)
    if kind == "bm":
        matcher_left.setTextureThreshold(0)
        matcher_left.setUniquenessRatio(0)
        wls = DisparityWLSFilter(True, max(0, min_disp + num_disp) + wsize2, max(0, -min_disp) + wsize2, wsize2, wsize2, ctx=ctx)
        wls.setDepthDiscontinuityRadius(int(math.ceil(0.33 * wsize)))
    elif kind == "sgbm":
        matcher_left.setUniquenessRatio(0)
        wls = DisparityWLSFilter(True, max(0, min_disp + num_disp), max(0, -min_disp), 0, 0, ctx=ctx)
        wls.setDepthDiscontinuityRadius(int(math.ceil(0.5 * wsize)))
    else:
        raise DwlsError(_lib.DWLS_BAD_ARG, "DisparityWLSFilter natively supports only StereoBM and StereoSGBM")
    return wls


def createRightMatcher(matcher_left):
    """disparity_filters.cpp:416-449 (needs cv2 for its matcher classes, exactly like the reference needs calib3d; the
    device matcher addingdisparityfiltering_b200.StereoBM needs nothing)."""
    try:
        import cv2
    except ImportError:
        cv2 = None
    min_disp = matcher_left.getMinDisparity()
    num_disp = matcher_left.getNumDisparities()
    wsize = matcher_left.getBlockSize()
    kind = _matcher_kind(matcher_left)
    if kind == "bm":
        from .stereo import StereoBM
        # the device matcher gets a device right matcher (same class, same context), a cv2 matcher a cv2 one
        right = StereoBM.create(num_disp, wsize, ctx=matcher_left._ctx) if isinstance(matcher_left, StereoBM) else cv2.StereoBM_create(num_disp, wsize)
        if isinstance(matcher_left, StereoBM):
            right.setPreFilterCap(matcher_left.getPreFilterCap())
        right.setMinDisparity(-(min_disp + num_disp) + 1)
        right.setTextureThreshold(0)
        right.setUniquenessRatio(0)
        right.setDisp12MaxDiff(1000000)
        right.setSpeckleWindowSize(0)
        return right
    if kind == "sgbm":
        right = cv2.StereoSGBM_create(-(min_disp + num_disp) + 1, num_disp, wsize)
        right.setUniquenessRatio(0)
        right.setP1(matcher_left.getP1())
        right.setP2(matcher_left.getP2())
        right.setMode(matcher_left.getMode())
        right.setPreFilterCap(matcher_left.getPreFilterCap())
        right.setDisp12MaxDiff(1000000)
        right.setSpeckleWindowSize(0)
        return right
    raise DwlsError(_lib.DWLS_BAD_ARG, "createRightMatcher supports only StereoBM and StereoSGBM")


class FastGlobalSmootherFilter:
    """FastGlobalSmootherFilterImpl (fgs_filter.cpp:61-233): guide + parameters, filter(src) -> dst."""

    def __init__(self, guide, lambda_, sigma_color, lambda_attenuation=0.25, num_iter=3, ctx=None):
        guide = np.asarray(guide)
        if guide.size == 0 or not (lambda_ >= 0) or not (sigma_color >= 0) or num_iter < 1:
            raise DwlsError(_lib.DWLS_BAD_ARG, "guide must be non-empty, lambda >= 0, sigma_color >= 0, num_iter >= 1")
        if guide.dtype != np.uint8 or not (guide.ndim == 2 or (guide.ndim == 3 and guide.shape[2] in (1, 3))):
            raise DwlsError(_lib.DWLS_BAD_ARG, "guide must be CV_8UC1 or CV_8UC3")
        self._guide = np.ascontiguousarray(guide)
        self._args = (float(lambda_), float(sigma_color), float(lambda_attenuation), int(num_iter))
        self._ctx = ctx

    def filter(self, src, dst=None):
        ctx = self._ctx or _lib.default_context()
        src = np.ascontiguousarray(src)
        if src.size == 0 or src.dtype not in _lib.DEPTH or src.ndim not in (2, 3) or (src.ndim == 3 and src.shape[2] > 4):
            raise DwlsError(_lib.DWLS_BAD_ARG, "src must be non-empty CV_8U / CV_16S / CV_32F with at most 4 channels")
        out = np.empty_like(src)
        i_g, i_s, i_d = image_of(self._guide), image_of(src), image_of(out)
        lam, sig, att, it = self._args
        ctx.check(_lib.lib().dwls_fgs_host(ctx.handle, C.byref(i_g), C.byref(i_s), C.byref(i_d), C.c_double(lam), C.c_double(sig),
                                           C.c_double(att), C.c_int(it)))
        return out


def createFastGlobalSmootherFilter(guide, lambda_, sigma_color, lambda_attenuation=0.25, num_iter=3, ctx=None):
    """edge_filter.hpp:393 / fgs_filter.cpp:680-684."""
    return FastGlobalSmootherFilter(guide, lambda_, sigma_color, lambda_attenuation, num_iter, ctx=ctx)


def fastGlobalSmootherFilter(guide, src, lambda_, sigma_color, lambda_attenuation=0.25, num_iter=3, ctx=None):
    """edge_filter.hpp:413 / fgs_filter.cpp:686-691."""
    return createFastGlobalSmootherFilter(guide, lambda_, sigma_color, lambda_attenuation, num_iter, ctx=ctx).filter(src)


def fgs_weights(guide, sigma_color, ctx=None):
    """Parity export of the Chor / Cvert planes (fgs_filter.cpp:593-661)."""
    ctx = ctx or _lib.default_context()
    guide = np.ascontiguousarray(guide)
    h, w = guide.shape[:2]
    chor = np.empty((h, w), np.float32)
    cvert = np.empty((h, w), np.float32)
    i_g = image_of(guide)
    ctx.check(_lib.lib().dwls_fgs_weights_host(ctx.handle, C.byref(i_g), C.c_double(sigma_color), chor.ctypes.data_as(C.c_void_p),
                                               cvert.ctypes.data_as(C.c_void_p)))
    return chor, cvert


# ---- ground-truth / metric helpers (disparity_filters.cpp:460-556), host-side like the reference ----
def readGT(src_path):
    """Returns (status, dst): Middlebury (1-channel) or MPI-Sintel (3-channel) ground truth -> 16 x disparity."""
    import cv2
    src = cv2.imread(src_path, cv2.IMREAD_UNCHANGED)
    if src is None or src.dtype != np.uint8:
        return 1, None
    if src.ndim == 3 and src.shape[2] == 3:
        return 0, (64 * src[:, :, 2].astype(np.int32) + src[:, :, 1].astype(np.int32) // 4).astype(np.int16)
    if src.ndim == 2:
        v = src.astype(np.int16)
        return 0, np.where(v == 0, np.int16(UNKNOWN_DISPARITY), (16 * v).astype(np.int16))
    return 1, None


def _roi_views(GT, src, ROI):
    GT, src = np.asarray(GT), np.asarray(src)
    if GT.dtype != np.int16 or src.dtype != np.int16 or GT.ndim != 2 or GT.shape != src.shape:
        raise DwlsError(_lib.DWLS_BAD_ARG, "GT and src must be CV_16SC1 images of equal size")
    x, y, w, h = [int(v) for v in ROI]
    return GT[y:y + h, x:x + w].astype(np.int64), src[y:y + h, x:x + w].astype(np.int64)


def computeMSE(GT, src, ROI):
    g, s = _roi_views(GT, src, ROI)
    known = g != UNKNOWN_DISPARITY
    return float(((g - s)[known] ** 2).sum() / (int(known.sum()) * 256))


def computeBadPixelPercent(GT, src, ROI, thresh=24):
    g, s = _roi_views(GT, src, ROI)
    known = g != UNKNOWN_DISPARITY
    return float(100.0 * int((np.abs(g - s)[known] >= thresh).sum()) / int(known.sum()))


def getDisparityVis(src, scale=1.0):
    src = np.asarray(src)
    if src.dtype != np.int16 or src.ndim != 2 or src.size == 0:
        raise DwlsError(_lib.DWLS_BAD_ARG, "src must be a non-empty CV_16SC1 image")
    v = np.rint(scale * src.astype(np.float64) / 16.0)          # saturate_cast<uchar>(double) = cvRound + clamp
    out = np.clip(v, 0, 255).astype(np.uint8)
    out[src == UNKNOWN_DISPARITY] = 0
    return out
