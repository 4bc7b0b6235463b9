"""Device stereo block matcher with cv2.StereoBM's interface (the producer of the filter's inputs).

`StereoBM.create(numDisparities, blockSize)` mirrors cv::StereoBM::create; the object is accepted by
createDisparityWLSFilter / createRightMatcher exactly like a cv2 matcher (disparity_filters.cpp:385-449), and
`compute(left, right)` returns the CV_16S map cv2.StereoBM.compute returns - bit for bit - computed by CUDA kernels
(csrc/dwls_stereo.cuh) through the C-ABI (dwls_stereo_bm_host / dwls_stereo_bm_dev).  There is no CPU path.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import DwlsError, image_of


class BmParams(C.Structure):
    _fields_ = [("num_disparities", C.c_int), ("block_size", C.c_int), ("min_disparity", C.c_int), ("pre_filter_cap", C.c_int),
                ("texture_threshold", C.c_int), ("uniqueness_ratio", C.c_int)]


class Views(C.Structure):
    _fields_ = [("n_frames", C.c_int), ("left", C.c_void_p), ("left_step", C.c_int64), ("left_frame", C.c_int64),
                ("right", C.c_void_p), ("right_step", C.c_int64), ("right_frame", C.c_int64),
                ("cols", C.c_int), ("rows", C.c_int), ("channels", C.c_int)]


class StereoBM:
    """cv2.StereoBM look-alike running on the GPU.  `kind` lets the WLS factories recognise it."""
    kind = "bm"

    def __init__(self, numDisparities=0, blockSize=21, ctx=None):
        self._p = dict(MinDisparity=0, NumDisparities=numDisparities if numDisparities else 64, BlockSize=blockSize, SpeckleWindowSize=0,
                       SpeckleRange=0, Disp12MaxDiff=-1, PreFilterCap=31, TextureThreshold=10, UniquenessRatio=15)
        self._ctx = ctx

    @staticmethod
    def create(numDisparities=0, blockSize=21, ctx=None):
        return StereoBM(numDisparities, blockSize, ctx=ctx)

    def __getattr__(self, name):
        if name.startswith("get") and name[3:] in self.__dict__.get("_p", {}):
            return lambda: self._p[name[3:]]
        if name.startswith("set") and name[3:] in self.__dict__.get("_p", {}):
            return lambda v: self._p.__setitem__(name[3:], int(v))
        raise AttributeError(name)

    def _c_params(self):
        p = self._p
        if p["SpeckleWindowSize"] > 0:
            raise DwlsError(_lib.DWLS_UNSUPPORTED, "speckle filtering is not implemented by the device matcher (the WLS factories switch it off)")
        if 0 <= p["Disp12MaxDiff"] < 1000000:
            raise DwlsError(_lib.DWLS_UNSUPPORTED, "disp12MaxDiff is not implemented by the device matcher (the WLS factories switch it off)")
        return BmParams(p["NumDisparities"], p["BlockSize"], p["MinDisparity"], p["PreFilterCap"], p["TextureThreshold"], p["UniquenessRatio"])

    def compute(self, left, right, with_right=False):
        """StereoBM::compute(left, right) -> CV_16S disparity (x16).  with_right=True also returns the map of the matcher
        createRightMatcher(self) would build, computed in the same launches: compute(right, left) with minDisparity
        -(min + num) + 1."""
        ctx = self._ctx or _lib.default_context()
        left, right = np.ascontiguousarray(left), np.ascontiguousarray(right)
        if left.dtype != np.uint8 or right.dtype != np.uint8 or left.shape != right.shape or left.ndim not in (2, 3):
            raise DwlsError(_lib.DWLS_BAD_ARG, "left and right must be CV_8U images of equal size and type")
        h, w = left.shape[:2]
        dl = np.empty((h, w), np.int16)
        dr = np.empty((h, w), np.int16) if with_right else None
        p = self._c_params()
        i_l, i_r, i_dl = image_of(left), image_of(right), image_of(dl)
        i_dr = image_of(dr) if with_right else None
        ctx.check(_lib.lib().dwls_stereo_bm_host(ctx.handle, C.byref(i_l), C.byref(i_r), C.byref(p), C.byref(i_dl), C.byref(i_dr) if with_right else None))
        return (dl, dr) if with_right else dl

    def compute_device(self, left, right, disp_left, disp_right=None, stream=None):
        """Batch on the device: left/right uint8 CUDA tensors [n,H,W] or [n,H,W,3] (BGR), disp_* int16 CUDA tensors [n,H,W].
        Asynchronous on `stream` (default: torch's current stream)."""
        import torch
        ctx = self._ctx or _lib.default_context()
        if stream is None:
            stream = torch.cuda.current_stream(left.device).cuda_stream
        if not stream:
            stream = 1
        v = views_of(left, right, ctx.device)
        n, H, W = left.shape[:3]
        for name, d in (("disp_left", disp_left), ("disp_right", disp_right)):
            if d is None:
                continue
            if d.dtype != torch.int16 or tuple(d.shape) != (n, H, W) or not d.is_cuda or d.device.index != ctx.device or d.stride(2) != 1:
                raise DwlsError(_lib.DWLS_BAD_ARG, "%s must be an int16 CUDA tensor [n,H,W] on the context's device with contiguous rows" % name)
        p = self._c_params()
        L = _lib.lib()
        ctx.check(L.dwls_stereo_bm_dev(ctx.handle, C.byref(v), C.byref(p), C.c_void_p(disp_left.data_ptr()), C.c_int64(disp_left.stride(1) * 2),
                                       C.c_int64(disp_left.stride(0) * 2), C.c_void_p(disp_right.data_ptr() if disp_right is not None else 0),
                                       C.c_int64(disp_right.stride(1) * 2 if disp_right is not None else 0),
                                       C.c_int64(disp_right.stride(0) * 2 if disp_right is not None else 0), C.c_void_p(stream)))
        return disp_left, disp_right


def views_of(left, right, device_index):
    """dwls_views header of two uint8 CUDA tensors [n,H,W] or [n,H,W,3]."""
    import torch
    for name, t in (("left", left), ("right", right)):
        if t.dtype != torch.uint8 or t.dim() not in (3, 4) or (t.dim() == 4 and t.shape[3] not in (1, 3)) or not t.is_cuda or t.device.index != device_index:
            raise DwlsError(_lib.DWLS_BAD_ARG, "%s must be a uint8 CUDA tensor [n,H,W] or [n,H,W,3] on the context's device" % name)
        cn = 1 if t.dim() == 3 else t.shape[3]
        if t.stride(2) != cn or (t.dim() == 4 and t.stride(3) != 1):
            raise DwlsError(_lib.DWLS_BAD_ARG, "%s: the pixels of a row must be contiguous" % name)
    if left.shape != right.shape:
        raise DwlsError(_lib.DWLS_BAD_ARG, "left and right must have the same shape")
    n, H, W = left.shape[:3]
    cn = 1 if left.dim() == 3 else left.shape[3]
    return Views(n, left.data_ptr(), left.stride(1), left.stride(0), right.data_ptr(), right.stride(1), right.stride(0), W, H, cn)


def halve_views_device(left, right, ctx=None, stream=None):
    """The sample's downscale step on the device (cv::resize by 0.5, even sizes): returns (left_half, right_half)."""
    import torch
    ctx = ctx or _lib.default_context()
    if stream is None:
        stream = torch.cuda.current_stream(left.device).cuda_stream
    if not stream:
        stream = 1
    v = views_of(left, right, ctx.device)
    shape = (left.shape[0], left.shape[1] // 2, left.shape[2] // 2) + tuple(left.shape[3:])
    lo, ro = torch.empty(shape, dtype=torch.uint8, device=left.device), torch.empty(shape, dtype=torch.uint8, device=left.device)
    ctx.check(_lib.lib().dwls_halve_views_dev(ctx.handle, C.byref(v), C.c_void_p(lo.data_ptr()), C.c_void_p(ro.data_ptr()), C.c_int64(lo.stride(1)),
                                              C.c_int64(lo.stride(0)), C.c_void_p(stream)))
    return lo, ro
