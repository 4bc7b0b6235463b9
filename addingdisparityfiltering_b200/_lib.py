"""ctypes binding of libdwls_b200.so (C-ABI in include/dwls_b200.h).

The CUDA library is the only implementation: if it is missing or no CUDA device is present the
calls raise — there is no CPU path in this package.
"""
import ctypes as C
import os
import subprocess
import threading

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("DWLS_LIB", os.path.join(_HERE, "libdwls_b200.so"))      # DWLS_LIB: tuning builds only
CSRC = os.path.join(_HERE, "csrc")

DWLS_OK, DWLS_BAD_ARG, DWLS_BAD_SIZE, DWLS_CUDA_ERROR, DWLS_UNSUPPORTED = 0, 1, 2, 3, 4
DEPTH = {np.dtype(np.uint8): 0, np.dtype(np.int16): 3, np.dtype(np.float32): 5}


class DwlsError(RuntimeError):
    """Maps the C-ABI status codes to the exceptions the reference raises (cv::Exception)."""

    def __init__(self, status, msg):
        super().__init__("dwls_b200 status %d: %s" % (status, msg))
        self.status = status


class Image(C.Structure):
    _fields_ = [("data", C.c_void_p), ("step", C.c_int64), ("rows", C.c_int), ("cols", C.c_int),
                ("depth", C.c_int), ("channels", C.c_int)]


class Params(C.Structure):
    _fields_ = [("lambda_", C.c_double), ("sigma_color", C.c_double), ("lrc_thresh", C.c_int), ("disc_radius", C.c_int),
                ("left_offset", C.c_int), ("right_offset", C.c_int), ("top_offset", C.c_int), ("bottom_offset", C.c_int),
                ("use_confidence", C.c_int), ("lambda_attenuation", C.c_double), ("num_iter", C.c_int)]


class Batch(C.Structure):
    _fields_ = [("n_frames", C.c_int),
                ("disp_left", C.c_void_p), ("dl_step", C.c_int64), ("dl_frame", C.c_int64),
                ("disp_right", C.c_void_p), ("dr_step", C.c_int64), ("dr_frame", C.c_int64),
                ("disp_cols", C.c_int), ("disp_rows", C.c_int),
                ("view", C.c_void_p), ("view_step", C.c_int64), ("view_frame", C.c_int64),
                ("view_cols", C.c_int), ("view_rows", C.c_int), ("view_channels", C.c_int),
                ("out", C.c_void_p), ("out_step", C.c_int64), ("out_frame", C.c_int64),
                ("conf", C.c_void_p), ("conf_step", C.c_int64), ("conf_frame", C.c_int64)]


EXPORTS = ["dwls_default_params", "dwls_ctx_create", "dwls_ctx_destroy", "dwls_last_error", "dwls_version",
           "dwls_kernel_launches", "dwls_set_group_frames", "dwls_set_lanes", "dwls_set_kernel_timing", "dwls_kernel_timing_report", "dwls_set_cuda_graphs", "dwls_set_plane_interleave", "dwls_graph_replays", "dwls_transfer_bytes", "dwls_filter_host", "dwls_filter_batch_dev",
           "dwls_filter_batch_host", "dwls_fgs_host", "dwls_fgs_dev", "dwls_lrc_mask_host", "dwls_fgs_weights_host",
           "dwls_fgs_dev_noalloc", "dwls_fgs_noalloc_scratch_bytes", "dwls_fgs_noalloc_required_bytes",
           "dwls_peer_alloc", "dwls_peer_free", "dwls_peer_open", "dwls_peer_close", "dwls_peer_push", "dwls_peer_flush", "dwls_peer_bytes",
           "dwls_bm_default_params", "dwls_stereo_bm_dev", "dwls_stereo_bm_host", "dwls_halve_views_dev"]

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared"]


def build(force=False, verbose=False):
    """Compile csrc/ into libdwls_b200.so for sm_100a (in-tree, so the .so travels with the repo)."""
    srcs = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith((".cu", ".cuh"))]
    srcs.append(os.path.join(_HERE, "..", "include", "dwls_b200.h"))
    if not force and os.path.exists(LIB_PATH) and all(os.path.getmtime(s) <= os.path.getmtime(LIB_PATH) for s in srcs):
        return LIB_PATH
    tmp = LIB_PATH + ".build%d.so" % os.getpid()        # build beside the target, then rename: a reader never sees a half-written library
    cmd = ["nvcc"] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", tmp, os.path.join(CSRC, "dwls_capi.cu")]
    try:
        subprocess.check_call(cmd)
        os.replace(tmp, LIB_PATH)
    finally:
        if os.path.exists(tmp):
            os.remove(tmp)
    return LIB_PATH


_lib = None
_lock = threading.Lock()


def lib():
    global _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise ImportError("libdwls_b200.so is not built (run `python -c 'import __graft_entry__ as g; g.build()'`); "
                                  "there is no CPU fallback")
            L = C.CDLL(LIB_PATH)
            L.dwls_last_error.restype = C.c_char_p
            L.dwls_last_error.argtypes = [C.c_void_p]
            L.dwls_version.restype = C.c_char_p
            L.dwls_kernel_launches.restype = C.c_int64
            L.dwls_kernel_launches.argtypes = [C.c_void_p]
            L.dwls_ctx_create.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
            L.dwls_ctx_destroy.argtypes = [C.c_void_p]
            L.dwls_set_group_frames.argtypes = [C.c_void_p, C.c_int]
            L.dwls_set_lanes.argtypes = [C.c_void_p, C.c_int]
            L.dwls_peer_alloc.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p), C.c_void_p]
            L.dwls_peer_free.argtypes = [C.c_void_p, C.c_void_p]
            L.dwls_peer_open.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p)]
            L.dwls_peer_close.argtypes = [C.c_void_p, C.c_void_p]
            L.dwls_peer_push.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
            L.dwls_peer_flush.argtypes = [C.c_void_p, C.c_void_p]
            L.dwls_peer_bytes.restype = C.c_int64
            L.dwls_peer_bytes.argtypes = [C.c_void_p]
            L.dwls_set_kernel_timing.argtypes = [C.c_void_p, C.c_int]
            L.dwls_set_cuda_graphs.argtypes = [C.c_void_p, C.c_int]
            L.dwls_set_plane_interleave.argtypes = [C.c_void_p, C.c_int]
            L.dwls_graph_replays.restype = C.c_int64
            L.dwls_graph_replays.argtypes = [C.c_void_p]
            L.dwls_transfer_bytes.restype = None
            L.dwls_transfer_bytes.argtypes = [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
            _lib = L
    return _lib


def image_of(a):
    """dwls_image header of a numpy array (rows x cols [x channels]); the array must stay alive."""
    if a.ndim not in (2, 3) or a.dtype not in DEPTH:
        raise DwlsError(DWLS_BAD_ARG, "unsupported array type %s / ndim %d" % (a.dtype, a.ndim))
    cn = 1 if a.ndim == 2 else a.shape[2]
    es = a.dtype.itemsize
    if a.strides[1] != es * cn or (a.ndim == 3 and a.strides[2] != es):
        raise DwlsError(DWLS_BAD_ARG, "pixels of a row must be contiguous")
    return Image(a.ctypes.data, a.strides[0], a.shape[0], a.shape[1], DEPTH[a.dtype], cn)


class Context:
    """One CUDA device + stream + workspace (dwls_ctx).  Not thread-safe, like the reference's filter objects."""

    def __init__(self, device=0):
        self._h = C.c_void_p()
        L = lib()
        st = L.dwls_ctx_create(int(device), C.byref(self._h))
        if st != DWLS_OK:
            raise DwlsError(st, L.dwls_last_error(None).decode())
        self.device = device

    def close(self):
        if self._h:
            lib().dwls_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def check(self, status):
        if status != DWLS_OK:
            raise DwlsError(status, lib().dwls_last_error(self._h).decode())

    @property
    def handle(self):
        return self._h

    def kernel_launches(self):
        return int(lib().dwls_kernel_launches(self._h))

    def set_group_frames(self, n):
        self.check(lib().dwls_set_group_frames(self._h, int(n)))

    def set_lanes(self, n):
        self.check(lib().dwls_set_lanes(self._h, int(n)))

    def set_plane_interleave(self, on):
        self.check(lib().dwls_set_plane_interleave(self._h, int(bool(on))))

    def set_cuda_graphs(self, on):
        self.check(lib().dwls_set_cuda_graphs(self._h, int(bool(on))))

    def graph_replays(self):
        return int(lib().dwls_graph_replays(self._h))

    def transfer_bytes(self):
        """(host->device, device->host) bytes moved by the host-buffer entry points since the context was created."""
        a, b = C.c_int64(), C.c_int64()
        lib().dwls_transfer_bytes(self._h, C.byref(a), C.byref(b))
        return int(a.value), int(b.value)

    def set_kernel_timing(self, on):
        self.check(lib().dwls_set_kernel_timing(self._h, int(bool(on))))

    def kernel_timing_report(self):
        """{kernel name: (launches, total_ms)} since the last report (synchronises the device)."""
        cap = 32
        names = (C.c_char_p * cap)()
        n = (C.c_int64 * cap)()
        ms = (C.c_double * cap)()
        k = lib().dwls_kernel_timing_report(self._h, names, n, ms, cap)
        return {names[i].decode(): (int(n[i]), float(ms[i])) for i in range(k)}


_default_ctx = {}


def default_context(device=0):
    if device not in _default_ctx:
        _default_ctx[device] = Context(device)
    return _default_ctx[device]
