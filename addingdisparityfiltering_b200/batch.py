"""Batched DisparityWLSFilter::filter: many independent stereo frames per call.

Device entry: torch CUDA tensors (plumbing only: device memory + stream) -> dwls_filter_batch_dev.
Host entry: pinned torch / numpy host buffers -> dwls_filter_batch_host (pipelined H2D / kernels / D2H).
Multi-GPU: frames are independent (disparity_filters.cpp keeps no state between filter() calls
except the getters), so a batch is sharded into contiguous blocks of frames, one block per rank,
with no collective on the data path; results are gathered with NCCL only on request.
"""
import ctypes as C

from . import _lib
from ._lib import Batch
from .ximgproc import _params


def _ptr(t):
    return t.data_ptr() if hasattr(t, "data_ptr") else t.ctypes.data


def _strides_bytes(t):
    if hasattr(t, "element_size"):
        return [int(s) * t.element_size() for s in t.stride()]
    return [int(s) for s in t.strides]


def _make_batch(disp_left, disp_right, view, out, conf):
    n, dh, dw = disp_left.shape
    H, W = view.shape[1], view.shape[2]
    cn = 1 if view.ndim == 3 else view.shape[3]
    b = Batch()
    b.n_frames = n
    sl = _strides_bytes(disp_left)
    b.disp_left, b.dl_frame, b.dl_step = _ptr(disp_left), sl[0], sl[1]
    if disp_right is not None:
        sr = _strides_bytes(disp_right)
        b.disp_right, b.dr_frame, b.dr_step = _ptr(disp_right), sr[0], sr[1]
    b.disp_cols, b.disp_rows = dw, dh
    sv = _strides_bytes(view)
    b.view, b.view_frame, b.view_step = _ptr(view), sv[0], sv[1]
    b.view_cols, b.view_rows, b.view_channels = W, H, cn
    so = _strides_bytes(out)
    b.out, b.out_frame, b.out_step = _ptr(out), so[0], so[1]
    if conf is not None:
        sc = _strides_bytes(conf)
        b.conf, b.conf_frame, b.conf_step = _ptr(conf), sc[0], sc[1]
    return b


class BatchFilter:
    """Parameters of a DisparityWLSFilter applied to a batch (same meaning as the single-frame setters)."""

    def __init__(self, use_confidence=True, lam=8000.0, sigma_color=1.0, lrc_thresh=24, disc_radius=5, offsets=(0, 0, 0, 0),
                 ctx=None, device=0):
        self.use_confidence, self.lam, self.sigma_color = use_confidence, lam, sigma_color
        self.lrc_thresh, self.disc_radius, self.offsets = lrc_thresh, disc_radius, offsets
        self.ctx = ctx or _lib.default_context(device)
        self.roi = (0, 0, 0, 0)

    def _call(self, fn, b, ROI, *extra):
        p = _params(self.lam, self.sigma_color, self.lrc_thresh, self.disc_radius, self.offsets, self.use_confidence)
        roi_in = (C.c_int * 4)(*[int(v) for v in ROI]) if ROI is not None else None
        roi_out = (C.c_int * 4)()
        st = fn(self.ctx.handle, C.byref(b), roi_in, C.byref(p), roi_out, *extra)
        self.roi = tuple(roi_out)
        self.ctx.check(st)

    def filter_device(self, disp_left, view, out, disp_right=None, conf=None, ROI=None, stream=None):
        """All arguments are CUDA tensors: disp [n,dh,dw] int16, view [n,H,W(,c)] uint8, out [n,H,W] int16,
        conf [n,H,W] float32 or None.  Asynchronous on `stream` (default: torch's current stream)."""
        import torch
        if stream is None:
            stream = torch.cuda.current_stream(disp_left.device).cuda_stream
        if not stream:
            stream = 1          # cudaStreamLegacy: torch's default stream, not the ctx's private one
        b = _make_batch(disp_left, disp_right, view, out, conf)
        self._call(_lib.lib().dwls_filter_batch_dev, b, ROI, C.c_void_p(stream))
        return out

    def filter_host(self, disp_left, view, out, disp_right=None, conf=None, ROI=None):
        """Same with host buffers (numpy arrays or pinned torch tensors); returns when `out` is complete."""
        b = _make_batch(disp_left, disp_right, view, out, conf)
        self._call(_lib.lib().dwls_filter_batch_host, b, ROI)
        return out


def shard_frames(n_frames, world_size, rank):
    """Contiguous block of frames owned by `rank`: [begin, end)."""
    base, rem = divmod(n_frames, world_size)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def gather_results(local_out, group=None):
    """all_gather of the CV_16S outputs over NCCL (GPU tensors) or gloo (CPU tensors).  Every rank must hold
    the same number of frames (pad the tail shard).  Returns [world * n_local, H, W]."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    full = torch.empty((world * local_out.shape[0],) + tuple(local_out.shape[1:]), dtype=local_out.dtype, device=local_out.device)
    # gathered as raw bytes: the collective only moves data, and gloo has no int16 kernels
    dist.all_gather_into_tensor(full.view(torch.uint8), local_out.contiguous().view(torch.uint8), group=group)
    return full
