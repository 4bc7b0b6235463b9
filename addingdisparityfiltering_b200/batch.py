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


def _bad(msg):
    raise _lib.DwlsError(_lib.DWLS_BAD_ARG, msg)


def _is_torch(t):
    return hasattr(t, "data_ptr")


def _dtype_name(t):
    return str(t.dtype).replace("torch.", "")


def _check_array(name, t, dtype, shape, on_device, device_index):
    """The checks image_of() makes for single frames, for one array of a batch: dtype, rank, shape, pixels of a row
    contiguous, non-negative strides, placement (CUDA tensor on the ctx's device / host memory)."""
    if _dtype_name(t) != dtype:
        _bad("%s must be %s, got %s" % (name, dtype, _dtype_name(t)))
    if tuple(t.shape) != tuple(shape):
        _bad("%s must have shape %s, got %s" % (name, tuple(shape), tuple(t.shape)))
    st = _strides_bytes(t)
    es = t.element_size() if _is_torch(t) else t.itemsize
    px = es * (t.shape[3] if len(t.shape) == 4 else 1)             # bytes per pixel
    if st[2] != px or (len(t.shape) == 4 and st[3] != es):
        _bad("%s: the pixels of a row must be contiguous" % name)
    if (t.shape[1] > 1 and st[1] < px * t.shape[2]) or (t.shape[0] > 1 and st[0] < st[1] * (t.shape[1] - 1) + px * t.shape[2]):
        _bad("%s: rows and frames must not overlap (positive strides, frame after frame)" % name)
    if on_device:
        if not (_is_torch(t) and t.is_cuda):
            _bad("%s must be a CUDA tensor" % name)
        if t.device.index != device_index:
            _bad("%s lives on cuda:%s but the context owns cuda:%s" % (name, t.device.index, device_index))
    elif _is_torch(t) and t.is_cuda:
        _bad("%s must be host memory (numpy array or pinned torch tensor)" % name)


def _make_batch(disp_left, disp_right, view, out, conf, on_device=False, device_index=0):
    if len(disp_left.shape) != 3 or disp_left.shape[0] <= 0:
        _bad("disp_left must be [n, rows, cols] with n > 0")
    if len(view.shape) not in (3, 4) or (len(view.shape) == 4 and view.shape[3] not in (1, 3)):
        _bad("view must be [n, H, W] or [n, H, W, 1 or 3]")
    n, dh, dw = disp_left.shape
    H, W = view.shape[1], view.shape[2]
    cn = 1 if len(view.shape) == 3 else view.shape[3]
    _check_array("disp_left", disp_left, "int16", (n, dh, dw), on_device, device_index)
    if disp_right is not None:
        _check_array("disp_right", disp_right, "int16", (n, dh, dw), on_device, device_index)      # disparity_filters.cpp:263-264
    _check_array("view", view, "uint8", (n, H, W) if len(view.shape) == 3 else (n, H, W, cn), on_device, device_index)
    _check_array("out", out, "int16", (n, H, W), on_device, device_index)
    if conf is not None:
        _check_array("conf", conf, "float32", (n, H, W), on_device, device_index)
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
        b = _make_batch(disp_left, disp_right, view, out, conf, on_device=True, device_index=self.ctx.device)
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


def shard_frames_weighted(n_frames, world_size, rank, root=0, root_extra=0):
    """Contiguous blocks like shard_frames, but the gather root owns `root_extra` frames more (> 0) or fewer (< 0) than an
    even share; the other ranks split the rest evenly.  Fewer is the useful direction when many ranks send to one root: the
    root's own results need no transfer, but while it computes it also receives everybody else's (the receive kernels take SMs
    and HBM bandwidth from a bandwidth-bound filter), so it should finish its own, smaller block together with the others."""
    if world_size == 1 or root_extra == 0:
        return shard_frames(n_frames, world_size, rank)
    even = n_frames // world_size
    root_size = max(0, min(n_frames, even + root_extra))
    base, rem = divmod(n_frames - root_size, world_size - 1)
    sizes, j = [], 0
    for r in range(world_size):
        if r == root:
            sizes.append(root_size)
        else:
            sizes.append(base + (1 if j < rem else 0))
            j += 1
    begin = sum(sizes[:rank])
    return begin, begin + sizes[rank]


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


def chunk_plan(n_local, chunk, tail_min=0):
    """Splits a shard of n_local frames into launch / gather chunks [(first, count)]: `chunk` frames each; with tail_min > 0
    the last chunk is halved repeatedly down to tail_min frames, because the gather of the LAST chunk is the only one
    that is not hidden behind the next chunk's kernels."""
    plan, f = [], 0
    while n_local - f > chunk:
        plan.append((f, chunk))
        f += chunk
    left = n_local - f
    while tail_min > 0 and left > tail_min:
        nf = (left + 1) // 2
        plan.append((f, nf))
        f += nf
        left -= nf
    if left > 0:
        plan.append((f, left))
    return plan


class _DeviceArray:
    """A raw device allocation presented through __cuda_array_interface__ so that torch can view it without copying."""

    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"shape": tuple(int(v) for v in shape), "typestr": typestr, "data": (int(ptr), False), "version": 2, "strides": None}


class ShardedBatch:
    """A batch of n_total independent frames sharded over the ranks of a process group (contiguous blocks,
    shard_frames), filtered chunk by chunk, with the CV_16S results gathered over NCCL WHILE the next chunk is computed.

    mode "root": grouped ncclSend / ncclRecv to rank `dst` (every byte crosses NVLink once: (world-1)/world of the batch
    enters the root, each other rank only reads its own shard); mode "all": every rank ends up with the whole batch
    (grouped send/recv between all pairs, (world-1) x the bytes).  The collective of chunk k is issued right after the
    kernels of chunk k: torch's NCCL stream waits for them through an event and runs beside the kernels of chunk k+1 that
    are launched next on the compute stream; wait() joins the two streams at the end of the step.  The receives land
    directly at their final position in `full` (frame order of the unsharded batch) - no staging copy, no reorder pass;
    a rank's own shard is a view of `full`, so the root's own results need no copy either.

    mode "peer": the same gather to rank `dst` without NCCL on the data path - the root's buffer is a CUDA IPC allocation every
    rank maps, and a rank pushes each finished chunk to its final place with a peer-to-peer copy on a copy stream
    (dwls_peer_push: copy engines over NVLink, no SM of either side), again overlapping the next chunk's kernels; wait() makes
    the compute stream wait for the pushes and then synchronises the ranks with one 4-byte all-reduce, after which the root may
    read.  Measured at 8 GPUs, NCCL send / recv from 7 peers into one root reaches about 0.4 TB/s, which bounds the step
    (DESIGN.md section 6); int16 CUDA batches only."""

    def __init__(self, n_total, frame_shape, chunk=16, tail_min=4, mode="root", dst=0, group=None, device=None, dtype=None, root_extra=0, ctx=None):
        import torch
        import torch.distributed as dist
        self.dist = dist if dist.is_available() and dist.is_initialized() else None
        self.group = group
        self.world = self.dist.get_world_size(group) if self.dist else 1
        self.rank = self.dist.get_rank(group) if self.dist else 0
        self.n_total, self.mode, self.dst = n_total, mode, dst
        self.spans = [shard_frames_weighted(n_total, self.world, r, dst, root_extra if mode in ("root", "peer") else 0) for r in range(self.world)]
        self.begin, self.end = self.spans[self.rank]
        self.plans = [chunk_plan(e - b, chunk, tail_min) for b, e in self.spans]      # every rank can derive every peer's plan
        self.plan = self.plans[self.rank]
        holds_all = self.world == 1 or mode == "all" or self.rank == dst
        dtype = dtype or torch.int16
        n_alloc = n_total if holds_all else self.end - self.begin
        self._peer = None
        if mode == "peer" and self.world > 1:
            self._peer_setup(n_total, frame_shape, dtype, device, ctx, holds_all)
        else:
            self.full = torch.empty((n_alloc,) + tuple(frame_shape), dtype=dtype, device=device)
        self.local = self.full[self.begin:self.end] if holds_all else self.full
        self.holds_all = holds_all
        self._works = []
        self.bytes_sent = 0

    def _peer_setup(self, n_total, frame_shape, dtype, device, ctx, is_root):
        import torch
        if dtype != torch.int16 or device is None or torch.device(device).type != "cuda":
            raise _lib.DwlsError(_lib.DWLS_BAD_ARG, "gather mode 'peer' moves int16 CUDA batches")
        device = torch.device(device)
        self.ctx = ctx or _lib.default_context(device.index or 0)
        L = _lib.lib()
        self._frame_bytes = 2
        for v in frame_shape:
            self._frame_bytes *= int(v)
        handle = torch.zeros(64, dtype=torch.uint8, device=device)
        # A rank whose allocation / mapping fails must not leave the others hanging in a collective: everybody takes part in the
        # handle broadcast and in one agreement all-reduce, and then ALL ranks raise together.
        error = None
        if is_root:
            try:
                ptr, hbuf = C.c_void_p(), (C.c_ubyte * 64)()
                self.ctx.check(L.dwls_peer_alloc(self.ctx.handle, C.c_size_t(n_total * self._frame_bytes), C.byref(ptr), hbuf))
                self.full = torch.as_tensor(_DeviceArray(ptr.value, (n_total,) + tuple(frame_shape), "<i2"), device=device)
                assert self.full.data_ptr() == ptr.value and self.full.dtype == torch.int16
                handle.copy_(torch.tensor(list(hbuf), dtype=torch.uint8))
                self._peer = ("root", ptr.value)
            except Exception as e:          # noqa: BLE001 - reported collectively below
                error = e
        self.dist.broadcast(handle, src=self.dst, group=self.group)
        if not is_root:
            try:
                hb = (C.c_ubyte * 64)(*handle.cpu().tolist())
                mapped = C.c_void_p()
                self.ctx.check(L.dwls_peer_open(self.ctx.handle, hb, C.byref(mapped)))
                self.full = torch.empty((self.end - self.begin,) + tuple(frame_shape), dtype=dtype, device=device)
                self._peer = ("rank", mapped.value)
            except Exception as e:          # noqa: BLE001
                error = e
        ok = torch.tensor([0 if error is not None else 1], dtype=torch.int32, device=device)
        self.dist.all_reduce(ok, op=self.dist.ReduceOp.MIN, group=self.group)
        if int(ok.item()) == 0:
            kind_ptr, self._peer = self._peer, None
            if kind_ptr is not None:        # undo what this rank did set up
                if kind_ptr[0] == "rank":
                    L.dwls_peer_close(self.ctx.handle, C.c_void_p(kind_ptr[1]))
                else:
                    self.full = None
                    L.dwls_peer_free(self.ctx.handle, C.c_void_p(kind_ptr[1]))
            raise _lib.DwlsError(_lib.DWLS_CUDA_ERROR, "peer-memory gather could not be set up on every rank (this rank: %s)" % (error if error is not None else "ok"))
        self._token = torch.zeros(1, dtype=torch.int32, device=device)
        self._device = device

    def close(self):
        """Releases the peer mapping / the root's IPC allocation (collective: every rank calls it once the buffer is no longer read)."""
        if self._peer is None:
            return
        import torch
        torch.cuda.synchronize(self._device)
        self.dist.barrier(group=self.group)
        kind, ptr = self._peer
        self._peer = None
        if kind == "rank":
            self.ctx.check(_lib.lib().dwls_peer_close(self.ctx.handle, C.c_void_p(ptr)))
            self.dist.barrier(group=self.group)
        else:
            self.dist.barrier(group=self.group)            # the other ranks have unmapped
            self.full = self.local = None
            self.ctx.check(_lib.lib().dwls_peer_free(self.ctx.handle, C.c_void_p(ptr)))

    def _stream(self):
        import torch
        return torch.cuda.current_stream(self._device).cuda_stream or 1      # 1 = cudaStreamLegacy, torch's default stream

    def _wire(self, t):
        # the collective only moves bytes (NCCL has no int16 type, gloo no int16 kernels)
        import torch
        return t.view(torch.uint8).view(-1)

    def _frames_of(self, r, k):
        b = self.spans[r][0]
        f0, nf = self.plans[r][k]
        return self._wire(self.full[b + f0:b + f0 + nf])

    def gather_chunk(self, k):
        """Queues the collective for my k-th chunk (and, on receiving ranks, the k-th chunk of every peer that has one)."""
        if self.world == 1:
            return
        dist = self.dist
        ops = []
        f0, nf = self.plan[k] if k < len(self.plan) else (0, 0)
        if self._peer is not None:
            if self._peer[0] == "rank" and nf:
                src = self.local[f0:f0 + nf]
                nbytes = nf * self._frame_bytes
                self.ctx.check(_lib.lib().dwls_peer_push(self.ctx.handle, C.c_void_p(self._peer[1] + (self.begin + f0) * self._frame_bytes),
                                                         C.c_void_p(src.data_ptr()), C.c_size_t(nbytes), C.c_void_p(self._stream())))
                self.bytes_sent += nbytes
            return
        mine = self._wire(self.local[f0:f0 + nf]) if nf else None
        if self.mode == "root":
            if self.rank == self.dst:
                ops = [dist.P2POp(dist.irecv, self._frames_of(r, k), r, self.group) for r in range(self.world)
                       if r != self.dst and k < len(self.plans[r])]
            elif nf:
                ops = [dist.P2POp(dist.isend, mine, self.dst, self.group)]
                self.bytes_sent += mine.numel() * mine.element_size()
        else:
            for r in range(self.world):
                if r == self.rank:
                    continue
                if nf:
                    ops.append(dist.P2POp(dist.isend, mine, r, self.group))
                    self.bytes_sent += mine.numel() * mine.element_size()
                if k < len(self.plans[r]):
                    ops.append(dist.P2POp(dist.irecv, self._frames_of(r, k), r, self.group))
        if ops:
            self._works += dist.batch_isend_irecv(ops)

    @property
    def n_chunks(self):
        return max(len(p) for p in self.plans)

    def wait(self):
        """Joins the communication stream into the caller's stream (GPU) / blocks until the transfers are done (CPU)."""
        for w in self._works:
            w.wait()
        self._works = []
        if self._peer is not None:
            if self._peer[0] == "rank":
                self.ctx.check(_lib.lib().dwls_peer_flush(self.ctx.handle, C.c_void_p(self._stream())))
            self.dist.all_reduce(self._token, group=self.group)      # stream-ordered after every rank's pushes: the root may read

    def run(self, filter_chunk, gather=True):
        """filter_chunk(first_local_frame, count, out_view) computes one chunk into out_view (asynchronously on the current
        stream); chunks of peers that are longer than mine still need my receives, hence n_chunks."""
        for k in range(self.n_chunks):
            if k < len(self.plan):
                f0, nf = self.plan[k]
                filter_chunk(f0, nf, self.local[f0:f0 + nf])
            if gather:
                self.gather_chunk(k)
        if gather:
            self.wait()
        return self.full
