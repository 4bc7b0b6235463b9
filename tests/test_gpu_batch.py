"""GPU parity of the batch entry points (dwls_filter_batch_dev / dwls_filter_batch_host) against the CPU oracle,
frame by frame: independent frames of a batch must come out exactly as single filter() calls would
(disparity_filters.cpp keeps no state between calls except the getters, :138-139)."""
import numpy as np
import pytest

import addingdisparityfiltering_b200 as A
from addingdisparityfiltering_b200 import synth
from addingdisparityfiltering_b200.batch import BatchFilter
from oracle import oracle as O
from tests.parity import assert_s16_parity

pytestmark = pytest.mark.gpu


def _frames(n, w, h, cn, seed):
    fr = [synth.make_textured_scene(w, h, cn, seed=seed + i, max_disp=max(32, w // 8)) for i in range(n)]
    return (np.stack([f[0] for f in fr]), np.stack([f[1] for f in fr]), np.stack([f[2] for f in fr]))


@pytest.mark.parametrize("size,radius", [((320, 240), 2), ((328, 122), 5), ((127, 61), 3)])
@pytest.mark.parametrize("strided", [False, True])
def test_batch_device_matches_oracle(ctx, size, radius, strided):
    import torch
    w, h = size
    n = 5
    view, dl, dr = _frames(n, w, h, 3, seed=100 + w)
    md = max(32, w // 8)
    roi = (md + 3, 2, w - md - 7, h - 5)
    dev = torch.device("cuda", 0)

    def put(a, pad_rows, pad_cols):
        # strided: frames and rows live inside a larger allocation (cv::Mat ROI views / padded batches)
        t = torch.from_numpy(a).to(dev)
        if not strided:
            return t
        shape = list(t.shape)
        big = torch.zeros([shape[0], shape[1] + pad_rows, shape[2] + pad_cols] + shape[3:], dtype=t.dtype, device=dev)
        v = big[:, 1:1 + shape[1], 8:8 + shape[2]]
        v.copy_(t)
        return v
    tv, tl, tr = put(view, 3, 24), put(dl, 2, 16), put(dr, 1, 8)
    out = put(np.zeros((n, h, w), np.int16), 2, 16)
    conf = put(np.zeros((n, h, w), np.float32), 2, 16)
    bf = BatchFilter(True, 8000.0, 1.5, 24, radius, ctx=ctx)
    ctx.set_group_frames(2)                      # 5 frames -> groups 2 + 2 + 1
    try:
        bf.filter_device(tl, tv, out, tr, conf, roi)
        torch.cuda.synchronize()
    finally:
        ctx.set_group_frames(0)
    assert tuple(bf.roi) == roi
    for i in range(n):
        ref = O.wls_filter(dl[i], view[i], dr[i], roi=roi, lam=8000.0, sigma=1.5, lrc_thresh=24, disc_radius=radius)
        assert np.abs(conf[i].cpu().numpy() - ref["conf"]).max() <= 1e-4
        assert_s16_parity(out[i].cpu().numpy(), ref["out"], roi)


@pytest.mark.parametrize("use_conf", [True, False])
def test_batch_host_matches_device_and_oracle(ctx, use_conf):
    import torch
    w, h, n = 320, 240, 11                       # 11 frames: groups 8, 2, 1 on the host path
    view, dl, dr = _frames(n, w, h, 3, seed=7)
    roi = (40, 0, w - 40, h)
    bf = BatchFilter(use_conf, 8000.0, 1.5, 24, 2, ctx=ctx)
    out_h = np.zeros((n, h, w), np.int16)
    conf_h = np.zeros((n, h, w), np.float32) if use_conf else None
    bf.filter_host(dl, view, out_h, dr if use_conf else None, conf_h, roi)
    dev = torch.device("cuda", 0)
    out_d = torch.zeros((n, h, w), dtype=torch.int16, device=dev)
    bf.filter_device(torch.from_numpy(dl).to(dev), torch.from_numpy(view).to(dev), out_d,
                     torch.from_numpy(dr).to(dev) if use_conf else None, None, roi)
    torch.cuda.synchronize()
    assert (out_d.cpu().numpy() == out_h).all()  # same kernels, same result bit for bit
    for i in (0, 5, 10):
        ref = O.wls_filter(dl[i], view[i], dr[i], roi=roi, use_confidence=use_conf, lam=8000.0, sigma=1.5, lrc_thresh=24, disc_radius=2)
        assert_s16_parity(out_h[i], ref["out"], roi)
        if use_conf:
            assert np.abs(conf_h[i] - ref["conf"]).max() <= 1e-4


def test_batch_half_res_device(ctx):
    import cv2
    import torch
    w, h, n = 320, 240, 3
    view, dl, dr = _frames(n, w, h, 3, seed=21)
    dl2 = np.stack([(cv2.resize(d, None, fx=0.5, fy=0.5) // 2).astype(np.int16) for d in dl])
    dr2 = np.stack([(cv2.resize(d, None, fx=0.5, fy=0.5) // 2).astype(np.int16) for d in dr])
    roi = (20, 1, w // 2 - 23, h // 2 - 2)
    dev = torch.device("cuda", 0)
    out = torch.zeros((n, h, w), dtype=torch.int16, device=dev)
    conf = torch.zeros((n, h, w), dtype=torch.float32, device=dev)
    bf = BatchFilter(True, 8000.0, 1.5, 24, 3, ctx=ctx)
    bf.filter_device(torch.from_numpy(dl2).to(dev), torch.from_numpy(view).to(dev), out, torch.from_numpy(dr2).to(dev), conf, roi)
    torch.cuda.synchronize()
    froi = (roi[0] * 2, roi[1] * 2, roi[2] * 2, roi[3] * 2)
    for i in range(n):
        ref = O.wls_filter(dl2[i], view[i], dr2[i], roi=roi, lam=8000.0, sigma=1.5, lrc_thresh=24, disc_radius=3)
        assert np.abs(conf[i].cpu().numpy() - ref["conf"]).max() <= 1e-4
        assert_s16_parity(out[i].cpu().numpy(), ref["out"], froi)


@pytest.mark.gpu
def test_graph_replay_matches_plain_launches():
    """Third and later calls with the same buffers replay a captured CUDA graph: results and kernel counts are unchanged,
    and changing an argument (lambda, ROI) never replays a stale graph."""
    import torch
    dev = torch.device("cuda", 0)
    ctx = A.Context(0)
    frames = [synth.make_artificial_example(320, 240, 3, seed=40 + i) for i in range(3)]
    roi = frames[0][3]
    view = torch.from_numpy(np.stack([f[0] for f in frames])).to(dev)
    dl = torch.from_numpy(np.stack([f[1] for f in frames])).to(dev)
    dr = torch.from_numpy(np.stack([f[2] for f in frames])).to(dev)
    out = torch.empty((3, 240, 320), dtype=torch.int16, device=dev)
    conf = torch.empty((3, 240, 320), dtype=torch.float32, device=dev)
    bf = BatchFilter(True, 8000.0, 1.5, 24, 2, ctx=ctx)
    ctx.set_cuda_graphs(False)
    bf.filter_device(dl, view, out, dr, conf, roi)          # builds the LUT for this sigma once
    n0 = ctx.kernel_launches()
    bf.filter_device(dl, view, out, dr, conf, roi)
    torch.cuda.synchronize()
    per_call = ctx.kernel_launches() - n0
    ref_out, ref_conf = out.clone(), conf.clone()
    ctx.set_cuda_graphs(True)
    for i in range(4):
        out.fill_(7); conf.fill_(-1)
        n0 = ctx.kernel_launches()
        bf.filter_device(dl, view, out, dr, conf, roi)
        torch.cuda.synchronize()
        assert ctx.kernel_launches() - n0 == per_call
        assert torch.equal(out, ref_out) and torch.equal(conf, ref_conf), i
    assert ctx.graph_replays() >= 2
    # a changed parameter is a different key
    bf2 = BatchFilter(True, 2000.0, 1.5, 24, 2, ctx=ctx)
    ctx.set_cuda_graphs(False)
    bf2.filter_device(dl, view, out, dr, conf, roi)
    torch.cuda.synchronize()
    want = out.clone()
    ctx.set_cuda_graphs(True)
    for _ in range(3):
        out.fill_(7)
        bf2.filter_device(dl, view, out, dr, conf, roi)
        torch.cuda.synchronize()
        assert torch.equal(out, want)
    assert not torch.equal(want, ref_out)
    # and the single-frame host entry point goes through the same cache
    f = A.createDisparityWLSFilterGeneric(True, ctx=ctx)
    f.setSigmaColor(1.5); f.setDepthDiscontinuityRadius(2)
    outs = [f.filter(frames[0][1], frames[0][0], None, frames[0][2], roi).copy() for _ in range(4)]
    assert all((o == outs[0]).all() for o in outs)
    assert (outs[0] == ref_out[0].cpu().numpy()).all()


def test_config5_batch_of_512_frames_1080p(ctx):
    """BASELINE.json configs[4] at full size: 512 independent 1080p stereo pairs in one device-resident batch call
    (16 launch groups of 32 frames).  The oracle needs ~75 ms per frame, so the full batch is checked through
    size-independent properties: every frame is one of four distinct stereo pairs, so (a) copies of a pair must come out
    bit-identical wherever they sit in the batch and in its launch groups, (b) the four distinct results must equal
    single-frame filter() calls bit for bit (frames are independent, disparity_filters.cpp:138-139), one of which is
    checked against the oracle, and (c) everything outside the ROI is the -16 fill / zero confidence."""
    import torch
    W, H, B, NB = 1920, 1080, 512, 4
    roi = (160, 0, 1760, 1080)
    dev = torch.device("cuda", 0)
    base = [synth.make_textured_scene(W, H, 3, seed=900 + i, max_disp=160) for i in range(NB)]
    order = np.random.default_rng(5).integers(0, NB, B)
    order[:NB] = np.arange(NB)
    idx = torch.from_numpy(order).to(dev)
    tv = torch.stack([torch.from_numpy(b[0]).to(dev) for b in base])[idx].contiguous()
    tl = torch.stack([torch.from_numpy(b[1]).to(dev) for b in base])[idx].contiguous()
    tr = torch.stack([torch.from_numpy(b[2]).to(dev) for b in base])[idx].contiguous()
    out = torch.full((B, H, W), 7, dtype=torch.int16, device=dev)
    conf = torch.full((B, H, W), -1.0, dtype=torch.float32, device=dev)
    bf = BatchFilter(True, 8000.0, 1.5, 24, 2, ctx=ctx)
    bf.filter_device(tl, tv, out, tr, conf, roi)
    torch.cuda.synchronize()
    assert tuple(bf.roi) == roi
    for k in range(NB):                                            # (a) position independence
        same = idx == k
        assert bool((out[same] == out[k]).all()) and bool((conf[same] == conf[k]).all())
    assert bool((out[:, :, :160] == -16).all()) and bool((conf[:, :, :160] == 0).all())   # (c)
    f = A.createDisparityWLSFilterGeneric(True, ctx=ctx)           # (b) batch == single calls
    f.setLambda(8000.0); f.setSigmaColor(1.5); f.setLRCthresh(24); f.setDepthDiscontinuityRadius(2)
    for k in range(NB):
        single = f.filter(base[k][1], base[k][0], None, base[k][2], roi)
        assert (out[k].cpu().numpy() == single).all()
        assert (conf[k].cpu().numpy() == f.getConfidenceMap()).all()
    ref = O.wls_filter(base[0][1], base[0][0], base[0][2], roi=roi, lam=8000.0, sigma=1.5, lrc_thresh=24, disc_radius=2)
    assert np.abs(conf[0].cpu().numpy() - ref["conf"]).max() <= 1e-4
    assert_s16_parity(out[0].cpu().numpy(), ref["out"], roi)


def test_calls_on_different_streams_do_not_race_on_the_shared_workspace(ctx):
    """Two asynchronous device calls on different torch streams share ctx workspace, LUT and graphs: the library orders
    the second after the first (event recorded by each call), so both give the bytes of serialised runs even with
    different sigma (the LUT is rewritten in place) and different batch sizes (other workspace layout)."""
    import torch
    dev = torch.device("cuda", 0)
    w, h = 640, 480
    roi = (80, 0, w - 80, h)
    va, la, ra = _frames(12, w, h, 3, seed=300)
    vb, lb, rb = _frames(5, w, h, 3, seed=400)
    ta = [torch.from_numpy(x).to(dev) for x in (la, va, ra)]
    tb = [torch.from_numpy(x).to(dev) for x in (lb, vb, rb)]
    fa = BatchFilter(True, 8000.0, 1.5, 24, 2, ctx=ctx)
    fb = BatchFilter(True, 500.0, 4.0, 24, 3, ctx=ctx)
    want = []
    for f, t in ((fa, ta), (fb, tb)):
        o = torch.empty((t[0].shape[0], h, w), dtype=torch.int16, device=dev)
        f.filter_device(t[0], t[1], o, t[2], None, roi)
        torch.cuda.synchronize()
        want.append(o.clone())
    sa, sb = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    for rep in range(6):
        oa = torch.full((12, h, w), 7, dtype=torch.int16, device=dev)
        ob = torch.full((5, h, w), 7, dtype=torch.int16, device=dev)
        torch.cuda.synchronize()
        with torch.cuda.stream(sa):
            fa.filter_device(ta[0], ta[1], oa, ta[2], None, roi)
        with torch.cuda.stream(sb):
            fb.filter_device(tb[0], tb[1], ob, tb[2], None, roi)       # no host sync in between
        with torch.cuda.stream(sa):
            fa.filter_device(ta[0], ta[1], oa, ta[2], None, roi)
        torch.cuda.synchronize()
        assert torch.equal(oa, want[0]) and torch.equal(ob, want[1]), rep
