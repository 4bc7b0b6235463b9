"""Device stereo block matcher against the real cv2.StereoBM.compute: bit-exact CV_16S maps, for the matcher
pair the WLS factories configure (samples/disparity_filtering.cpp:148-160, disparity_filters.cpp:389-403, 419-428) and for
OpenCV's default texture / uniqueness tests; then the whole device-born pipeline match -> filter."""
import numpy as np
import pytest

import addingdisparityfiltering_b200 as A
from addingdisparityfiltering_b200 import synth
from tests.parity import record
from tests.test_stereo_model import BM_CASES, cv2_bm, gray_pair

pytestmark = pytest.mark.gpu


def _ours(nd, ws, md, tex, un, ctx):
    m = A.StereoBM.create(nd, ws, ctx=ctx)
    m.setMinDisparity(md); m.setTextureThreshold(tex); m.setUniquenessRatio(un); m.setDisp12MaxDiff(1000000); m.setSpeckleWindowSize(0)
    return m


@pytest.mark.parametrize("case", BM_CASES)
@pytest.mark.parametrize("size", [(640, 480), (320, 240), (203, 97)])
def test_device_stereobm_equals_cv2(ctx, case, size):
    nd, ws, md, tex, un = case
    gl, gr, _, _ = gray_pair(size[0], size[1], seed=3 + ws, max_disp=min(60, nd - 4))
    a, b = (gl, gr) if md >= 0 else (gr, gl)
    want = cv2_bm(nd, ws, md, tex, un).compute(a, b)
    got = _ours(nd, ws, md, tex, un, ctx).compute(a, b)
    assert got.dtype == np.int16 and got.shape == want.shape
    assert (got == want).all(), (int((got != want).sum()), np.argwhere(got != want)[:5])


@pytest.mark.parametrize("size,nd,ws", [((960, 540), 80, 7), ((1920, 1080), 160, 15), ((640, 480), 160, 15), ((1024, 436), 80, 7)])
def test_device_matcher_pair_equals_cv2_pair(ctx, size, nd, ws):
    """Left matcher + createRightMatcher's matcher in one launch, on colour views (BGR -> gray on the device) at the sizes
    of the sample's paths: 1080p halved, 1080p --no-downscale, 640x480 --no-downscale, the tutorial pair's size."""
    cv2 = pytest.importorskip("cv2")
    left, right, _ = synth.make_stereo_pair(size[0], size[1], seed=11, max_disp=min(140, nd - 8))
    lm = cv2.StereoBM_create(nd, ws)
    A.createDisparityWLSFilter(lm, ctx=ctx)                       # switches texture / uniqueness / disp12 / speckle off (389-400)
    rm = A.createRightMatcher(lm)
    gl, gr = cv2.cvtColor(left, cv2.COLOR_BGR2GRAY), cv2.cvtColor(right, cv2.COLOR_BGR2GRAY)
    want_l, want_r = lm.compute(gl, gr), rm.compute(gr, gl)
    ours = A.StereoBM.create(nd, ws, ctx=ctx)
    A.createDisparityWLSFilter(ours, ctx=ctx)
    got_l, got_r = ours.compute(left, right, with_right=True)     # colour in: the conversion is part of the device path
    assert (got_l == want_l).all() and (got_r == want_r).all()
    ours_r = A.createRightMatcher(ours)                           # and the right matcher object on its own
    assert isinstance(ours_r, A.StereoBM) and ours_r.getMinDisparity() == -nd + 1
    assert (ours_r.compute(gr, gl) == want_r).all()
    record("stereobm_%dx%d_nd%d_w%d_vs_cv2" % (size[0], size[1], nd, ws), identical=1.0, invalid_share=float((want_l < 0).mean()))


def test_device_batch_and_halving(ctx):
    torch = pytest.importorskip("torch")
    cv2 = pytest.importorskip("cv2")
    from addingdisparityfiltering_b200.stereo import halve_views_device
    dev = torch.device("cuda", 0)
    pairs = [synth.make_stereo_pair(640, 480, seed=20 + i, max_disp=70) for i in range(3)]
    tl = torch.from_numpy(np.stack([p[0] for p in pairs])).to(dev)
    tr = torch.from_numpy(np.stack([p[1] for p in pairs])).to(dev)
    hl, hr = halve_views_device(tl, tr, ctx=ctx)
    torch.cuda.synchronize()
    for i, p in enumerate(pairs):
        assert (hl[i].cpu().numpy() == cv2.resize(p[0], None, fx=0.5, fy=0.5)).all()
        assert (hr[i].cpu().numpy() == cv2.resize(p[1], None, fx=0.5, fy=0.5)).all()
    m = A.StereoBM.create(80, 7, ctx=ctx)
    A.createDisparityWLSFilter(m, ctx=ctx)
    dl = torch.empty((3, 240, 320), dtype=torch.int16, device=dev)
    dr = torch.empty((3, 240, 320), dtype=torch.int16, device=dev)
    m.compute_device(hl, hr, dl, dr)
    torch.cuda.synchronize()
    for i in range(3):
        wl, wr = m.compute(hl[i].cpu().numpy(), hr[i].cpu().numpy(), with_right=True)
        assert (dl[i].cpu().numpy() == wl).all() and (dr[i].cpu().numpy() == wr).all()


def test_errors(ctx):
    g = np.zeros((64, 64), np.uint8)
    for nd, ws, md in ((24, 7, 0), (32, 6, 0), (32, 3, 0), (32, 7, 5)):
        m = A.StereoBM.create(nd, ws, ctx=ctx)
        m.setMinDisparity(md)
        with pytest.raises(A.DwlsError):
            m.compute(g, g)
    m = A.StereoBM.create(32, 7, ctx=ctx)
    with pytest.raises(A.DwlsError):
        m.compute(g, g[:, :-1])
    m.setSpeckleWindowSize(100)
    with pytest.raises(A.DwlsError):
        m.compute(g, g)
    wide = A.StereoBM.create(128, 7, ctx=ctx)                     # more disparities than columns: everything FILTERED, like cv2
    assert (wide.compute(g, g) == -16).all()


@pytest.mark.parametrize("downscale", [True, False])
def test_match_and_filter_pipeline_equals_the_cv2_matcher_pipeline(ctx, downscale):
    """configs[0] with the matcher on the device: the sample's wls_conf path (samples/disparity_filtering.cpp:128-190) with
    A.StereoBM in place of cv2.StereoBM gives the same filtered map, confidence map and ROI bit for bit."""
    cv2 = pytest.importorskip("cv2")
    from tests.stereo_inputs import sample_matching
    left, right, _ = synth.make_stereo_pair(640, 480, seed=42, max_disp=140)
    f_cv, dl_cv, dr_cv = sample_matching(left, right, "bm", downscale=downscale, max_disp=160, ctx=ctx)
    f_cv.setLambda(8000.0); f_cv.setSigmaColor(1.5)
    want = f_cv.filter(dl_cv, left, None, dr_cv)
    max_disp, wsize = (80, 7) if downscale else (160, 15)
    lm, rm = (cv2.resize(left, None, fx=0.5, fy=0.5), cv2.resize(right, None, fx=0.5, fy=0.5)) if downscale else (left, right)
    matcher = A.StereoBM.create(max_disp, wsize, ctx=ctx)
    f = A.createDisparityWLSFilter(matcher, ctx=ctx)
    dl, dr = matcher.compute(lm, rm, with_right=True)
    assert (dl == dl_cv).all() and (dr == dr_cv).all()
    f.setLambda(8000.0); f.setSigmaColor(1.5)
    got = f.filter(dl, left, None, dr)
    assert f.getROI() == f_cv.getROI() and (got == want).all() and (f.getConfidenceMap() == f_cv.getConfidenceMap()).all()
