"""GPU parity tests: the CUDA path (through the C-ABI) against the CPU oracle on the same inputs.

Gates (BASELINE.json north_star): ROI and LRC mask bit-exact; confidence map within 1e-4
absolute; CV_16S filtered disparity within 1 LSB on >= 99.9 % of pixels; outside-ROI fill exact.
The envelopes of the reference's own tests are re-stated where they apply
(test_fgs_filter.cpp:59-87, 109-151; test_disparity_wls_filter.cpp:99-153).
"""
import numpy as np
import pytest

import addingdisparityfiltering_b200 as A
from addingdisparityfiltering_b200 import synth
from oracle import oracle as O
from tests.parity import assert_s16_parity, s16_stats

pytestmark = pytest.mark.gpu


def _guide(rng, h, w, cn):
    g = rng.integers(0, 256, (h, w, cn)).astype(np.float32)
    g = (g + np.roll(g, 1, 0) + np.roll(g, 1, 1)) / 3
    g = g.astype(np.uint8)
    return g[:, :, 0] if cn == 1 else g


def test_lut_and_weights_match_oracle(ctx):
    rng = np.random.default_rng(1)
    for cn in (1, 3):
        g = _guide(rng, 61, 127, cn)
        for sigma in (0.5, 1.5, 10.0):
            ch, cv = A.fgs_weights(g, sigma, ctx=ctx)
            oh, ov = O.fgs_weights(g, sigma)
            # LUT: float sqrt/div are IEEE on both sides; exp is correctly rounded here, <=0.502 ulp in glibc
            for a, b in ((ch, oh), (cv, ov)):
                rel = np.abs(a - b) / np.maximum(np.abs(b), 1e-30)
                assert (a == b).mean() > 0.99
                assert rel.max() < 2.5e-7


@pytest.mark.parametrize("size", [(127, 61), (320, 240), (640, 480), (33, 9), (8, 8), (1000, 700)])
@pytest.mark.parametrize("cn", [1, 3])
def test_fgs_float_matches_oracle(ctx, size, cn):
    w, h = size
    rng = np.random.default_rng(w * 7 + h + cn)
    g = _guide(rng, h, w, cn)
    src = (rng.random((h, w)) * 255).astype(np.float32)
    for lam, sigma in ((1000.0, 10.0), (8000.0, 1.5), (100.0, 50.0)):
        got = A.fastGlobalSmootherFilter(g, src, lam, sigma, ctx=ctx)
        ref = O.fgs_filter(g, src, lam, sigma)
        err = np.abs(got - ref)
        # the reference's own float32 solve sits ~3e-4 of the range away from the exact (float64) solution at
        # lambda=8000 (condition number ~4*lambda); 1e-3 of the range bounds GPU-vs-oracle rounding differences
        assert err.max() <= 255 * 1e-3, (lam, sigma, err.max())


def test_fgs_constant_image_invariance(ctx):
    # FastGlobalSmootherTest.SplatSurfaceAccuracy (test_fgs_filter.cpp:59-87): mean abs err <= 1/64
    rng = np.random.default_rng(3)
    for i in range(3):
        w, h = int(rng.integers(512, 1024)), int(rng.integers(512, 1024))
        cn = 1 if i % 2 else 3
        g = _guide(rng, h, w, cn)
        val = int(rng.integers(-32000, 32000))
        src = np.full((h, w), val, np.int16)
        out = A.fastGlobalSmootherFilter(g, src, 1000.0, 10.0, ctx=ctx)
        assert np.abs(out.astype(np.int64) - val).mean() <= 1.0 / 64


@pytest.mark.parametrize("dtype,cn", [(np.uint8, 1), (np.uint8, 3), (np.uint8, 4), (np.int16, 1), (np.int16, 3), (np.float32, 1), (np.float32, 2)])
def test_fgs_type_matrix(ctx, dtype, cn):
    # src types of FastGlobalSmootherTest.MultiThreadReproducibility (test_fgs_filter.cpp:54, 127-132)
    rng = np.random.default_rng(11 + cn)
    w, h = 320, 240
    g = _guide(rng, h, w, 3)
    shape = (h, w) if cn == 1 else (h, w, cn)
    if dtype == np.uint8:
        src = rng.integers(0, 256, shape).astype(np.uint8)
    elif dtype == np.int16:
        src = rng.integers(-32767, 32767, shape).astype(np.int16)
    else:
        src = ((rng.random(shape) * 2 - 1) * 1e5).astype(np.float32)
    lam, sigma = 5000.0, 20.0
    got = A.fastGlobalSmootherFilter(g, src, lam, sigma, ctx=ctx)
    ref = O.fgs_filter(g, src, lam, sigma)
    assert got.dtype == src.dtype and got.shape == src.shape
    if dtype == np.float32:
        assert np.abs(got - ref).max() <= 2e5 * 1e-3
    else:
        d = np.abs(got.astype(np.int64) - ref.astype(np.int64))
        assert d.max() <= 1 and d.sum() <= d.size / 64          # L-inf <= 1, L1 <= total/64 (114-115)


def _check_wls(ctx, view, dl, dr, roi, use_conf, lam=8000.0, sigma=1.5, radius=2, lrc=24):
    f = A.createDisparityWLSFilterGeneric(use_conf, ctx=ctx)
    f.setLambda(lam); f.setSigmaColor(sigma); f.setDepthDiscontinuityRadius(radius); f.setLRCthresh(lrc)
    got = f.filter(dl, view, None, dr, roi)
    ref = O.wls_filter(dl, view, dr, roi=roi, use_confidence=use_conf, lam=lam, sigma=sigma, lrc_thresh=lrc, disc_radius=radius, want_mask=True)
    assert ref["rc"] == 0
    assert tuple(f.getROI()) == tuple(ref["roi"])                       # ROI bit-exact
    H, W = view.shape[:2]
    dh, dw = dl.shape
    sx, sy = W / np.float32(dw), H / np.float32(dh)
    x, y, w, h = ref["roi"]
    froi = (int(np.float32(x) * sx), int(np.float32(y) * sy), int(np.float32(w) * sx), int(np.float32(h) * sy)) if (W, H) != (dw, dh) else ref["roi"]
    outside = np.ones((H, W), bool)
    outside[froi[1]:froi[1] + froi[3], froi[0]:froi[0] + froi[2]] = False
    assert (got[outside] == -16).all() and (ref["out"][outside] == -16).all()   # fill exact
    if use_conf:
        mask = f.lrc_mask(dl, dr, view_cols=W, ROI=roi)
        assert (mask == ref["mask"]).all()                                # LRC decisions bit-exact
        conf = f.getConfidenceMap()
        assert conf.shape == (H, W)
        assert np.abs(conf - ref["conf"]).max() <= 1e-4                   # confidence within 1e-4 absolute (0..255 scale)
    return assert_s16_parity(got, ref["out"], froi)


@pytest.mark.parametrize("size", [(127, 61), (320, 240), (640, 480)])
@pytest.mark.parametrize("cn", [1, 3])
@pytest.mark.parametrize("use_conf", [True, False])
def test_wls_full_res_matches_oracle(ctx, size, cn, use_conf):
    w, h = size
    view, dl, dr, roi = synth.make_artificial_example(w, h, cn, seed=w + h + cn)
    _check_wls(ctx, view, dl, dr, roi, use_conf, radius=5)
    view, dl, dr = synth.make_textured_scene(w, h, cn, seed=w + cn, max_disp=max(32, w // 8))
    md = max(32, w // 8)
    _check_wls(ctx, view, dl, dr, (md, 0, w - md, h), use_conf, radius=2)


@pytest.mark.parametrize("size", [(127, 61), (320, 240), (640, 480)])
@pytest.mark.parametrize("use_conf", [True, False])
def test_wls_half_res_matches_oracle(ctx, size, use_conf):
    import cv2
    w, h = size
    view, dl, dr, roi = synth.make_artificial_example(w, h, 3, seed=5 * w + h)
    dl2 = (cv2.resize(dl, None, fx=0.5, fy=0.5) // 2).astype(np.int16)
    dr2 = (cv2.resize(dr, None, fx=0.5, fy=0.5) // 2).astype(np.int16)
    roi2 = (roi[0] // 2, roi[1] // 2, roi[2] // 2, roi[3] // 2)
    _check_wls(ctx, view, dl2, dr2, roi2, use_conf, radius=3)


def test_wls_1080p_config2(ctx):
    # BASELINE.json configs[1]: 1920x1080, SGBM-style geometry (ROI (160,0,1760,1080), radius 2, LRC 24)
    view, dl, dr = synth.make_textured_scene(1920, 1080, 3, seed=42, max_disp=160)
    st = _check_wls(ctx, view, dl, dr, (160, 0, 1760, 1080), True, lam=8000.0, sigma=1.5, radius=2)
    print("cfg2 parity", st)


def test_wls_matcher_factory_geometry(ctx):
    class FakeBM:
        kind = "bm"
        def __init__(s): s.calls = {}
        def getMinDisparity(s): return 0
        def getNumDisparities(s): return 80
        def getBlockSize(s): return 7
        def __getattr__(s, n):
            if n.startswith("set"):
                return lambda v: s.calls.__setitem__(n, v)
            raise AttributeError(n)
    m = FakeBM()
    f = A.createDisparityWLSFilter(m, ctx=ctx)
    assert m.calls == {"setDisp12MaxDiff": 1000000, "setSpeckleWindowSize": 0, "setTextureThreshold": 0, "setUniquenessRatio": 0}
    assert f.getDepthDiscontinuityRadius() == 3 and f.getLRCthresh() == 24 and f.getLambda() == 8000.0 and f.getSigmaColor() == 1.0
    view, dl, dr, _ = synth.make_artificial_example(320, 240, 3, seed=9)
    f.filter(dl, view, None, dr)
    assert f.getROI() == (83, 3, 234, 234)                               # SURVEY section 8 config 1 geometry


def test_error_behaviour(ctx):
    view, dl, dr, roi = synth.make_artificial_example(64, 48, 3, seed=1)
    f = A.createDisparityWLSFilterGeneric(True, ctx=ctx)
    with pytest.raises(A.DwlsError):
        f.filter(dl.astype(np.float32), view, None, dr, roi)             # CV_Assert depth (221)
    with pytest.raises(A.DwlsError):
        f.filter(dl, view, None, None, roi)                              # right map required (262)
    with pytest.raises(A.DwlsError):
        f.filter(dl, view, None, dr[:, :-1], roi)                        # size mismatch (263)
    with pytest.raises(A.DwlsError) as e:
        A.fastGlobalSmootherFilter(view, np.zeros((10, 10), np.float32), 100.0, 1.0, ctx=ctx)
    assert e.value.status == 2                                           # StsBadSize (185-189)
    with pytest.raises(A.DwlsError):
        A.fastGlobalSmootherFilter(view, dl, -1.0, 1.0, ctx=ctx)         # lambda >= 0 (143)


@pytest.mark.parametrize("radius", [0, 1, 2, 3, 4, 5, 6, 7, 8, 9])
@pytest.mark.parametrize("size", [(320, 96), (127, 61)])          # 16-byte-aligned rows (TMA tile path) / unaligned rows (plain-load path)
def test_confidence_stage_radii_and_alignments(ctx, radius, size):
    # box-variance + LRC stage alone (disparity_filters.cpp:161-210, 306-373): every radius of the register kernel (0..7),
    # the shared-memory kernel (8, 9), ROI origins at every residue mod 8 (in-place realignment after the aligned TMA
    # fetch), ROIs touching the image border and ROIs narrower/shorter than the halo (multi-bounce reflect-101)
    w, h = size
    view, dl, dr, _ = synth.make_artificial_example(w, h, 1, seed=31 * radius + w)
    rng = np.random.default_rng(radius + w)
    dl = (dl.astype(np.int32) + rng.integers(-40, 40, dl.shape)).astype(np.int16)
    dr = (dr.astype(np.int32) + rng.integers(-40, 40, dr.shape)).astype(np.int16)
    rois = [(0, 0, w, h), (3 + radius, 1, w - 11 - radius, h - 2), (w // 3 + 5, h // 4, w // 2 + 2, h // 2 + 1), (w - 13, h - 7, 13, 7),
            (17, 5, 3, 2), (40, 2, w - 47, 1)]
    f = A.createDisparityWLSFilterGeneric(True, ctx=ctx)
    f.setDepthDiscontinuityRadius(radius)
    for roi in rois:
        got = f.filter(dl, view, None, dr, roi)
        ref = O.wls_filter(dl, view, dr, roi=roi, lam=8000.0, sigma=1.0, lrc_thresh=24, disc_radius=radius, want_mask=True)
        assert ref["rc"] == 0
        assert (f.lrc_mask(dl, dr, ROI=roi) == ref["mask"]).all(), (radius, roi)
        assert np.abs(f.getConfidenceMap() - ref["conf"]).max() <= 1e-4, (radius, roi)
        assert_s16_parity(got, ref["out"], roi)


@pytest.mark.parametrize("radius", [1, 2, 3])
@pytest.mark.parametrize("size", [(960, 80), (1003, 70)])           # aligned rows (TMA tile path) / odd row length (plain-load path)
def test_lrc_wide_rois_large_disparities(ctx, radius, size):
    # LRC check (disparity_filters.cpp:306-341) on ROIs several tiles wide: ROI origins at every residue mod 8 (parity of the
    # gathered right-disparity words, alignment of the gathered confidences), ROIs ending at the image's right / left edge,
    # rows with disparities of -64 .. 399 px (right columns far from the pixel's own tile), rows of arbitrary int16 values
    # (mostly misses).  Written for the shared-memory window variant of the LRC step (exp/lrc_window_mode.patch, dropped:
    # profiles/README.md); kept because nothing else drives the gathers this far from the tile.
    w, h = size
    view, dl, dr, _ = synth.make_artificial_example(w, h, 1, seed=7 * radius + w)
    rng = np.random.default_rng(radius * 1000 + w)
    dl = (dl.astype(np.int32) + rng.integers(-40, 40, dl.shape)).astype(np.int16)
    dr = (dr.astype(np.int32) + rng.integers(-40, 40, dr.shape)).astype(np.int16)
    dl[10:20] = (rng.integers(-64, 400, (10, w)) * 16).astype(np.int16)          # -64 .. 399 px: partly outside the window
    dl[30:34] = rng.integers(-32768, 32768, (4, w)).astype(np.int16)             # mostly misses
    dl[40:44, ::3] = 300 * 16
    rois = [(0, 0, w, h)] + [(k, 1, w - 2 * k - (k & 1), h - 2) for k in range(1, 9)] + [(w - 416, 3, 416, h - 5), (0, 0, 417, h), (37, 2, 421, 33)]
    f = A.createDisparityWLSFilterGeneric(True, ctx=ctx)
    f.setDepthDiscontinuityRadius(radius)
    for roi in rois:
        got = f.filter(dl, view, None, dr, roi)
        ref = O.wls_filter(dl, view, dr, roi=roi, lam=8000.0, sigma=1.0, lrc_thresh=24, disc_radius=radius, want_mask=True)
        assert ref["rc"] == 0
        assert (f.lrc_mask(dl, dr, ROI=roi) == ref["mask"]).all(), (radius, roi)
        assert np.abs(f.getConfidenceMap() - ref["conf"]).max() <= 1e-4, (radius, roi)
        assert_s16_parity(got, ref["out"], roi)


def test_confidence_extreme_values(ctx):
    # |d| up to 32767 / -32768: window sums of squares need more than 32 bits; the LRC sum L + R must not wrap
    w, h = 320, 64
    rng = np.random.default_rng(5)
    view = rng.integers(0, 256, (h, w)).astype(np.uint8)
    dl = rng.integers(-32768, 32768, (h, w)).astype(np.int16)
    dr = rng.integers(-32768, 32768, (h, w)).astype(np.int16)
    dl[:8] = -32768; dr[:8] = -32768; dl[8:16] = 32767; dr[8:16] = -32767
    f = A.createDisparityWLSFilterGeneric(True, ctx=ctx)
    for radius in (2, 7):
        f.setDepthDiscontinuityRadius(radius)
        roi = (8, 0, w - 16, h)
        f.filter(dl, view, None, dr, roi)
        ref = O.wls_filter(dl, view, dr, roi=roi, lam=8000.0, sigma=1.0, lrc_thresh=24, disc_radius=radius, want_mask=True)
        assert (f.lrc_mask(dl, dr, ROI=roi) == ref["mask"]).all()
        assert np.abs(f.getConfidenceMap() - ref["conf"]).max() <= 1e-4


@pytest.mark.parametrize("size", [(2500, 40), (4096, 33), (3840, 48), (5000, 24), (8192, 16), (64, 2000), (40, 1153), (200, 2176),
                                  (64, 2177), (37, 3000), (130, 4100), (24, 8192)])
@pytest.mark.parametrize("planes", [1, 2])
def test_fgs_wide_and_tall_images(ctx, size, planes):
    # rows wider than 2048 take the 16-elements-per-thread row pass (> 4096: 512 threads, looped level 2);
    # more than 1152 rows take the 17-rows-per-thread column pass, more than 2176 rows the transposed column pass
    w, h = size
    rng = np.random.default_rng(w + h)
    g = _guide(rng, h, w, 3)
    shape = (h, w) if planes == 1 else (h, w, planes)
    src = (rng.random(shape) * 255).astype(np.float32)
    got = A.fastGlobalSmootherFilter(g, src, 8000.0, 1.5, ctx=ctx)
    ref = O.fgs_filter(g, src, 8000.0, 1.5)
    assert np.abs(got - ref).max() <= 255 * 1e-3


@pytest.mark.parametrize("conf", [True, False])
def test_wls_tall_roi(ctx, conf):
    # ROI taller than 2176 rows: column sweeps run over transposed planes, the normalisation is fused into the transpose back
    w, h = 96, 2500
    view, dl, dr, roi = synth.make_artificial_example(w, h, 3, seed=3)
    f = A.createDisparityWLSFilterGeneric(conf, ctx=ctx)
    f.setSigmaColor(1.5)
    got = f.filter(dl, view, None, dr if conf else None, roi)
    ref = O.wls_filter(dl, view, dr if conf else None, roi=roi, lam=8000.0, sigma=1.5, use_confidence=conf)
    assert ref["rc"] == 0
    if conf:
        assert np.abs(f.getConfidenceMap() - ref["conf"]).max() <= 1e-4
    assert_s16_parity(got, ref["out"], roi)


def test_size_limit_is_reported(ctx):
    g = np.zeros((8, 8200), np.uint8)
    with pytest.raises(A.DwlsError) as e:
        A.fastGlobalSmootherFilter(g, g.astype(np.float32), 100.0, 1.0, ctx=ctx)
    assert e.value.status == 4


@pytest.mark.parametrize("size", [(1, 1), (7, 1), (1, 7), (2, 2), (3, 5), (5, 3), (9, 2), (33, 1), (1, 40)])
@pytest.mark.parametrize("cn", [1, 3])
def test_fgs_tiny_images(ctx, size, cn):
    # degenerate lines: a single row has no vertical neighbour (fgs_filter.cpp:637 would read row 1; the last weight of a
    # line is 0, 614/658-660), rows narrower than the reference's 4-wide SIMD blocks (316, 393)
    w, h = size
    rng = np.random.default_rng(10 * w + h)
    g = _guide(rng, h, w, cn)
    src = (rng.random((h, w)) * 255).astype(np.float32)
    got = A.fastGlobalSmootherFilter(g, src, 900.0, 3.0, ctx=ctx)
    ref = O.fgs_filter(g, src, 900.0, 3.0)
    assert got.shape == ref.shape and np.isfinite(got).all()
    assert np.abs(got - ref).max() <= 255 * 1e-3


@pytest.mark.parametrize("roi", [(0, 0, 64, 48), (5, 0, 1, 48), (0, 7, 64, 1), (63, 47, 1, 1), (10, 10, 2, 3)])
def test_wls_roi_touching_borders_and_slivers(ctx, roi):
    # ROI equal to the image, one column, one row, one pixel in the corner: the box filter reflects inside the crop
    # (disparity_filters.cpp:168-174) and every line solve degenerates
    w, h = 64, 48
    view, dl, dr, _ = synth.make_artificial_example(w, h, 3, seed=21)
    f = A.createDisparityWLSFilterGeneric(True, ctx=ctx)
    f.setDepthDiscontinuityRadius(2)
    got = f.filter(dl, view, None, dr, roi)
    ref = O.wls_filter(dl, view, dr, roi=roi, lam=8000.0, sigma=1.0, disc_radius=2, want_mask=True)
    assert ref["rc"] == 0 and tuple(f.getROI()) == tuple(ref["roi"])
    assert (f.lrc_mask(dl, dr, ROI=roi) == ref["mask"]).all()
    assert np.abs(f.getConfidenceMap() - ref["conf"]).max() <= 1e-4
    assert_s16_parity(got, ref["out"], roi)


def test_wls_zero_confidence_everywhere(ctx):
    # left and right maps disagree at every pixel -> confidence 0 everywhere -> the reference divides 0 by (0 + 1e-43f):
    # 0 * (1 / 1e-43f) = 0 * inf = NaN -> saturate_cast<short>(NaN) = -32768 (disparity_filters.cpp:295-296)
    w, h = 96, 40
    rng = np.random.default_rng(4)
    view = rng.integers(0, 256, (h, w, 3)).astype(np.uint8)
    dl = np.full((h, w), 16 * 4, np.int16)
    dr = np.full((h, w), 16 * 40, np.int16)          # L + R far above the threshold
    roi = (8, 0, w - 8, h)
    f = A.createDisparityWLSFilterGeneric(True, ctx=ctx)
    got = f.filter(dl, view, None, dr, roi)
    ref = O.wls_filter(dl, view, dr, roi=roi, lam=8000.0, sigma=1.0)
    x, y, rw, rh = roi
    hit = ref["conf"][y:y + rh, x:x + rw] == 0
    assert hit.any()
    assert (f.getConfidenceMap() == ref["conf"]).all()
    assert (got == ref["out"]).all(), (np.unique(got), np.unique(ref["out"]))


@pytest.mark.parametrize("case", ["full", "half", "odd_rows", "narrow", "radius9"])
def test_plane_layouts_give_identical_results(ctx, case):
    """The two solver planes of the confidence path travel between the sweeps either as two row-major planes or as one
    sector-interleaved array (dwls_set_plane_interleave): same arithmetic per thread, so the outputs must be bit-identical,
    and both are checked against the oracle."""
    import cv2
    w, h, radius = {"full": (640, 480, 2), "half": (640, 480, 3), "odd_rows": (328, 123, 5), "narrow": (45, 31, 1), "radius9": (320, 96, 9)}[case]
    view, dl, dr, roi = synth.make_artificial_example(w, h, 3, seed=11 * w + h)
    if case == "half":
        dl = (cv2.resize(dl, None, fx=0.5, fy=0.5) // 2).astype(np.int16)
        dr = (cv2.resize(dr, None, fx=0.5, fy=0.5) // 2).astype(np.int16)
        roi = (roi[0] // 2, roi[1] // 2, roi[2] // 2, roi[3] // 2)
    outs, confs = [], []
    try:
        for on in (False, True):
            ctx.set_plane_interleave(on)
            _check_wls(ctx, view, dl, dr, roi, True, radius=radius)
            f = A.createDisparityWLSFilterGeneric(True, ctx=ctx)
            f.setSigmaColor(1.5); f.setDepthDiscontinuityRadius(radius)
            outs.append(f.filter(dl, view, None, dr, roi).copy())
            confs.append(f.getConfidenceMap().copy())
    finally:
        ctx.set_plane_interleave(False)          # the layout the product ships with
    assert (outs[0] == outs[1]).all()
    assert (confs[0] == confs[1]).all()


def test_wls_zero_confidence_regions_next_to_sharp_edges(ctx):
    """Half of the ROI fails the LRC check under a noisy guide (weights ~ 0), so the filtered confidence stays 0 or
    subnormal over whole regions: 1 / (c~ + 1e-43f) overflows there and the reference writes -32768
    (disparity_filters.cpp:295-296).  The final kernel takes that range without a branch (scaled reciprocal)."""
    rng = np.random.default_rng(3)
    w, h = 640, 480
    view = rng.integers(0, 256, (h, w, 3)).astype(np.uint8)
    dl = (np.full((h, w), 160) + rng.integers(-3, 4, (h, w))).astype(np.int16)
    dr = (-dl).astype(np.int16)
    dr[:, : w // 2] = 5000
    roi = (32, 0, w - 32, h)
    st = _check_wls(ctx, view, dl, dr, roi, True, radius=2)
    ref = O.wls_filter(dl, view, dr, roi=roi, lam=8000.0, sigma=1.5, disc_radius=2)
    assert (ref["out"][:, 32:] == -32768).mean() > 0.3          # the case really exercises the overflow range
    print("zero-confidence regions parity", st)


def test_large_strided_host_images_take_the_staged_transfer_path(ctx):
    """Host images above 2 MB travel through pinned staging filled by several host threads (dwls_filter_host / dwls_fgs_host).
    Row-strided inputs and outputs (cv::Mat ROI views inside larger allocations), passed straight to the C-ABI, must give
    the same bytes as dense ones."""
    import ctypes as C
    from addingdisparityfiltering_b200 import _lib
    from addingdisparityfiltering_b200.ximgproc import _params
    w, h = 1280, 720
    view, dl, dr, roi = synth.make_artificial_example(w, h, 3, seed=77)

    def strided(a, pad):
        big = np.zeros((a.shape[0] + 3, a.shape[1] + 5 + pad) + a.shape[2:], a.dtype)
        v = big[2:2 + a.shape[0], 5:5 + a.shape[1]]
        v[...] = a
        assert not v.flags["C_CONTIGUOUS"]
        return v

    def run(dl_, dr_, view_, out_, conf_):
        p = _params(8000.0, 1.5, 24, 3, (0, 0, 0, 0), True)
        imgs = [_lib.image_of(x) for x in (dl_, dr_, view_, out_, conf_)]
        roi_in, roi_out = (C.c_int * 4)(*roi), (C.c_int * 4)()
        ctx.check(_lib.lib().dwls_filter_host(ctx.handle, C.byref(imgs[0]), C.byref(imgs[1]), C.byref(imgs[2]), roi_in, C.byref(p),
                                              C.byref(imgs[3]), C.byref(imgs[4]), roi_out))
    out_d, conf_d = np.zeros((h, w), np.int16), np.zeros((h, w), np.float32)
    run(dl, dr, view, out_d, conf_d)
    out_s, conf_s = strided(np.zeros((h, w), np.int16), 11), strided(np.zeros((h, w), np.float32), 3)
    run(strided(dl, 9), strided(dr, 13), strided(view, 7), out_s, conf_s)
    assert (out_s == out_d).all() and (conf_s == conf_d).all()
    ref = O.wls_filter(dl, view, dr, roi=roi, lam=8000.0, sigma=1.5, disc_radius=3)
    assert np.abs(conf_d - ref["conf"]).max() <= 1e-4
    assert_s16_parity(out_d, ref["out"], roi)
    # standalone FGS, 8UC3 source of 2.7 MB: dense vs strided guide / source / destination
    lib = _lib.lib()
    dst_d, dst_s = np.zeros_like(view), strided(np.zeros_like(view), 5)
    g_s, s_s = strided(view, 7), strided(view, 3)
    for g_, s_, d_ in ((view, view, dst_d), (g_s, s_s, dst_s)):
        ig, is_, id_ = _lib.image_of(g_), _lib.image_of(s_), _lib.image_of(d_)
        ctx.check(lib.dwls_fgs_host(ctx.handle, C.byref(ig), C.byref(is_), C.byref(id_), C.c_double(1000.0), C.c_double(10.0), C.c_double(0.25), 3))
    assert (dst_d == dst_s).all()
