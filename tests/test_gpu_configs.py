"""GPU parity at the FULL sizes of every BASELINE.json config and on inputs from REAL stereo matchers.

configs[0]  640x480 StereoBM sample path (downscale and --no-downscale)       samples/disparity_filtering.cpp:128-190
configs[1]  1920x1080 StereoSGBM disparities, confidence + LRC                 disparity_filters.cpp:404-408
configs[2]  fastGlobalSmootherFilter on 3840x2160, 3-channel guide             fgs_filter.cpp:182-233
configs[3]  half-resolution filtering with upscaled output at 1080p            disparity_filters.cpp:268-277
plus the reference's own tutorial pair (ambush_5) against outputs of the compiled reference, and upscale ratios
other than 2x (resize_factor is arbitrary, disparity_filters.cpp:224-227).
Every case records its measured maxima (tests/parity.record -> gpurun_out/parity_stats.jsonl -> BASELINE.md section 3).
"""
import numpy as np
import pytest

import addingdisparityfiltering_b200 as A
from addingdisparityfiltering_b200 import synth
from oracle import oracle as O
from tests.parity import assert_s16_parity, record, s16_stats
from tests.stereo_inputs import ambush_fixture, sample_matching

pytestmark = pytest.mark.gpu


def _view_roi(roi, dw, dh, W, H):
    """disparity_filters.cpp:245-246: the ROI scaled to the view with float ratios and (int) truncation."""
    if (W, H) == (dw, dh):
        return tuple(roi)
    sx, sy = np.float32(W) / np.float32(dw), np.float32(H) / np.float32(dh)
    return tuple(int(np.float32(v) * s) for v, s in zip(roi, (sx, sy, sx, sy)))


def _check(name, f, dl, view, dr, roi=None, max_abs=2):
    """All four north-star gates for one filter() call of filter object f against the oracle with f's parameters."""
    use_conf = f._use_confidence
    got = f.filter(dl, view, None, dr if use_conf else None, roi)
    ref = O.wls_filter(dl, view, dr if use_conf else None, roi=roi, use_confidence=use_conf, lam=f.getLambda(), sigma=f.getSigmaColor(),
                       lrc_thresh=f.getLRCthresh(), disc_radius=f.getDepthDiscontinuityRadius(), offsets=f._offsets, want_mask=True)
    assert ref["rc"] == 0
    assert tuple(f.getROI()) == tuple(ref["roi"])                                  # gate (i): ROI bit-exact
    H, W = view.shape[:2]
    dh, dw = dl.shape
    froi = _view_roi(ref["roi"], dw, dh, W, H)
    outside = np.ones((H, W), bool)
    outside[froi[1]:froi[1] + froi[3], froi[0]:froi[0] + froi[2]] = False
    assert (got[outside] == -16).all() and (ref["out"][outside] == -16).all()      # fill exact
    dconf, mask_equal = 0.0, None
    if use_conf:
        mask = f.lrc_mask(dl, dr, view_cols=W, ROI=roi)
        mask_equal = bool((mask == ref["mask"]).all())
        assert mask_equal                                                           # gate (ii): LRC decisions bit-exact
        conf = f.getConfidenceMap()
        assert conf.shape == (H, W)
        dconf = float(np.abs(conf - ref["conf"]).max())
        assert dconf <= 1e-4                                                        # gate (iii)
    st = assert_s16_parity(got, ref["out"], froi, max_abs=max_abs)                  # gate (iv)
    zero_conf = float((ref["conf"][froi[1]:froi[1] + froi[3], froi[0]:froi[0] + froi[2]] == 0).mean()) if use_conf else None
    record(name, roi=list(ref["roi"]), view_roi=list(froi), identical=st["exact"], within1=st["within1"], max_abs=st["max"], nan_corner_px=st["corner"],
           max_dconf=dconf, lrc_mask_equal=mask_equal, zero_conf_share=zero_conf)
    return got, ref, st


# ---------------------------------------------------------------------------------------------------------------
# configs[0]: 640x480, StereoBM, the sample's default (downscale) and --no-downscale paths, REAL cv2.StereoBM maps
# ---------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("downscale", [True, False])
def test_config0_stereobm_sample_path(ctx, downscale):
    pytest.importorskip("cv2")
    left, right, _ = synth.make_stereo_pair(640, 480, seed=42, max_disp=140)
    f, dl, dr = sample_matching(left, right, "bm", downscale=downscale, max_disp=160, ctx=ctx)
    f.setLambda(8000.0); f.setSigmaColor(1.5)                                       # sample defaults (31-32)
    assert dl.dtype == np.int16 and (dl == -16).any() or (dl < 0).any()             # a real matcher marks invalid pixels
    _check("cfg0_bm_%s" % ("downscale" if downscale else "fullres"), f, dl, left, dr)
    assert f.getROI() == ((83, 3, 234, 234) if downscale else (167, 7, 466, 466))   # SURVEY section 8 config 1 / 1b geometry
    assert f.getDepthDiscontinuityRadius() == (3 if downscale else 5)


# ---------------------------------------------------------------------------------------------------------------
# configs[1]: 1920x1080, REAL cv2.StereoSGBM(0, 160, 3, P1=216, P2=864, cap 63, 3WAY) left + right maps
# ---------------------------------------------------------------------------------------------------------------
def test_config1_stereosgbm_1080p(ctx):
    pytest.importorskip("cv2")
    left, right, _ = synth.make_stereo_pair(1920, 1080, seed=42)
    f, dl, dr = sample_matching(left, right, "sgbm", downscale=False, max_disp=160, ctx=ctx)
    f.setLambda(8000.0); f.setSigmaColor(1.5)
    _check("cfg1_sgbm_1080p", f, dl, left, dr)
    assert f.getROI() == (160, 0, 1760, 1080) and f.getDepthDiscontinuityRadius() == 2 and f.getLRCthresh() == 24


def test_config1_stereosgbm_1080p_batch_of_one_equals_single_call(ctx):
    """The device batch entry point on the same real-matcher frame gives the bytes of the host single call."""
    torch = pytest.importorskip("torch")
    pytest.importorskip("cv2")
    from addingdisparityfiltering_b200.batch import BatchFilter
    left, right, _ = synth.make_stereo_pair(1920, 1080, seed=43)
    f, dl, dr = sample_matching(left, right, "sgbm", downscale=False, max_disp=160, ctx=ctx)
    f.setLambda(8000.0); f.setSigmaColor(1.5)
    single = f.filter(dl, left, None, dr)
    dev = torch.device("cuda", 0)
    out = torch.empty((1, 1080, 1920), dtype=torch.int16, device=dev)
    conf = torch.empty((1, 1080, 1920), dtype=torch.float32, device=dev)
    bf = BatchFilter(True, 8000.0, 1.5, 24, 2, offsets=f._offsets, ctx=ctx)
    bf.filter_device(torch.from_numpy(dl)[None].to(dev), torch.from_numpy(left)[None].to(dev), out, torch.from_numpy(dr)[None].to(dev), conf)
    torch.cuda.synchronize()
    assert (out[0].cpu().numpy() == single).all() and (conf[0].cpu().numpy() == f.getConfidenceMap()).all()


# ---------------------------------------------------------------------------------------------------------------
# configs[2]: fastGlobalSmootherFilter standalone, 3840x2160, 3-channel guide, num_iter 3, attenuation 0.25
# ---------------------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def guide_4k():
    rng = np.random.default_rng(7)
    g = synth._gauss_blur(rng.integers(0, 256, (2160, 3840, 3)).astype(np.float32), 2.0)
    return np.clip(np.rint(128 + (g - 128) * 3.0), 0, 255).astype(np.uint8)


@pytest.mark.parametrize("lam,sigma", [(1000.0, 10.0), (8000.0, 1.5), (1000.0, 1.5), (8000.0, 10.0)])
def test_config2_fgs_4k_float_source(ctx, guide_4k, lam, sigma):
    rng = np.random.default_rng(int(lam + sigma))
    src = (rng.random((2160, 3840)) * 255).astype(np.float32)
    got = A.fastGlobalSmootherFilter(guide_4k, src, lam, sigma, 0.25, 3, ctx=ctx)
    ref = O.fgs_filter(guide_4k, src, lam, sigma, 0.25, 3)
    err = np.abs(got - ref)
    record("cfg2_fgs4k_32FC1_lam%g_sig%g" % (lam, sigma), max_abs_err=float(err.max()), mean_abs_err=float(err.mean()), value_range=255.0)
    assert err.max() <= 255 * 1e-3                                                  # float tolerance: 1e-3 of the value range


@pytest.mark.parametrize("lam,sigma", [(1000.0, 10.0), (8000.0, 1.5)])
def test_config2_fgs_4k_8uc3_source(ctx, guide_4k, lam, sigma):
    got = A.fastGlobalSmootherFilter(guide_4k, guide_4k, lam, sigma, 0.25, 3, ctx=ctx)
    ref = O.fgs_filter(guide_4k, guide_4k, lam, sigma, 0.25, 3)
    d = np.abs(got.astype(np.int64) - ref.astype(np.int64))
    record("cfg2_fgs4k_8UC3_lam%g_sig%g" % (lam, sigma), identical=float((d == 0).mean()), max_abs=int(d.max()), l1_per_px=float(d.mean()))
    assert d.max() <= 1 and d.sum() <= d.size / 64                                  # the reference's envelope (test_fgs_filter.cpp:114-115)


@pytest.mark.parametrize("dtype,cn", [(np.int16, 1), (np.float32, 3), (np.float32, 4), (np.uint8, 4), (np.int16, 3)])
def test_config2_fgs_4k_other_types(ctx, guide_4k, dtype, cn):
    """The rest of the type matrix at 4K (fgs_filter.cpp:184; CV_32FC3 is the one round 1 never ran)."""
    rng = np.random.default_rng(cn + 100)
    shape = (2160, 3840) if cn == 1 else (2160, 3840, cn)
    if dtype == np.uint8:
        src = rng.integers(0, 256, shape).astype(np.uint8)
    elif dtype == np.int16:
        src = rng.integers(-32767, 32767, shape).astype(np.int16)
    else:
        src = (rng.random(shape) * 255).astype(np.float32)
    got = A.fastGlobalSmootherFilter(guide_4k, src, 500.0, 1.5, ctx=ctx)            # the interpolator's parameters (sparse_match_interpolators.cpp:203)
    ref = O.fgs_filter(guide_4k, src, 500.0, 1.5)
    if dtype == np.float32:
        err = np.abs(got - ref)
        record("cfg2_fgs4k_32FC%d_lam500_sig1.5" % cn, max_abs_err=float(err.max()), value_range=255.0)
        assert err.max() <= 255 * 1e-3
    else:
        d = np.abs(got.astype(np.int64) - ref.astype(np.int64))
        record("cfg2_fgs4k_%s_c%d" % (np.dtype(dtype).name, cn), identical=float((d == 0).mean()), max_abs=int(d.max()))
        assert d.max() <= 1 and d.sum() <= d.size / 64


# ---------------------------------------------------------------------------------------------------------------
# configs[3]: half-resolution maps (960x540) filtered at 1920x1080, ROI (83,3,874,534), radius 3, LRC 24
# ---------------------------------------------------------------------------------------------------------------
def _half_res_1080p(seed):
    import cv2
    view, dl, dr, roi = synth.make_artificial_example(1920, 1080, 3, seed=seed)
    # perf_disparity_wls_filter.cpp:76-83: bilinear resize by 0.5, values halved
    dl2 = (cv2.resize(dl, None, fx=0.5, fy=0.5) // 2).astype(np.int16)
    dr2 = (cv2.resize(dr, None, fx=0.5, fy=0.5) // 2).astype(np.int16)
    return view, dl2, dr2


@pytest.mark.parametrize("use_conf", [True, False])
def test_config3_half_res_1080p(ctx, use_conf):
    pytest.importorskip("cv2")
    view, dl2, dr2 = _half_res_1080p(1000)
    f = A.createDisparityWLSFilterGeneric(use_conf, ctx=ctx)
    f.setLambda(8000.0); f.setSigmaColor(1.5); f.setDepthDiscontinuityRadius(3); f.setLRCthresh(24)
    _check("cfg3_half_1080p_%s" % ("conf" if use_conf else "noconf"), f, dl2, view, dr2, roi=(83, 3, 874, 534))


def test_config3_half_res_1080p_real_stereobm(ctx):
    """The sample's default path at 1080p: views halved, StereoBM(80, 7) on 960x540, filtered at full resolution."""
    pytest.importorskip("cv2")
    left, right, _ = synth.make_stereo_pair(1920, 1080, seed=44)
    f, dl, dr = sample_matching(left, right, "bm", downscale=True, max_disp=160, ctx=ctx)
    f.setLambda(8000.0); f.setSigmaColor(1.5)
    assert dl.shape == (540, 960)
    _check("cfg3_half_1080p_real_bm", f, dl, left, dr)
    assert f.getROI() == (83, 3, 874, 534) and f.getDepthDiscontinuityRadius() == 3


# ---------------------------------------------------------------------------------------------------------------
# the reference's tutorial pair through the sample's default path, against outputs of the COMPILED REFERENCE
# ---------------------------------------------------------------------------------------------------------------
def test_ambush_pair_against_reference_outputs(ctx):
    pytest.importorskip("cv2")
    z, left, right = ambush_fixture()
    if int(left.astype(np.int64).sum()) != int(z["left_sum"]):
        pytest.skip("this cv2 build decodes the JPEG differently from the one that made the fixture")
    dl, dr = z["dl"], z["dr"]
    f = A.DisparityWLSFilter(True, *[int(v) for v in z["offsets"]], ctx=ctx)
    f.setLambda(8000.0); f.setSigmaColor(1.5); f.setDepthDiscontinuityRadius(int(z["radius"]))
    got = f.filter(dl, left, None, dr)
    assert tuple(f.getROI()) == tuple(int(v) for v in z["roi"])
    froi = _view_roi(f.getROI(), dl.shape[1], dl.shape[0], left.shape[1], left.shape[0])
    dconf = float(np.abs(f.getConfidenceMap() - z["conf"]).max())
    assert dconf <= 1e-4
    st = assert_s16_parity(got, z["out"], froi)
    outside = np.ones(got.shape, bool)
    outside[froi[1]:froi[1] + froi[3], froi[0]:froi[0] + froi[2]] = False
    assert (got[outside] == z["out"][outside]).all()
    record("ambush_5_bm_vs_compiled_reference", identical=st["exact"], within1=st["within1"], max_abs=st["max"], nan_corner_px=st["corner"], max_dconf=dconf)
    # and the matcher stage reproduces the stored maps when this cv2 build matches like the one that made them
    f2, dl2, dr2 = sample_matching(left, right, "bm", downscale=True, max_disp=160, ctx=ctx)
    if (dl2 == dl).all() and (dr2 == dr).all():
        f2.setLambda(8000.0); f2.setSigmaColor(1.5)
        assert (f2.filter(dl2, left, None, dr2) == got).all()


# ---------------------------------------------------------------------------------------------------------------
# upscale ratios other than 2x (general lin_coeff branch of the resize, different ratios in x and y)
# ---------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("disp_size,view_size", [((214, 160), (321, 240)), ((107, 80), (321, 240)), ((160, 120), (640, 480)),
                                                 ((200, 75), (500, 300)), ((127, 61), (400, 100)), ((640, 360), (1920, 1080))])
@pytest.mark.parametrize("use_conf", [True, False])
def test_upscale_ratios(ctx, disp_size, view_size, use_conf):
    import cv2
    dw, dh = disp_size
    W, H = view_size
    view, dl, dr, roi = synth.make_artificial_example(W, H, 3, seed=W + dh)
    fx = dw / W
    dl2 = np.rint(cv2.resize(dl, (dw, dh), interpolation=cv2.INTER_LINEAR).astype(np.float32) * fx).astype(np.int16)
    dr2 = np.rint(cv2.resize(dr, (dw, dh), interpolation=cv2.INTER_LINEAR).astype(np.float32) * fx).astype(np.int16)
    roi2 = (int(roi[0] * fx) + 1, 1, int(roi[2] * fx) - 2, dh - 2)
    f = A.createDisparityWLSFilterGeneric(use_conf, ctx=ctx)
    f.setLambda(8000.0); f.setSigmaColor(1.5); f.setDepthDiscontinuityRadius(3)
    _check("upscale_%dx%d_to_%dx%d_%s" % (dw, dh, W, H, "conf" if use_conf else "noconf"), f, dl2, view, dr2, roi=roi2)
