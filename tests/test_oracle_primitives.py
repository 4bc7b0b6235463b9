"""CPU tests (no GPU): the oracle's restatement of the OpenCV primitives the reference calls
(boxFilter, sqrBoxFilter, resize, convertTo — external to the reference tree) against cv2 4.x on
this machine, plus the data-free known-answer properties of the reference's own tests."""
import numpy as np
import pytest

from oracle import oracle as O

cv2 = pytest.importorskip("cv2")


@pytest.fixture(autouse=True)
def _no_ipp():
    # OpenCV's own code path (IPP substitutes a differently-rounding resize); restored afterwards
    old = cv2.ipp.useIPP()
    cv2.ipp.setUseIPP(False)
    yield
    cv2.ipp.setUseIPP(old)


@pytest.mark.parametrize("w,h,r", [(203, 57, 2), (216, 40, 5), (64, 33, 3), (1760, 48, 2), (234, 60, 3), (17, 9, 5)])
def test_box_filters_match_cv2(w, h, r):
    rng = np.random.default_rng(w + h + r)
    a = rng.integers(-3000, 3000, (h, w)).astype(np.int16)
    k = 2 * r + 1
    m, q = O.box_mean_16s(a, r), O.box_meansq_16s(a, r)
    m2 = cv2.boxFilter(a, cv2.CV_32F, (k, k))
    q2 = cv2.sqrBoxFilter(a, cv2.CV_32F, (k, k))
    assert (q == q2).all()                                    # double sum * double scale -> float: bit exact
    main = 4 * (w // 4)                                       # cv2's SIMD body: float scale (the oracle's canonical form)
    assert (m[:, :main] == m2[:, :main]).all()
    tail = np.abs(m[:, main:] - m2[:, main:])                 # scalar tail of cv2 uses a double scale: <= 1 ulp
    assert (tail <= np.spacing(np.abs(m2[:, main:]))).all()


def test_box_filter_extreme_values_exact():
    a = np.full((40, 50), -32768, np.int16)
    a[::3, ::2] = 32767
    q = O.box_meansq_16s(a, 5)
    assert (q == cv2.sqrBoxFilter(a, cv2.CV_32F, (11, 11))).all()


@pytest.mark.parametrize("src,dst", [((61, 45), (122, 90)), ((63, 30), (127, 61)), ((160, 120), (320, 240)), ((50, 40), (50, 40))])
def test_resize_matches_cv2_bit_exact(src, dst):
    rng = np.random.default_rng(src[0] * dst[1])
    a = rng.integers(-3000, 3000, (src[1], src[0])).astype(np.int16)
    assert (O.resize_linear(a, dst) == cv2.resize(a, dst, interpolation=cv2.INTER_LINEAR)).all()
    f = (rng.random((src[1], src[0])) * 255).astype(np.float32)
    assert (O.resize_linear(f, dst) == cv2.resize(f, dst, interpolation=cv2.INTER_LINEAR)).all()


def test_lut_matches_formula():
    for sigma in (0.5, 1.0, 1.5, 10.0):
        lut = O.fgs_lut(sigma, 4096)
        i = np.arange(4096, dtype=np.float32)
        ref = -np.exp(-np.sqrt(i) / np.float32(sigma)).astype(np.float32)
        assert np.abs(lut - ref).max() <= np.spacing(np.float32(1.0))


def test_constant_image_invariance():
    # FastGlobalSmootherTest.SplatSurfaceAccuracy (modules/ximgproc/test/test_fgs_filter.cpp:59-87)
    rng = np.random.default_rng(0)
    for cn in (1, 3):
        g = rng.integers(0, 256, (97, 131, cn)).astype(np.uint8)
        g = g[:, :, 0] if cn == 1 else g
        val = int(rng.integers(-32000, 32000))
        out = O.fgs_filter(g, np.full((97, 131), val, np.int16), 1000.0, 10.0)
        assert np.abs(out.astype(np.int64) - val).mean() <= 1.0 / 64


def test_association_envelope():
    # FastGlobalSmootherTest.MultiThreadReproducibility (109-151): the reference's 1-thread and N-thread
    # runs associate the denominator differently; both forms of the oracle stay inside L-inf <= 1, L1 <= total/64
    rng = np.random.default_rng(5)
    g = rng.integers(0, 256, (120, 160, 3)).astype(np.uint8)
    src = rng.integers(-32767, 32767, (120, 160)).astype(np.int16)
    a = O.fgs_filter(g, src, 5000.0, 20.0, assoc=0, nstripes=1)
    b = O.fgs_filter(g, src, 5000.0, 20.0, assoc=1, nstripes=8, threaded=True)
    d = np.abs(a.astype(np.int64) - b.astype(np.int64))
    assert d.max() <= 1 and d.sum() <= d.size / 64


def test_wls_roi_rules_and_fill():
    from addingdisparityfiltering_b200 import synth
    view, dl, dr, roi = synth.make_artificial_example(160, 120, 3, seed=2)
    r = O.wls_filter(dl, view, dr, roi=None, offsets=(83, 0, 3, 3), lam=8000.0, sigma=1.5, disc_radius=3, want_mask=True)
    assert r["rc"] == 0 and r["roi"] == (83, 3, 77, 114)             # disparity_filters.cpp:231-233
    x, y, w, h = r["roi"]
    outside = np.ones(dl.shape, bool)
    outside[y:y + h, x:x + w] = False
    assert (r["out"][outside] == -16).all() and (r["conf"][outside] == 0).all() and (r["mask"][outside] == 0).all()
    assert set(np.unique(r["mask"][~outside])) <= {1, 2, 3}
    assert O.wls_filter(dl, view, dr, roi=(100, 0, 100, 120))["rc"] == 2     # ROI outside the map: cv::Mat(m, roi) asserts
    # half-resolution call: ROI reported at disparity resolution, confidence map at view size (139, 274)
    r2 = O.wls_filter(dl[::2, ::2].copy(), view, dr[::2, ::2].copy(), roi=(24, 0, 56, 60), lam=8000.0, sigma=1.5, disc_radius=3)
    assert r2["rc"] == 0 and r2["roi"] == (24, 0, 56, 60) and r2["conf"].shape == (120, 160)
    assert (r2["out"][:, :48] == -16).all()


def test_all_zero_confidence_gives_int16_min():
    # c~ == 0 -> 0 * (1/1e-43f) = 0 * inf = NaN -> cvRound -> INT_MIN -> saturate: -32768 (disparity_filters.cpp:295-296)
    view = np.zeros((32, 48, 3), np.uint8)
    dl = np.full((32, 48), 160, np.int16)
    dr = np.full((32, 48), 3000, np.int16)                    # every LRC check fails -> confidence 0 wherever it is evaluated
    r = O.wls_filter(dl, view, dr, roi=(10, 0, 38, 32), lam=100.0, sigma=1.0, disc_radius=1, want_mask=True)
    assert r["rc"] == 0
    if (r["mask"][:, 10:] == 3).all():
        assert (r["out"][:, 10:] == -32768).all()
