"""Golden vectors produced by the reference itself (tests/golden/make_golden.py: the reference's two
source files compiled unmodified, 4 threads).  CPU: the oracle reproduces them bit-exactly, which is
what pins the oracle.  GPU: the CUDA path stays within the north_star tolerances of them."""
import os

import numpy as np
import pytest

from oracle import oracle as O
from tests.golden import make_golden as G
from tests.parity import assert_s16_parity

HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _load(name):
    return np.load(os.path.join(HERE, name + ".npz"))


@pytest.mark.parametrize("i", range(len(G.FGS_CASES)))
def test_oracle_reproduces_reference_fgs_bit_exact(i):
    name, w, h, cn, kind, lam, sigma = G.FGS_CASES[i]
    g, src = G.fgs_inputs(w, h, cn, kind, 100 + i)
    out = O.fgs_filter(g, src, lam, sigma, assoc=1, nstripes=G.THREADS)
    assert (out == _load(name)["out"]).all()


@pytest.mark.parametrize("i", range(len(G.WLS_CASES)))
def test_oracle_reproduces_reference_wls_bit_exact(i):
    name, w, h, cn, conf, half, radius, seed = G.WLS_CASES[i]
    view, dl, dr, roi = G.wls_inputs(w, h, cn, half, seed)
    gold = _load(name)
    r = O.wls_filter(dl, view, dr, roi=roi, use_confidence=conf, lam=8000.0, sigma=1.5, disc_radius=radius, assoc=1, nstripes=G.THREADS)
    assert r["rc"] == 0 and tuple(gold["roi"]) == r["roi"]
    assert (r["out"] == gold["out"]).all()
    if conf:
        assert (r["conf"] == gold["conf"]).all()


@pytest.mark.gpu
@pytest.mark.parametrize("i", range(len(G.FGS_CASES)))
def test_gpu_fgs_against_reference_golden(ctx, i):
    import addingdisparityfiltering_b200 as A
    name, w, h, cn, kind, lam, sigma = G.FGS_CASES[i]
    g, src = G.fgs_inputs(w, h, cn, kind, 100 + i)
    out = A.fastGlobalSmootherFilter(g, src, lam, sigma, ctx=ctx)
    gold = _load(name)["out"]
    if kind == "f32":
        assert np.abs(out - gold).max() <= 255 * 1e-3
    else:
        d = np.abs(out.astype(np.int64) - gold.astype(np.int64))
        assert d.max() <= 1 and d.sum() <= d.size / 64           # the reference's own envelope (test_fgs_filter.cpp:114-115)


@pytest.mark.gpu
@pytest.mark.parametrize("i", range(len(G.WLS_CASES)))
def test_gpu_wls_against_reference_golden(ctx, i):
    import addingdisparityfiltering_b200 as A
    name, w, h, cn, conf, half, radius, seed = G.WLS_CASES[i]
    view, dl, dr, roi = G.wls_inputs(w, h, cn, half, seed)
    gold = _load(name)
    f = A.createDisparityWLSFilterGeneric(conf, ctx=ctx)
    f.setLambda(8000.0); f.setSigmaColor(1.5); f.setDepthDiscontinuityRadius(radius)
    out = f.filter(dl, view, None, dr, roi)
    assert tuple(f.getROI()) == tuple(gold["roi"])                                  # ROI bit-exact
    s = 2 if half else 1
    froi = tuple(int(v) * s for v in roi)
    outside = np.ones(out.shape, bool)
    outside[froi[1]:froi[1] + froi[3], froi[0]:froi[0] + froi[2]] = False
    assert (out[outside] == gold["out"][outside]).all()                             # fill exact
    assert_s16_parity(out, gold["out"], froi)                                       # <= 1 LSB on >= 99.9 %
    if conf:
        assert np.abs(f.getConfidenceMap() - gold["conf"]).max() <= 1e-4            # confidence within 1e-4


def test_oracle_reproduces_reference_on_the_tutorial_pair_bit_exact():
    """ambush_5 through the sample's default path (real cv2.StereoBM maps, half resolution, radius 3): the stored outputs
    are the compiled reference's (tests/golden/make_golden_real.py)."""
    pytest.importorskip("cv2")
    from tests.stereo_inputs import ambush_fixture
    z, left, _ = ambush_fixture()
    if int(left.astype(np.int64).sum()) != int(z["left_sum"]):
        pytest.skip("this cv2 build decodes the JPEG differently from the one that made the fixture")
    r = O.wls_filter(z["dl"], left, z["dr"], lam=8000.0, sigma=1.5, disc_radius=int(z["radius"]), offsets=tuple(int(v) for v in z["offsets"]),
                     assoc=1, nstripes=4)
    assert r["rc"] == 0 and r["roi"] == tuple(int(v) for v in z["roi"])
    assert (r["out"] == z["out"]).all() and (r["conf"] == z["conf"]).all()
    assert (z["dl"] < 0).any()                      # a real matcher's invalid-disparity markers reach the LRC gather
