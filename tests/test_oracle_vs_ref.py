"""CPU tests: the oracle against the reference's own code (oracle/_ref, compiled from the reference
sources where they lie; the prebuilt library travels with the repo).  Bit-exact, for every stripe
count, because the oracle emulates which rows/columns fall into the reference's SIMD blocks."""
import numpy as np
import pytest

from addingdisparityfiltering_b200 import synth
from oracle import oracle as O
from oracle import ref as R

pytestmark = pytest.mark.skipif(not R.available(), reason="oracle/_ref not built (needs /root/reference at build time)")


@pytest.mark.parametrize("threads", [1, 3, 8])
def test_fgs_bit_exact(threads):
    R.set_num_threads(threads)
    rng = np.random.default_rng(threads)
    for (w, h, cn) in [(127, 61, 1), (203, 77, 3), (64, 5, 3), (9, 40, 1)]:
        g = rng.integers(0, 256, (h, w, cn)).astype(np.uint8)
        g = g[:, :, 0] if cn == 1 else g
        for src in ((rng.random((h, w)) * 255).astype(np.float32), rng.integers(-32767, 32767, (h, w, 3)).astype(np.int16),
                    rng.integers(0, 256, (h, w, 4)).astype(np.uint8)):
            lam, sigma = float(rng.uniform(100, 10000)), float(rng.uniform(1, 100))
            assert (R.fgs_filter(g, src, lam, sigma) == O.fgs_filter(g, src, lam, sigma, assoc=1, nstripes=threads)).all()


@pytest.mark.parametrize("threads", [1, 4])
@pytest.mark.parametrize("use_conf", [True, False])
@pytest.mark.parametrize("half", [False, True])
def test_wls_bit_exact(threads, use_conf, half):
    R.set_num_threads(threads)
    view, dl, dr, roi = synth.make_artificial_example(320, 240, 3, seed=3 + threads)
    if half:
        dl, dr, roi = np.ascontiguousarray(dl[::2, ::2] // 2), np.ascontiguousarray(dr[::2, ::2] // 2), tuple(v // 2 for v in roi)
    a = R.wls_filter(dl, view, dr, roi, use_conf, 8000.0, 1.5, 24, 3)
    b = O.wls_filter(dl, view, dr, roi=roi, use_confidence=use_conf, lam=8000.0, sigma=1.5, disc_radius=3, assoc=1, nstripes=threads)
    assert a["rc"] == 0 and b["rc"] == 0 and a["roi"] == b["roi"]
    assert (a["out"] == b["out"]).all()
    if use_conf:
        assert (a["conf"] == b["conf"]).all()


def test_factory_geometry_from_reference():
    # createDisparityWLSFilter / createRightMatcher of the reference itself (disparity_filters.cpp:385-449)
    import addingdisparityfiltering_b200 as A
    from tests.test_host_logic import _FakeMatcher
    for kind, mn, nd, bs, dw, dh in (("bm", 0, 80, 7, 320, 240), ("bm", 0, 160, 15, 640, 480), ("sgbm", 0, 160, 3, 1920, 1080), ("sgbm", 16, 64, 5, 400, 300)):
        roi, radius, rmin = R.factory_geometry(kind, mn, nd, bs, dw, dh)
        f = A.createDisparityWLSFilter(_FakeMatcher(kind, mn, nd, bs))
        lo, ro, to, bo = f._offsets
        assert roi == (lo, to, dw - lo - ro, dh - to - bo) and radius == f.getDepthDiscontinuityRadius() and rmin == -(mn + nd) + 1
