"""CPU tests (no GPU): the C-ABI library loads and exports every declared symbol, the host-side
mirror validates arguments like the reference, the matcher factories reproduce the reference's
geometry, the metric helpers match their definitions, and the multi-rank sharding / gather path runs
on gloo with world_size 2."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from addingdisparityfiltering_b200 import _lib
    _lib.build()
    header = open(os.path.join(ROOT, "include", "dwls_b200.h")).read()
    declared = set(re.findall(r"\b(dwls_[a-z0-9_]+)\s*\(", header))
    assert len(declared) >= 16
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), name
    assert declared == set(_lib.EXPORTS)
    lib.dwls_version.restype = ctypes.c_char_p
    assert b"sm_100a" in lib.dwls_version()


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    import addingdisparityfiltering_b200 as A
    with pytest.raises(A.DwlsError) as e:
        A.Context(0)
    assert e.value.status == 3 and "no CPU fallback" in str(e.value)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "addingdisparityfiltering_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                txt = open(os.path.join(dirpath, f), errors="ignore").read().lower()
                assert "oracle" not in txt, os.path.join(dirpath, f)


class _FakeMatcher:
    def __init__(self, kind, min_disp, num_disp, bs):
        self.kind, self.v, self.calls = kind, (min_disp, num_disp, bs), {}

    def getMinDisparity(self): return self.v[0]
    def getNumDisparities(self): return self.v[1]
    def getBlockSize(self): return self.v[2]

    def __getattr__(self, n):
        if n.startswith("set"):
            return lambda v: self.calls.__setitem__(n, v)
        raise AttributeError(n)


def test_factory_geometry_matches_reference_rules():
    import addingdisparityfiltering_b200 as A
    # StereoBM (disparity_filters.cpp:397-403): margins (min+num)+w/2, -min+w/2, w/2, w/2 ; radius ceil(0.33 w)
    f = A.createDisparityWLSFilter(_FakeMatcher("bm", 0, 160, 15))
    assert f._offsets == (167, 7, 7, 7) and f.getDepthDiscontinuityRadius() == 5
    # StereoSGBM (404-408): margins (min+num), -min, 0, 0 ; radius ceil(0.5 w)
    m = _FakeMatcher("sgbm", 0, 160, 3)
    f = A.createDisparityWLSFilter(m)
    assert f._offsets == (160, 0, 0, 0) and f.getDepthDiscontinuityRadius() == 2
    assert m.calls == {"setDisp12MaxDiff": 1000000, "setSpeckleWindowSize": 0, "setUniquenessRatio": 0}
    assert (f.getLambda(), f.getSigmaColor(), f.getLRCthresh()) == (8000.0, 1.0, 24)
    with pytest.raises(A.DwlsError):
        A.createDisparityWLSFilter(_FakeMatcher("other", 0, 16, 3))


def test_right_matcher_with_real_cv2_matchers():
    cv2 = pytest.importorskip("cv2")
    import addingdisparityfiltering_b200 as A
    bm = cv2.StereoBM_create(80, 7)
    r = A.createRightMatcher(bm)
    assert (r.getMinDisparity(), r.getNumDisparities(), r.getBlockSize()) == (-79, 80, 7)          # 424-425
    sg = cv2.StereoSGBM_create(0, 160, 3, P1=216, P2=864, preFilterCap=63, mode=cv2.STEREO_SGBM_MODE_SGBM_3WAY)
    r = A.createRightMatcher(sg)
    assert (r.getMinDisparity(), r.getNumDisparities(), r.getP1(), r.getP2(), r.getMode()) == (-159, 160, 216, 864, sg.getMode())
    f = A.createDisparityWLSFilter(sg)
    assert f._offsets == (160, 0, 0, 0) and sg.getUniquenessRatio() == 0 and sg.getDisp12MaxDiff() == 1000000


def test_argument_validation_before_any_device_call():
    import addingdisparityfiltering_b200 as A
    f = A.createDisparityWLSFilterGeneric(True)
    with pytest.raises(A.DwlsError):
        f.filter(np.zeros((4, 4), np.float32), np.zeros((4, 4, 3), np.uint8))          # CV_Assert depth == CV_16S (221)
    with pytest.raises(A.DwlsError):
        f.filter(np.zeros((4, 4), np.int16), np.zeros((4, 4, 2), np.uint8))            # 1 or 3 channels (222)
    with pytest.raises(A.DwlsError):
        A.createFastGlobalSmootherFilter(np.zeros((4, 4), np.float32), 1.0, 1.0)        # guide CV_8U (fgs_filter.cpp:144)
    with pytest.raises(A.DwlsError):
        A.createFastGlobalSmootherFilter(np.zeros((4, 4), np.uint8), -1.0, 1.0)         # lambda >= 0 (143)


def test_metric_helpers():
    import addingdisparityfiltering_b200 as A
    rng = np.random.default_rng(0)
    gt = rng.integers(0, 2000, (40, 60)).astype(np.int16)
    src = (gt + rng.integers(-40, 40, gt.shape)).astype(np.int16)
    gt[5:10, 5:10] = 16320                                                              # UNKNOWN_DISPARITY (460)
    roi = (4, 3, 50, 30)
    g, s = gt[3:33, 4:54].astype(np.int64), src[3:33, 4:54].astype(np.int64)
    k = g != 16320
    assert A.computeMSE(gt, src, roi) == pytest.approx(((g - s)[k] ** 2).sum() / (k.sum() * 256))      # 497-517
    assert A.computeBadPixelPercent(gt, src, roi) == pytest.approx(100.0 * (np.abs(g - s)[k] >= 24).sum() / k.sum())
    vis = A.getDisparityVis(gt, 2.0)
    assert vis.dtype == np.uint8 and (vis[5:10, 5:10] == 0).all() and vis[0, 0] == min(255, int(np.rint(2.0 * gt[0, 0] / 16.0)))


def test_shard_frames_covers_batch():
    from addingdisparityfiltering_b200.batch import shard_frames
    for n, world in ((512, 8), (512, 3), (5, 8), (1, 1)):
        spans = [shard_frames(n, world, r) for r in range(world)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
        assert max(e - b for b, e in spans) - min(e - b for b, e in spans) <= 1


def test_weighted_shards_cover_batch():
    from addingdisparityfiltering_b200.batch import shard_frames, shard_frames_weighted
    for n, world, extra in ((512, 8, 0), (512, 8, 8), (512, 8, -24), (512, 8, -64), (512, 8, -100), (512, 2, 5), (7, 4, 100), (512, 1, 9), (5, 3, -1)):
        spans = [shard_frames_weighted(n, world, r, 0, extra) for r in range(world)]
        assert spans[0][0] == 0 and spans[-1][1] == n and all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
        sizes = [e - b for b, e in spans]
        assert all(v >= 0 for v in sizes)
        if extra == 0 or world == 1:
            assert spans == [shard_frames(n, world, r) for r in range(world)]
        else:
            assert sizes[0] == max(0, min(n, n // world + extra)) and max(sizes[1:]) - min(sizes[1:]) <= 1


_WORKER = r'''
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from addingdisparityfiltering_b200.batch import shard_frames, gather_results
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%s" % sys.argv[2], rank=int(sys.argv[3]), world_size=2)
rank = dist.get_rank()
b, e = shard_frames(6, 2, rank)
local = torch.stack([torch.full((4, 5), f, dtype=torch.int16) for f in range(b, e)])
full = gather_results(local)
assert full.shape == (6, 4, 5) and [int(full[f, 0, 0]) for f in range(6)] == list(range(6)), full[:, 0, 0]
dist.barrier(); dist.destroy_process_group(); print("ok", rank)
'''


def test_two_rank_gather_on_gloo(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(_WORKER)
    port = str(29500 + os.getpid() % 2000)
    procs = [subprocess.Popen([sys.executable, str(script), ROOT, port, str(r)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
             for r in range(2)]
    outs = [p.communicate(timeout=180)[0] for p in procs]
    assert all(p.returncode == 0 for p in procs), outs


def test_batch_arguments_are_validated_before_reaching_the_c_abi():
    """The batch path hands raw pointers and strides to the kernels: wrong dtype / shape / layout / placement must raise
    BAD_ARG on the host (like image_of() does for single frames) instead of writing out of bounds on the device."""
    import torch
    from addingdisparityfiltering_b200 import DwlsError
    from addingdisparityfiltering_b200.batch import _make_batch
    n, h, w = 3, 8, 16
    dl, v = np.zeros((n, h, w), np.int16), np.zeros((n, h, w, 3), np.uint8)
    o, c = np.zeros((n, h, w), np.int16), np.zeros((n, h, w), np.float32)
    b = _make_batch(dl, dl, v, o, c)
    assert (b.n_frames, b.disp_cols, b.disp_rows, b.view_channels, b.dl_step, b.dl_frame, b.conf_step) == (n, w, h, 3, 2 * w, 2 * w * h, 4 * w)
    big = np.zeros((n, h + 2, w + 4), np.int16)
    b = _make_batch(big[:, 1:h + 1, 2:w + 2], None, torch.from_numpy(v)[..., 0].contiguous(), torch.from_numpy(o), None)   # row/frame-strided views are fine
    assert (b.dl_step, b.dl_frame, b.view_channels) == (2 * (w + 4), 2 * (w + 4) * (h + 2), 1)
    bad = [lambda: _make_batch(dl.astype(np.float32), dl, v, o, c),                 # CV_16S required (221)
           lambda: _make_batch(dl, dl[:, :, :-1], v, o, c),                          # right map of another size (263)
           lambda: _make_batch(dl, dl, v, o[:, :-1], c),                             # undersized output
           lambda: _make_batch(dl, dl, v, o, c.astype(np.float64)),
           lambda: _make_batch(dl[:, :, ::2], None, v[:, :, ::2], o[:, :, ::2], None),   # column-sliced view
           lambda: _make_batch(np.transpose(dl, (0, 2, 1)), None, np.zeros((n, w, h), np.uint8), np.zeros((n, w, h), np.int16), None),
           lambda: _make_batch(dl, dl, v, o, c, on_device=True),                     # host arrays handed to the device entry point
           lambda: _make_batch(dl[:0], None, v[:0], o[:0], None),                    # empty batch
           lambda: _make_batch(dl, None, v[..., 0], o, None),                        # channel-sliced guide
           lambda: _make_batch(dl[::-1], None, v, o, None)]                          # negative frame stride
    for f in bad:
        with pytest.raises(DwlsError) as e:
            f()
        assert e.value.status == 1


def test_bench_reference_arm_prints_exactly_one_json_line():
    # bench.py contract: ONE JSON line on stdout (libraries such as NCCL write banners there, so the bench parks file
    # descriptor 1 on stderr while it runs); the reference arm runs here on the CPU with a tiny sample
    import json
    import subprocess
    import sys
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0", "--ref-frames", "1"],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.split("\n") if l.strip()]
    assert len(lines) == 1, r.stdout[:2000]
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "Mpix/s" and d["value"] > 0
    for key in ("metric", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "config", "cpu_baseline", "e2e"):
        assert key in d, key
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["e2e"]["h2d_bytes_per_step"] == 0
    # a non-zero rank of a torchrun launch of this arm does no work and prints nothing
    r1 = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"],
                        capture_output=True, text=True, timeout=600, cwd=ROOT, env=dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1"))
    assert r1.returncode == 0 and r1.stdout.strip() == ""
