"""dwls_fgs_dev_noalloc (no counterpart in the reference): the
fastGlobalSmootherFilter device entry point whose scratch lives in a static device array, for callers that cannot call
cudaMalloc.  Called through the C-ABI and compared with the existing entry point dwls_fgs_dev on the same inputs, on calls
issued one after another (the shared array keeps the previous call's contents).  Collected last on purpose (pytest -x)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

FGS_ARGS = None


def _lib_and_ctx():
    import addingdisparityfiltering_b200 as A
    from addingdisparityfiltering_b200 import _lib
    L = _lib.lib()
    sig = [C.c_void_p, C.POINTER(_lib.Image), C.POINTER(_lib.Image), C.POINTER(_lib.Image), C.c_double, C.c_double, C.c_double, C.c_int, C.c_void_p]
    L.dwls_fgs_dev.argtypes = sig
    L.dwls_fgs_dev_noalloc.argtypes = sig
    L.dwls_fgs_noalloc_scratch_bytes.restype = C.c_size_t
    L.dwls_fgs_noalloc_required_bytes.restype = C.c_size_t
    L.dwls_fgs_noalloc_required_bytes.argtypes = [C.c_int, C.c_int, C.c_int]
    return A, _lib, L, A.Context(0)


def _img(_lib, t):
    import torch
    depth = {torch.uint8: 0, torch.int16: 3, torch.float32: 5}[t.dtype]
    cn = 1 if t.dim() == 2 else t.shape[2]
    return _lib.Image(t.data_ptr(), t.stride(0) * t.element_size(), t.shape[0], t.shape[1], depth, cn)


def _case(rng, w, h, kind):
    g = rng.integers(0, 256, (h, w, 3)).astype(np.float32)
    g = ((g + np.roll(g, 1, 0) + np.roll(g, 1, 1)) / 3).astype(np.uint8)
    if kind == "f32":
        src = (rng.random((h, w)) * 255).astype(np.float32)
    elif kind == "u8c3":
        src = rng.integers(0, 256, (h, w, 3)).astype(np.uint8)
    else:
        src = rng.integers(-32767, 32767, (h, w)).astype(np.int16)
    return g, src


def test_noalloc_entry_point_matches_dwls_fgs_dev_on_consecutive_calls():
    import torch
    A, _lib, L, ctx = _lib_and_ctx()
    dev = torch.device("cuda", 0)
    stream = torch.cuda.current_stream(dev).cuda_stream or 1
    rng = np.random.default_rng(2024)
    cap = L.dwls_fgs_noalloc_scratch_bytes()
    assert cap == 16 << 20
    # call 1 fills most of the array (1280x720 CV_32FC1 = 11.8 of 16 MiB), call 2 is a small odd-sized 3-channel image whose
    # planes land on call 1's leftovers (row padding 203 -> 208 included), call 3 another type, call 4 repeats call 1
    cases = [(1280, 720, "f32", 8000.0, 1.5), (203, 97, "u8c3", 1000.0, 10.0), (640, 480, "s16", 500.0, 1.5), (1280, 720, "f32", 8000.0, 1.5)]
    prepared = []
    for w, h, kind, lam, sigma in cases:
        g, src = _case(rng, w, h, kind)
        assert L.dwls_fgs_noalloc_required_bytes(w, h, 1 if src.ndim == 2 else src.shape[2]) <= cap
        tg, ts = torch.from_numpy(g).to(dev), torch.from_numpy(src).to(dev)
        want, got = torch.empty_like(ts), torch.empty_like(ts)
        prepared.append((tg, ts, want, got, lam, sigma, kind))
    # the existing entry point first (it allocates its workspace; this also loads the kernels in a fresh process)
    for tg, ts, want, got, lam, sigma, kind in prepared:
        G, S_, D = _img(_lib, tg), _img(_lib, ts), _img(_lib, want)
        ctx.check(L.dwls_fgs_dev(ctx.handle, C.byref(G), C.byref(S_), C.byref(D), lam, sigma, 0.25, 3, C.c_void_p(stream)))
    # Free device memory must not shrink over the measured calls.  What may legitimately take device memory the FIRST time
    # it runs in a process is code, not data: CUDA loads kernels lazily (the driver's own memset kernel behind
    # cudaMemsetAsync, torch's fill kernel used by this test).  So everything is run once, unmeasured, on the SMALLEST case -
    # a hidden grow-only allocation would then still have to grow for the 1280x720 call inside the measured region.
    small = min(prepared, key=lambda c: c[1].numel())
    G, S_, D = _img(_lib, small[0]), _img(_lib, small[1]), _img(_lib, small[3])
    ctx.check(L.dwls_fgs_dev_noalloc(ctx.handle, C.byref(G), C.byref(S_), C.byref(D), small[4], small[5], 0.25, 3, C.c_void_p(stream)))
    for c in prepared:
        c[3].fill_(0)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info(dev)[0]
    launches0 = ctx.kernel_launches()
    for tg, ts, want, got, lam, sigma, kind in prepared:          # back to back on one stream, no host sync in between
        G, S_, D = _img(_lib, tg), _img(_lib, ts), _img(_lib, got)
        ctx.check(L.dwls_fgs_dev_noalloc(ctx.handle, C.byref(G), C.byref(S_), C.byref(D), lam, sigma, 0.25, 3, C.c_void_p(stream)))
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info(dev)[0]
    assert ctx.kernel_launches() > launches0
    assert free_after >= free_before, "device memory was allocated during dwls_fgs_dev_noalloc (%d -> %d free bytes)" % (free_before, free_after)
    for i, (tg, ts, want, got, lam, sigma, kind) in enumerate(prepared):
        a, b = got.cpu().numpy(), want.cpu().numpy()
        if kind == "f32":
            err = float(np.abs(a - b).max())
            assert err <= 255 * 1e-3, (i, err)                                  # the float tolerance of the FGS parity tests
        else:
            d = np.abs(a.astype(np.int64) - b.astype(np.int64))
            assert d.max() <= 1 and d.sum() <= d.size / 64, (i, int(d.max()))   # the reference's own integer envelope
        assert (a == b).all(), "call %d: same kernels on the same inputs should even agree bit for bit" % i
    ctx.close()


def test_noalloc_entry_point_rejects_what_does_not_fit():
    import torch
    A, _lib, L, ctx = _lib_and_ctx()
    dev = torch.device("cuda", 0)
    for w, h, cn in ((1920, 1080, 1), (1280, 720, 3), (64, 2500, 1)):
        g = torch.zeros((h, w, 3), dtype=torch.uint8, device=dev)
        s = torch.zeros((h, w) if cn == 1 else (h, w, cn), dtype=torch.float32, device=dev)
        d = torch.empty_like(s)
        G, S_, D = _img(_lib, g), _img(_lib, s), _img(_lib, d)
        st = L.dwls_fgs_dev_noalloc(ctx.handle, C.byref(G), C.byref(S_), C.byref(D), 100.0, 1.0, 0.25, 3, None)
        assert st == _lib.DWLS_UNSUPPORTED
        msg = L.dwls_last_error(ctx.handle).decode()
        assert "dwls_fgs_dev_noalloc" in msg and ("scratch" in msg), msg
    assert L.dwls_fgs_noalloc_required_bytes(1920, 1080, 1) > L.dwls_fgs_noalloc_scratch_bytes()
    # argument checks are the ones of dwls_fgs_dev (fgs_filter.cpp:143-144, 184-189)
    g = torch.zeros((32, 32, 3), dtype=torch.uint8, device=dev)
    s = torch.zeros((16, 16), dtype=torch.float32, device=dev)
    G, S_, D = _img(_lib, g), _img(_lib, s), _img(_lib, torch.empty_like(s))
    assert L.dwls_fgs_dev_noalloc(ctx.handle, C.byref(G), C.byref(S_), C.byref(D), 100.0, 1.0, 0.25, 3, None) == _lib.DWLS_BAD_SIZE
    ctx.close()
