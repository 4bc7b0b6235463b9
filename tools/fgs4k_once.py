#!/usr/bin/env python
"""fastGlobalSmootherFilter on one 3840x2160 CV_32FC1 image with a 3-channel guide, device buffers (configs[2]); a few calls."""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import addingdisparityfiltering_b200 as A
from addingdisparityfiltering_b200 import _lib
dev = torch.device("cuda", 0); ctx = A.Context(0); L = _lib.lib()
L.dwls_fgs_dev.argtypes = [C.c_void_p, C.POINTER(_lib.Image), C.POINTER(_lib.Image), C.POINTER(_lib.Image), C.c_double, C.c_double, C.c_double, C.c_int, C.c_void_p]
W, H = 3840, 2160
g = torch.randint(0, 256, (H, W, 3), dtype=torch.uint8, device=dev)
src = torch.rand((H, W), device=dev) * 255; dst = torch.empty_like(src)
def img(t, depth, cn): return _lib.Image(t.data_ptr(), t.stride(0) * t.element_size(), t.shape[0], t.shape[1], depth, cn)
G, S_, D = img(g, 0, 3), img(src, 5, 1), img(dst, 5, 1)
st = torch.cuda.current_stream().cuda_stream or 1
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
    assert L.dwls_fgs_dev(ctx.handle, C.byref(G), C.byref(S_), C.byref(D), 1000.0, 10.0, 0.25, 3, C.c_void_p(st)) == 0
torch.cuda.synchronize(); print("ok", float(dst.mean()))
