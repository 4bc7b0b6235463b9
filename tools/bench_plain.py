"""cfg 2 batch throughput with the per-kernel event brackets off (what a caller of filter_device sees)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench as Bn
import addingdisparityfiltering_b200 as A
from addingdisparityfiltering_b200.batch import BatchFilter

dev = torch.device("cuda", 0)
ctx = A.Context(0)
B = int(sys.argv[1]) if len(sys.argv) > 1 else 32
base = Bn.make_base_frames(2)
tb = [[torch.from_numpy(a).to(dev) for a in fr] for fr in base]
view = torch.stack([torch.roll(tb[f % 2][0], shifts=3 * f, dims=1) for f in range(B)]).contiguous()
dl = torch.stack([torch.roll(tb[f % 2][1], shifts=3 * f, dims=1) for f in range(B)]).contiguous()
dr = torch.stack([torch.roll(tb[f % 2][2], shifts=3 * f, dims=1) for f in range(B)]).contiguous()
out = torch.empty((B, Bn.H, Bn.W), dtype=torch.int16, device=dev)
conf = torch.empty((B, Bn.H, Bn.W), dtype=torch.float32, device=dev)
bf = BatchFilter(True, Bn.LAMBDA, Bn.SIGMA, Bn.LRC, Bn.RADIUS, ctx=ctx)
for graphs in (False, True):
    ctx.set_cuda_graphs(graphs)
    for _ in range(4):
        bf.filter_device(dl, view, out, dr, conf, Bn.ROI)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        bf.filter_device(dl, view, out, dr, conf, Bn.ROI)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    print("frames", B, "graphs", graphs, "overlap", os.environ.get("DWLS_NO_OVERLAP") is None, "ms/step", round(ms, 4),
          "us/frame", round(ms * 1e3 / B, 2), "Gpix/s", round(B * Bn.W * Bn.H / ms / 1e6, 2))
