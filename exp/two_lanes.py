"""Does running two half-batches on two streams (row pass of one beside the column pass of the other) beat one batch?"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench as Bn
import addingdisparityfiltering_b200 as A
from addingdisparityfiltering_b200.batch import BatchFilter
dev = torch.device("cuda", 0)
base = Bn.make_base_frames(2)
tb = [[torch.from_numpy(a).to(dev) for a in fr] for fr in base]
def mk(B):
    view = torch.stack([torch.roll(tb[f % 2][0], shifts=3 * f, dims=1) for f in range(B)]).contiguous()
    dl = torch.stack([torch.roll(tb[f % 2][1], shifts=3 * f, dims=1) for f in range(B)]).contiguous()
    dr = torch.stack([torch.roll(tb[f % 2][2], shifts=3 * f, dims=1) for f in range(B)]).contiguous()
    out = torch.empty((B, Bn.H, Bn.W), dtype=torch.int16, device=dev)
    conf = torch.empty((B, Bn.H, Bn.W), dtype=torch.float32, device=dev)
    return view, dl, dr, out, conf
for lanes, B in ((1, 32), (2, 16), (2, 32), (3, 16), (4, 8)):
    ctxs = [A.Context(0) for _ in range(lanes)]
    bfs = [BatchFilter(True, Bn.LAMBDA, Bn.SIGMA, Bn.LRC, Bn.RADIUS, ctx=c) for c in ctxs]
    data = [mk(B) for _ in range(lanes)]
    streams = [torch.cuda.Stream() for _ in range(lanes)]
    def once(stagger):
        for i in range(lanes):
            v, dl, dr, out, conf = data[i]
            bfs[i].filter_device(dl, v, out, dr, conf, Bn.ROI, stream=streams[i].cuda_stream)
    for _ in range(4): once(False)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for s in streams: s.wait_event(e0)
    for _ in range(20): once(False)
    for s in streams: torch.cuda.current_stream().wait_stream(s)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    print("lanes", lanes, "frames/lane", B, "us/frame", round(ms * 1e3 / (lanes * B), 2), "Gpix/s", round(lanes * B * Bn.W * Bn.H / ms / 1e6, 2), flush=True)
    del ctxs, bfs, data
