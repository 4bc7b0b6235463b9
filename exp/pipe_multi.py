"""Mimic of the host-batch pipeline (2 lanes: H2D -> busy kernel -> D2H) in plain torch, all ranks at once."""
import os, time, sys
import torch, torch.distributed as dist
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
FR_IN, FR_OUT, NF = 14515200, 4147200, 32
h = torch.empty(NF * FR_IN, dtype=torch.uint8).pin_memory(); h.fill_(1)
ho = torch.empty(NF * FR_OUT, dtype=torch.uint8).pin_memory(); ho.fill_(1)
work = torch.empty(64 << 20, dtype=torch.float32, device="cuda")

def run(G, lanes, d2h_stream, kern_us_per_frame=58):
    st = [torch.cuda.Stream() for _ in range(lanes)]
    sd = torch.cuda.Stream()
    d = [torch.empty(G * FR_IN, dtype=torch.uint8, device="cuda") for _ in range(lanes)]
    do = [torch.empty(G * FR_OUT, dtype=torch.uint8, device="cuda") for _ in range(lanes)]
    evs = [None] * lanes
    def once():
        ln = 0
        for f0 in range(0, NF, G):
            s = st[ln]
            with torch.cuda.stream(s):
                if d2h_stream and evs[ln] is not None: s.wait_event(evs[ln])
                d[ln].copy_(h[f0 * FR_IN:(f0 + G) * FR_IN], non_blocking=True)
                torch.cuda._sleep(int(G * kern_us_per_frame * 1965))
                if d2h_stream:
                    e = torch.cuda.Event(); e.record(s)
                    with torch.cuda.stream(sd):
                        sd.wait_event(e)
                        ho[f0 * FR_OUT:(f0 + G) * FR_OUT].copy_(do[ln], non_blocking=True)
                        evs[ln] = torch.cuda.Event(); evs[ln].record(sd)
                else:
                    ho[f0 * FR_OUT:(f0 + G) * FR_OUT].copy_(do[ln], non_blocking=True)
            ln = (ln + 1) % lanes
        torch.cuda.synchronize()
    once()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(5): once()
    dt = (time.perf_counter() - t0) / 5
    t = torch.tensor([dt * 1e3], device="cuda"); g = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(g, t)
    if rank == 0:
        ms = max(float(x) for x in g)
        print("G", G, "lanes", lanes, "d2h_stream", d2h_stream, "ms/step per rank", [round(float(x), 2) for x in g],
              "=> Gpix/s total", round(world * NF * 2.0736 / ms, 2), flush=True)

for G in (4, 8, 16):
    for lanes in (2, 3):
        for ds in (False, True):
            run(G, lanes, ds)
dist.destroy_process_group()
