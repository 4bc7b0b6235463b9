"""H2D/D2H pinned-copy bandwidth with all ranks copying at once, before and after binding each rank to its GPU's NUMA node."""
import os, time, glob
import torch, torch.distributed as dist

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
p = torch.cuda.get_device_properties(local)
bus = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
try:
    node = int(open("/sys/bus/pci/devices/%s/numa_node" % bus).read())
except Exception as e:
    node = "?" + repr(e)
nodes = sorted(glob.glob("/sys/devices/system/node/node[0-9]*"))
aff = sorted(os.sched_getaffinity(0))
print("rank", rank, "bus", bus, "numa", node, "nodes", len(nodes), "affinity", aff[:2], "..", aff[-1], len(aff), flush=True)
if rank == 0:
    for n in nodes:
        print(n, open(n + "/cpulist").read().strip(), flush=True)
    os.system("nvidia-smi topo -m 2>&1 | head -14")

def measure(tag):
    n, m = 464486400, 132710400
    h = torch.empty(n, dtype=torch.uint8).pin_memory(); h.fill_(1)
    d = torch.empty(n, dtype=torch.uint8, device="cuda")
    ho = torch.empty(m, dtype=torch.uint8).pin_memory(); ho.fill_(1)
    do = torch.empty(m, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    for mode in ("h2d", "both"):
        for _ in range(2):
            torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize(); t0 = time.perf_counter()
            for _ in range(5):
                with torch.cuda.stream(s1): d.copy_(h, non_blocking=True)
                if mode == "both":
                    with torch.cuda.stream(s2): ho.copy_(do, non_blocking=True)
            torch.cuda.synchronize(); dt = time.perf_counter() - t0
        t = torch.tensor([5 * n / dt / 1e9], device="cuda"); g = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(g, t)
        if rank == 0:
            print(tag, mode, "H2D GB/s per rank", [round(float(x), 1) for x in g], "sum", round(sum(float(x) for x in g), 1), flush=True)

measure("unbound")
if isinstance(node, int) and node >= 0:
    cpus = set()
    for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
        a, _, b = part.partition("-")
        cpus |= set(range(int(a), int(b or a) + 1))
    cpus &= os.sched_getaffinity(0)
    if cpus:
        os.sched_setaffinity(0, cpus)
measure("numa-bound")
dist.destroy_process_group()
