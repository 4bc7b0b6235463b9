"""Multi-rank path: frames sharded over ranks, results gathered chunk by chunk.
CPU: world_size 2 and 3 on gloo (placement logic).  GPU: world_size 2 on NCCL when the box has two GPUs
(`gpurun --gpus 2 -- python -m pytest tests/test_sharded.py -m gpu`), gathered batch == single-GPU batch bit for bit."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORKER = os.path.join(ROOT, "tests", "sharded_worker.py")


def _run(backend, world, mode, n_total, timeout=300):
    port = str(21000 + (os.getpid() * 7 + world * 131 + (17 if mode == "all" else 0) + n_total) % 20000)
    procs = [subprocess.Popen([sys.executable, WORKER, backend, port, str(r), str(world), mode, str(n_total)], stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(world)]
    outs = []
    for p in procs:
        try:
            outs.append(p.communicate(timeout=timeout)[0])
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
    assert all(p.returncode == 0 for p in procs), "\n".join(outs)


@pytest.mark.parametrize("world,mode,n_total", [(2, "root", 13), (2, "all", 13), (3, "root", 8), (2, "root", 1)])
def test_sharded_gather_placement_on_gloo(world, mode, n_total):
    _run("gloo", world, mode, n_total)


def test_chunk_plan_covers_shard():
    from addingdisparityfiltering_b200.batch import chunk_plan
    for n in (0, 1, 5, 16, 17, 64, 170, 512):
        for chunk, tail in ((16, 4), (32, 0), (3, 1), (8, 8)):
            plan = chunk_plan(n, chunk, tail)
            assert sum(c for _, c in plan) == n and all(c > 0 for _, c in plan)
            assert all(plan[i][0] + plan[i][1] == plan[i + 1][0] for i in range(len(plan) - 1))
            assert not plan or (plan[0][0] == 0 and max(c for _, c in plan) <= chunk)


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["root", "all", "peer"])
def test_sharded_filter_two_gpus_nccl(mode):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    _run("nccl", 2, mode, 13, timeout=600)
