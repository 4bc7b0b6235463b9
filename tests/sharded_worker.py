"""Worker of the multi-rank tests (one process per rank; launched by tests/test_sharded.py or by torchrun).

  python tests/sharded_worker.py <backend: gloo|nccl> <port> <rank> <world> <mode: root|all> [n_total]

gloo (CPU): the chunked, pipelined gather of ShardedBatch with a stand-in filter that stamps every frame with its global
index - checks the placement of every chunk of every rank, uneven shards included.
nccl (one GPU per rank): the real filter on every rank's shard of a batch, gathered over NCCL; rank `dst` then filters the
WHOLE batch alone and the gathered result must equal it bit for bit (frames are independent,
disparity_filters.cpp:138-139)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import numpy as np
    import torch
    import torch.distributed as dist
    from addingdisparityfiltering_b200.batch import ShardedBatch
    if len(sys.argv) >= 6:
        backend, port, rank, world, mode = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4]), sys.argv[5]
        n_total = int(sys.argv[6]) if len(sys.argv) > 6 else 13
        init = dict(init_method="tcp://127.0.0.1:%s" % port, rank=rank, world_size=world)
    else:                                                         # under torchrun: env:// rendezvous
        backend, mode, n_total = sys.argv[1], sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 13
        rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
        init = {}
    if backend == "nccl":
        local = int(os.environ.get("LOCAL_RANK", rank))
        torch.cuda.set_device(local)
        dev = torch.device("cuda", local)
        dist.init_process_group("nccl", device_id=dev, **init)
    else:
        dev = torch.device("cpu")
        dist.init_process_group("gloo", **init)
    H, W = (48, 64) if backend == "gloo" else (240, 320)
    ctx = None
    if backend == "nccl":
        import addingdisparityfiltering_b200 as A
        ctx = A.Context(dev.index)
    sb = ShardedBatch(n_total, (H, W), chunk=3, tail_min=1, mode=mode, dst=0, device=dev, ctx=ctx)
    assert sb.spans[rank] == (sb.begin, sb.end) and sum(e - b for b, e in sb.spans) == n_total

    if backend == "gloo":
        def filt(f0, nf, out):
            for i in range(nf):
                out[i].fill_(sb.begin + f0 + i)
        for rep in range(2):                                     # twice: the second step reuses the buffers
            sb.full.fill_(-1)
            full = sb.run(filt)
            if sb.holds_all:
                got = [int(full[f, 0, 0]) for f in range(n_total)]
                assert got == list(range(n_total)) and bool((full[:, -1, -1] == full[:, 0, 0]).all()), got
            else:
                assert full.shape[0] == sb.end - sb.begin
    else:
        from addingdisparityfiltering_b200 import synth
        from addingdisparityfiltering_b200.batch import BatchFilter
        frames = [synth.make_textured_scene(W, H, 3, seed=50 + f, max_disp=40) for f in range(n_total)]   # the same batch on every rank
        roi = (40, 0, W - 40, H)
        tv, tl, tr = [torch.from_numpy(np.stack([fr[k] for fr in frames])).to(dev) for k in range(3)]
        bf = BatchFilter(True, 8000.0, 1.5, 24, 2, ctx=ctx)
        lv, ll, lr = tv[sb.begin:sb.end], tl[sb.begin:sb.end], tr[sb.begin:sb.end]

        def filt(f0, nf, out):
            bf.filter_device(ll[f0:f0 + nf], lv[f0:f0 + nf], out, lr[f0:f0 + nf], None, roi)
        for rep in range(3):
            sb.full.fill_(7)
            full = sb.run(filt)
            torch.cuda.synchronize()
        if sb.holds_all:
            alone = torch.empty((n_total, H, W), dtype=torch.int16, device=dev)
            bf.filter_device(tl, tv, alone, tr, None, roi)
            torch.cuda.synchronize()
            assert torch.equal(full, alone), "gathered batch differs from the single-GPU batch"
            assert ctx.kernel_launches() > 0
        if mode == "peer":
            assert (sb.bytes_sent > 0) == (rank != 0)
        sb.close()
    dist.barrier()
    dist.destroy_process_group()
    print("ok rank %d/%d %s %s" % (rank, world, backend, mode))


if __name__ == "__main__":
    main()
