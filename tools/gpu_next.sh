#!/bin/bash
N=${1:-8}
tools/scale_variants.sh $N "16 4 0 peer" "32 0 0 peer" "16 4 0 root"
grep -v "^\*\|OMP_NUM\|^$" gpurun_out/b${N}_0.err | tail -4
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29777 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_r02_n$N.json 2> gpurun_out/bench_r02_n$N.err
python - "gpurun_out/bench_r02_n$N.json" <<'PY'
import json,sys
for l in open(sys.argv[1]):
    if l.startswith('{'):
        d=json.loads(l); print("DEFAULT", d["n_gpus"], "value", round(d["value"]), "compute_only", round(d["compute_only"]["value"]), "e2e", round(d["e2e"]["value"]), "mode", d["run"]["gather_mode"], d["run"]["gather_fallback"], d["run"]["chunk_plan_rank0"])
PY
grep -v "^\*\|OMP_NUM\|^$" gpurun_out/bench_r02_n$N.err | tail -3
