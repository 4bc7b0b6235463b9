#!/bin/bash
# usage: tools/scale_variants.sh N "chunk tail root_extra mode [ENV=val ...]" ...   -> gpurun_out/bN_<i>.json, one summary line each
N=$1; shift
i=0
for cfg in "$@"; do
  set -- $cfg; chunk=$1; tail=$2; extra=$3; mode=$4; shift 4
  env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600+i)) bench.py --gpus $N --steps 5 --warmup 3 --no-e2e --no-cpu --chunk $chunk --tail-min $tail --root-extra $extra --gather-mode $mode > gpurun_out/b${N}_$i.json 2> gpurun_out/b${N}_$i.err
  python - "gpurun_out/b${N}_$i.json" "$cfg" <<'PY'
import json,sys
ok=False
for l in open(sys.argv[1]):
    if l.startswith('{'):
        d=json.loads(l); ok=True
        print(sys.argv[2], "| value", round(d["value"]), "compute_only", round(d["compute_only"]["value"]), "nvlink", d["nvlink_bytes_per_step"], d["run"]["chunk_plan_rank0"], "pipe", round(d['roofline']['pipeline']['frac'],3))
if not ok: print(sys.argv[2], "FAILED")
PY
  i=$((i+1))
done
