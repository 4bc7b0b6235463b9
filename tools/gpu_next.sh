#!/bin/bash
N=${1:-8}
tools/scale_variants.sh $N "16 4 4 peer" "16 4 8 peer" "16 2 4 peer"
grep -v "^\*\|OMP_NUM\|^$" gpurun_out/b${N}_0.err | tail -3
