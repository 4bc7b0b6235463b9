#!/bin/bash
N=${1:-8}
tools/scale_variants.sh $N "32 4 0 peer" "32 8 0 peer" "32 4 -8 peer"
grep -v "^\*\|OMP_NUM\|^$" gpurun_out/b${N}_0.err | tail -3
