#!/bin/bash
# 8 GPUs, short: the gather root owns fewer frames than the other ranks (it also receives everybody's results)
N=${1:-8}
tools/scale_variants.sh $N "16 4 -16" "16 4 -32" "16 4 -48" "32 8 -32" "16 4 -32 NCCL_MAX_P2P_NCHANNELS=4"
grep -v "^\*\|OMP_NUM\|^$" gpurun_out/b${N}_4.err | tail -3
