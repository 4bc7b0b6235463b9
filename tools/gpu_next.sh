#!/bin/bash
# the job the next gpurun call executes: 2 GPUs - NCCL correctness test, gather variants at N = 2
python -m pytest tests/test_sharded.py -m gpu -x -q 2>&1 | tail -3
tools/scale_variants.sh 2 "32 8 0" "32 8 0 NCCL_P2P_USE_CUDA_MEMCPY=1" "32 0 0" "16 4 0" "32 8 4"
grep -v "^\*\|OMP_NUM\|^$" gpurun_out/b2_1.err | tail -5
