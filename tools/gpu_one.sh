#!/bin/bash
python tools/fgs4k_once.py 3 > gpurun_out/plain_fgs.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:hsolve_kernel -s 3 -c 1 -f -o gpurun_out/h4k_r02 python tools/fgs4k_once.py 3 > gpurun_out/ncu_h4k.log 2>&1
tail -2 gpurun_out/ncu_h4k.log
