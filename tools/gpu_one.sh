#!/bin/bash
# single-GPU job: matcher profile + e2e upload A/B
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python tools/bm_bench.py 32 5
python tools/e2e_bench.py 64 5
DWLS_E2E_WINDOWS=0 python tools/e2e_bench.py 64 5
python tools/e2e_bench.py 64 5 4
python tools/e2e_bench.py 64 5 16
python tools/bm_bench.py 4 1 > gpurun_out/plain_bm.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:bm_kernel -s 2 -c 1 -f -o gpurun_out/bm_r02 python tools/bm_bench.py 4 1 > gpurun_out/ncu_bm.log 2>&1
tail -2 gpurun_out/ncu_bm.log
