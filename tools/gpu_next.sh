#!/bin/bash
# the job the next gpurun call executes (edited between calls; gpurun snapshots the tree when the call starts)
echo "== single call, small-launch grid shapes on / off"
python tools/single_kernels.py 1
DWLS_SMALL_GRIDS=0 python tools/single_kernels.py 1
echo "== bench, default flags"
python bench.py > gpurun_out/bench_r02_n1.json 2> gpurun_out/bench_r02_n1.err; tail -c 1500 gpurun_out/bench_r02_n1.json; tail -3 gpurun_out/bench_r02_n1.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_r02_ref.json 2>&1
echo "== ncu"
CMD="python bench.py --frames 32 --steps 2 --warmup 1 --no-cpu --no-e2e --no-extras"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_r02.csv $CMD > gpurun_out/ncu_list.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:conf_run|weights_k|solve_k|border_k" -s 10 -c 10 -f -o gpurun_out/full_r02 $CMD > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
ls -la gpurun_out/ | tail -8
