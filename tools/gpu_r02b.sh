#!/bin/bash
# Round 2, visit B: ncu launch lists of configs 2-4 (which kernels carry the step on the paths config 5 never enters)
set -x
mkdir -p gpurun_out
for n in 2 3 4; do
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches_config$n.csv python bench.py --config $n --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launch_config$n.log 2>&1
done
MVBA_TILE_DIV=6 timeout 300 python bench.py --config 4 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench4_b.json 2>&1
tail -2 gpurun_out/ncu_launch_config4.log
