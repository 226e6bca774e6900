#!/bin/bash
# tile plan (threads, accumulators per warp, tracks per round, staging capacity, widest window) of configs 4 and 3 from the instrumented build
mkdir -p gpurun_out
cp multiview_b200/libmvba.so /tmp/libmvba_keep.so
cp multiview_b200/libmvba_prof.so multiview_b200/libmvba.so
for c in 4 3; do
  timeout 100 python bench.py --config $c --steps 1 --warmup 3 --no-cpu-baseline 2>&1 | grep "tile plan" | head -1 > gpurun_out/r02g_tile_plan_config$c.txt
done
cp /tmp/libmvba_keep.so multiview_b200/libmvba.so
cat gpurun_out/r02g_tile_plan_config4.txt gpurun_out/r02g_tile_plan_config3.txt
