#!/bin/bash
# ncu source-level capture of the round-1 tile kernel on config 4 (one launch)
mkdir -p gpurun_out
timeout 600 ncu --set full --import-source on --clock-control none -k regex:tile_kernel -s 2 -c 1 -f -o gpurun_out/tile_cfg4 python bench.py --config 4 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_tile_cfg4.log 2>&1
echo "ncu exit $?"; ls -la gpurun_out/*.ncu-rep; tail -3 gpurun_out/ncu_tile_cfg4.log
