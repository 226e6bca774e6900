#!/bin/bash
# ncu --set full of the staged tile kernel and pass A at the current commit
set -x
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'tile2_kernel|camq_kernel' -s 4 -c 2 -o gpurun_out/r02f_build_full -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_r02f.log 2>&1
tail -3 gpurun_out/ncu_r02f.log | cut -c1-300
