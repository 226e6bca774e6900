#!/bin/bash
# Round 2, visit E: ncu --set full of the staged tile kernel and the 27-accumulator pass A (source counters for the hot lines)
set -x
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'tile2_kernel|camq_kernel|camfold_kernel' -s 6 -c 3 -o gpurun_out/r02e_build_full -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_r02e.log 2>&1
tail -3 gpurun_out/ncu_r02e.log | cut -c1-300
ls -la gpurun_out/
