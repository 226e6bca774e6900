#!/bin/bash
# ncu source-level capture of camera_pass_kernel on config 4 (one launch)
mkdir -p gpurun_out
timeout 600 ncu --set full --import-source on --clock-control none -k regex:camera_pass_kernel -s 1 -c 1 -f -o gpurun_out/r02g_campass_cfg4 python bench.py --config 4 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r02g_ncu_campass_cfg4.log 2>&1
echo "ncu exit $?"; ls -la gpurun_out/r02g_campass_cfg4.ncu-rep
