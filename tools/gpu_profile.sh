#!/bin/bash
# Profile set of a milestone ($1 = tag, e.g. r02c): bench line, ncu launch list (serialised, cold cache: shares only), ncu --set full of the
# build kernels + PCG + back-substitution, all at config 5 full size on one GPU.
TAG=${1:-r02x}
set -x
mkdir -p gpurun_out
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/${TAG}_bench_config5.json 2> gpurun_out/${TAG}_bench_config5.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/${TAG}_launches_config5.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'tile2_kernel|camq_kernel|camfold_kernel|backsub_kernel|residual_kernel|pcg_fused_kernel' -s 12 -c 6 -o gpurun_out/${TAG}_ncu_full -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_ncu_full.log 2>&1
tail -2 gpurun_out/${TAG}_ncu_full.log | cut -c1-200
cut -c1-400 gpurun_out/${TAG}_bench_config5.json
