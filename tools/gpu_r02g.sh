#!/bin/bash
# tests + bench (new back-substitution, pass A kept over rejected steps) + A/B against the per-track back-substitution + ncu of the small kernels
bash tools/gpu_iter.sh "MVBA_BACKSUB_TRACKS=1 python bench.py --steps 20 --warmup 5 --no-cpu-baseline; timeout 900 ncu --set full --clock-control none --import-source on -k regex:camq_kernel\|backsub2_kernel\|residual_kernel -s 6 -c 3 -o gpurun_out/r02e_ncu_small -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r02e_ncu_small.log 2>&1"
