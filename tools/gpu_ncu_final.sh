#!/bin/bash
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/launches_final.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launch_final.log 2>&1
timeout 200 ncu --set full --clock-control none --import-source on -k regex:'backsub_kernel' -c 1 -o gpurun_out/backsub_full -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_backsub.log 2>&1
tail -2 gpurun_out/ncu_launch_final.log | cut -c1-300
