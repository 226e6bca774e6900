#!/bin/bash
# Round 2, visit D: staged tile kernel (tile2) -- whole GPU suite (no -x), bench lines, phase profile
set -x
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
MVBA_CAM_STAGED=1 MVBA_TILE_UNSTAGED=1 timeout 600 python -m pytest tests/test_gpu_converged.py -m gpu -q > gpurun_out/pytest_converged_old.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_converged_old.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_config5.json 2> gpurun_out/bench_config5.err; echo "bench exit $?" >> gpurun_out/bench_config5.err
MVBA_TILE_UNSTAGED=1 timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_config5_unstaged.json 2> gpurun_out/bench_config5_unstaged.err
for n in 1 2; do
  timeout 600 python bench.py --config $n --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_config$n.json 2> gpurun_out/bench_config$n.err
done
cp multiview_b200/libmvba_prof.so multiview_b200/libmvba.so
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>&1 | grep "tile" | sort | uniq -c | sort -rn | head -6 > gpurun_out/tile_prof.txt
tail -15 gpurun_out/pytest_gpu.log; tail -5 gpurun_out/pytest_converged_old.log; cat gpurun_out/tile_prof.txt
for n in 5 5_unstaged 1 2; do cut -c1-1200 gpurun_out/bench_config$n.json; tail -2 gpurun_out/bench_config$n.err; done
