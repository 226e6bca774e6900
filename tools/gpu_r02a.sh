#!/bin/bash
# Round 2, visit A: the new config 1-4 parity tests, the whole GPU suite, a bench line per config.
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/smi.txt 2>&1
nproc >> gpurun_out/smi.txt; free -g >> gpurun_out/smi.txt
timeout 900 python -m pytest tests/test_gpu_configs.py -m gpu -q -x > gpurun_out/pytest_configs.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_configs.log
timeout 900 python -m pytest tests -m gpu -q --deselect tests/test_gpu_configs.py > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
for n in 1 2 3 4; do
  timeout 600 python bench.py --config $n --steps 10 --warmup 3 > gpurun_out/bench_config$n.json 2> gpurun_out/bench_config$n.err; echo "bench $n exit $?" >> gpurun_out/bench_config$n.err
done
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_config5.json 2> gpurun_out/bench_config5.err; echo "bench exit $?" >> gpurun_out/bench_config5.err
tail -5 gpurun_out/pytest_configs.log; tail -5 gpurun_out/pytest_gpu.log
for n in 1 2 3 4 5; do cut -c1-600 gpurun_out/bench_config$n.json; tail -2 gpurun_out/bench_config$n.err; done
