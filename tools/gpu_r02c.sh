#!/bin/bash
# Round 2, visit C: FP64 peaks, the 27-accumulator pass A (camq/camfold) against the whole GPU suite, bench lines
set -x
mkdir -p gpurun_out
tools/bin/fp64_peak > gpurun_out/fp64_peak.json 2>&1
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_config5.json 2> gpurun_out/bench_config5.err; echo "bench exit $?" >> gpurun_out/bench_config5.err
MVBA_CAM_STAGED=1 timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_config5_staged.json 2> gpurun_out/bench_config5_staged.err
for n in 1 2; do
  timeout 600 python bench.py --config $n --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_config$n.json 2> gpurun_out/bench_config$n.err
done
cat gpurun_out/fp64_peak.json; tail -5 gpurun_out/pytest_gpu.log
for n in 5 5_staged 1 2; do cut -c1-1500 gpurun_out/bench_config$n.json; tail -2 gpurun_out/bench_config$n.err; done
