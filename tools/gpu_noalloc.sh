#!/bin/bash
# The static-scratch RPC fit: its test alone in a fresh process, then after the whole GPU suite (pytest -x as the round's end runs it)
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_noalloc_alone.log
timeout 2400 python -m pytest tests -x -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_config5.json 2> gpurun_out/bench_config5.err; echo "bench exit $?" >> gpurun_out/bench_config5.err
tail -15 gpurun_out/pytest_noalloc_alone.log; tail -8 gpurun_out/pytest_gpu.log; tail -c 600 gpurun_out/bench_config5.json
