#!/bin/bash
# final code: whole GPU suite, then one bench line per BASELINE config 1-4 (config 5 is the default line)
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -x -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
for n in 1 2 3 4; do
  timeout 600 python bench.py --config $n --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02f_bench_config$n.json 2> gpurun_out/r02f_bench_config$n.err
  python - <<PY
import json
try:
    d=json.load(open('gpurun_out/r02f_bench_config$n.json'))
    print($n, d['config']['workload'], round(d['value'],1), 'it/s e2e', round(d['e2e']['value'],1), {k:round(v*1e3/d['steps'],3) for k,v in d['phases_s'].items()}, 'frac', round(d['roofline']['frac'],3), d['roofline']['kernel'][:60])
except Exception as e: print($n, e)
PY
done
tail -4 gpurun_out/pytest_gpu.log
