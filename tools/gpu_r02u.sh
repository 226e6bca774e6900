#!/bin/bash
# bench.py smoke after the l2 note change: config 1 and config 5, short
mkdir -p gpurun_out
timeout 120 python bench.py --config 1 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_l2_c1.json 2> gpurun_out/bench_l2_c1.err; echo "c1 exit $?"
timeout 200 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_l2_c5.json 2> gpurun_out/bench_l2_c5.err; echo "c5 exit $?"
python - <<'PY'
import json
for c in ('c1', 'c5'):
    j = json.load(open(f'gpurun_out/bench_l2_{c}.json'))
    print(c, round(j['value'], 1), j['config']['l2'])
PY
