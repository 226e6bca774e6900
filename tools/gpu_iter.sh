#!/bin/bash
# Round 2 iteration visit: whole GPU suite, config-5 bench line, tile phase profile (developer build), optional extra command ($1)
set -x
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_config5.json 2> gpurun_out/bench_config5.err; echo "bench exit $?" >> gpurun_out/bench_config5.err
if [ -n "$1" ]; then bash -c "$1" > gpurun_out/extra.log 2>&1; fi
cp multiview_b200/libmvba.so /tmp/libmvba_keep.so
cp multiview_b200/libmvba_prof.so multiview_b200/libmvba.so
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>&1 | grep "tile" | grep -v metric | sort | uniq -c | sort -rn | head -9 > gpurun_out/tile_prof.txt
cp /tmp/libmvba_keep.so multiview_b200/libmvba.so
tail -8 gpurun_out/pytest_gpu.log; cat gpurun_out/tile_prof.txt
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_config5.json'))
print(round(d['value'],1), 'it/s; ms/step', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value'],1), {k:round(v*1e3/d['steps'],3) for k,v in d['phases_s'].items()}, 'build launch_ms', round(d['roofline']['launch_ms'],3), d['solve'])
PY
