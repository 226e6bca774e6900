#!/bin/bash
# config 4: 256-thread / 24-accumulator instantiation of the tile kernel against the 512-thread one; parity tests that reach it
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_configs.py tests/test_gpu_parity_eval.py tests/test_gpu_parity_solve.py tests/test_gpu_edges.py -m gpu -q > gpurun_out/pytest_direct.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_direct.log
for w in 0 1; do
  MVBA_TILE_256=$w timeout 300 python bench.py --config 4 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_config4_t$w.json 2> gpurun_out/bench_config4_t$w.err
done
timeout 300 python bench.py --config 3 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_config3.json 2> gpurun_out/bench_config3.err
cp multiview_b200/libmvba.so /tmp/libmvba_keep.so
cp multiview_b200/libmvba_prof.so multiview_b200/libmvba.so
timeout 300 python bench.py --config 4 --steps 3 --warmup 3 --no-cpu-baseline 2>&1 | grep "tile" | grep -v metric | sort | uniq -c | sort -rn | head -4 > gpurun_out/tile_prof_config4_t1.txt
cp /tmp/libmvba_keep.so multiview_b200/libmvba.so
tail -5 gpurun_out/pytest_direct.log
python - <<'PY'
import json
for w in ('4_t0', '4_t1', '3'):
    try:
        j = json.load(open(f'gpurun_out/bench_config{w}.json'))
        print('config', w, round(j['value'], 1), 'it/s', {k: round(v * 1e3 / j['steps'], 3) for k, v in j['phases_s'].items()}, j['solve'])
    except Exception as e:
        print('config', w, 'failed', e)
PY
cat gpurun_out/tile_prof_config4_t1.txt
