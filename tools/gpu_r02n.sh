#!/bin/bash
# config 4: phase c loop variants of the wide tile kernel (MVBA_TILE_CMODE 0 / 1); parity tests under CMODE=1
mkdir -p gpurun_out
MVBA_TILE_CMODE=1 timeout 900 python -m pytest tests/test_gpu_configs.py tests/test_gpu_parity_solve.py tests/test_gpu_edges.py -m gpu -q > gpurun_out/pytest_cmode1.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_cmode1.log
for w in 0 1; do
  MVBA_TILE_CMODE=$w timeout 300 python bench.py --config 4 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_config4_c$w.json 2> gpurun_out/bench_config4_c$w.err
done
cp multiview_b200/libmvba.so /tmp/libmvba_keep.so
cp multiview_b200/libmvba_prof.so multiview_b200/libmvba.so
for w in 0 1; do
  MVBA_TILE_CMODE=$w timeout 300 python bench.py --config 4 --steps 3 --warmup 3 --no-cpu-baseline 2>&1 | grep "tile" | grep -v metric | sort | uniq -c | sort -rn | head -2 > gpurun_out/tile_prof_config4_c$w.txt
done
cp /tmp/libmvba_keep.so multiview_b200/libmvba.so
tail -4 gpurun_out/pytest_cmode1.log
python - <<'PY'
import json
for w in ('4_c0', '4_c1'):
    try:
        j = json.load(open(f'gpurun_out/bench_config{w}.json'))
        print('config', w, round(j['value'], 1), 'it/s', {k: round(v * 1e3 / j['steps'], 3) for k, v in j['phases_s'].items()}, j['solve'])
    except Exception as e:
        print('config', w, 'failed', e)
PY
cat gpurun_out/tile_prof_config4_c0.txt gpurun_out/tile_prof_config4_c1.txt
