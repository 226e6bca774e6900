#!/bin/bash
# config 4: 12-accumulator instantiation, window caps; parity tests that reach the round-1 tile kernel
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_configs.py tests/test_gpu_parity_eval.py tests/test_gpu_parity_solve.py tests/test_gpu_edges.py -m gpu -q > gpurun_out/pytest_direct.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_direct.log
for w in 0 11 9; do
  if [ $w -eq 0 ]; then unset MVBA_TILE_WCAP; else export MVBA_TILE_WCAP=$w; fi
  timeout 300 python bench.py --config 4 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_config4_w$w.json 2> gpurun_out/bench_config4_w$w.err
done
unset MVBA_TILE_WCAP
cp multiview_b200/libmvba.so /tmp/libmvba_keep.so
cp multiview_b200/libmvba_prof.so multiview_b200/libmvba.so
for w in 0 11; do
  if [ $w -eq 0 ]; then unset MVBA_TILE_WCAP; else export MVBA_TILE_WCAP=$w; fi
  timeout 300 python bench.py --config 4 --steps 3 --warmup 3 --no-cpu-baseline 2>&1 | grep "tile" | grep -v metric | sort | uniq -c | sort -rn | head -3 > gpurun_out/tile_prof_config4_w$w.txt
done
unset MVBA_TILE_WCAP
cp /tmp/libmvba_keep.so multiview_b200/libmvba.so
tail -5 gpurun_out/pytest_direct.log
python - <<'PY'
import json
for w in (0, 11, 9):
    try:
        j = json.load(open(f'gpurun_out/bench_config4_w{w}.json'))
        print('wcap', w, round(j['value'], 1), 'it/s', {k: round(v * 1e3 / j['steps'], 3) for k, v in j['phases_s'].items()}, j['solve'])
    except Exception as e:
        print('wcap', w, 'failed', e)
PY
cat gpurun_out/tile_prof_config4_w0.txt gpurun_out/tile_prof_config4_w11.txt
