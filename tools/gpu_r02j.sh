#!/bin/bash
# config 4 (RPC distortion floating): direct mode of the tile kernel against the staged one; parity tests that reach that path
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_configs.py tests/test_gpu_parity_eval.py tests/test_gpu_parity_solve.py tests/test_gpu_edges.py tests/test_rpc_update.py tests/test_rig_pass.py -m gpu -q > gpurun_out/pytest_direct.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_direct.log
for d in 0 1; do
  MVBA_TILE_DIRECT=$d timeout 300 python bench.py --config 4 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_config4_direct$d.json 2> gpurun_out/bench_config4_direct$d.err
done
cp multiview_b200/libmvba.so /tmp/libmvba_keep.so
cp multiview_b200/libmvba_prof.so multiview_b200/libmvba.so
for d in 0 1; do
  MVBA_TILE_DIRECT=$d timeout 300 python bench.py --config 4 --steps 3 --warmup 3 --no-cpu-baseline 2>&1 | grep "tile" | grep -v metric | sort | uniq -c | sort -rn | head -4 > gpurun_out/tile_prof_config4_direct$d.txt
done
cp /tmp/libmvba_keep.so multiview_b200/libmvba.so
tail -5 gpurun_out/pytest_direct.log
python - <<'PY'
import json
for d in (0, 1):
    try:
        j = json.load(open(f'gpurun_out/bench_config4_direct{d}.json'))
        print('direct', d, round(j['value'], 1), 'it/s', {k: round(v * 1e3 / j['steps'], 3) for k, v in j['phases_s'].items()}, j['solve'])
    except Exception as e:
        print('direct', d, 'failed', e)
PY
cat gpurun_out/tile_prof_config4_direct0.txt gpurun_out/tile_prof_config4_direct1.txt
