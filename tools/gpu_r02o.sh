#!/bin/bash
# round 2, last visit on one GPU (tag r02g): whole suite as the driver runs it, smoke, bench lines of all five configs, launch list +
# phase profile + ncu --set full of the tile kernel on config 4
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -m gpu -q > gpurun_out/r02g_pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02g_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r02g_smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/r02g_smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r02g_bench_config5.json 2> gpurun_out/r02g_bench_config5.err
timeout 600 python bench.py --config 4 --steps 10 --warmup 3 > gpurun_out/r02g_bench_config4.json 2> gpurun_out/r02g_bench_config4.err
for c in 1 2 3; do
  timeout 300 python bench.py --config $c --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r02g_bench_config$c.json 2> gpurun_out/r02g_bench_config$c.err
done
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r02g_launches_config4.csv python bench.py --config 4 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r02g_ncu_launch4.log 2>&1
timeout 300 ncu --set full --import-source on --clock-control none -k regex:tile_kernel -s 2 -c 1 -f -o gpurun_out/r02g_tile_cfg4 python bench.py --config 4 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r02g_ncu_tile_cfg4.log 2>&1
cp multiview_b200/libmvba.so /tmp/libmvba_keep.so
cp multiview_b200/libmvba_prof.so multiview_b200/libmvba.so
timeout 300 python bench.py --config 4 --steps 3 --warmup 3 --no-cpu-baseline 2>&1 | grep "tile" | grep -v metric | sort | uniq -c | sort -rn | head -4 > gpurun_out/r02g_tile_prof_config4.txt
cp /tmp/libmvba_keep.so multiview_b200/libmvba.so
tail -4 gpurun_out/r02g_pytest_gpu.log; tail -2 gpurun_out/r02g_smoke.log
python - <<'PY'
import json
for c in (5, 4, 3, 2, 1):
    try:
        j = json.load(open(f'gpurun_out/r02g_bench_config{c}.json'))
        print('config', c, round(j['value'], 1), 'it/s e2e', round(j['e2e']['value'], 1), {k: round(v * 1e3 / j['steps'], 3) for k, v in j['phases_s'].items()})
    except Exception as e:
        print('config', c, 'failed', e)
PY
cat gpurun_out/r02g_tile_prof_config4.txt
