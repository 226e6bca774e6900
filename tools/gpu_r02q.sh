#!/bin/bash
# config 4 (and 3): tile planning knobs of the round-1 tile kernel (tracks a tile must hold before a wider track may close it)
mkdir -p gpurun_out
for m in 128 64 32 16 8; do
  MVBA_TILE_MIN=$m timeout 200 python bench.py --config 4 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_config4_min$m.json 2> gpurun_out/bench_config4_min$m.err
done
for m in 128 32 8; do
  MVBA_TILE_MIN=$m timeout 200 python bench.py --config 3 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_config3_min$m.json 2> gpurun_out/bench_config3_min$m.err
done
python - <<'PY'
import json
for c, ms in ((4, (128, 64, 32, 16, 8)), (3, (128, 32, 8))):
    for m in ms:
        try:
            j = json.load(open(f'gpurun_out/bench_config{c}_min{m}.json'))
            print('config', c, 'tile_min', m, round(j['value'], 1), 'it/s', {k: round(v * 1e3 / j['steps'], 3) for k, v in j['phases_s'].items()}, j['solve']['final_cost'])
        except Exception as e:
            print('config', c, m, 'failed', e)
PY
