#!/bin/bash
# gpu parity tests, then bench with a few values of the tile-count knob
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log; tail -2 gpurun_out/pytest_gpu.log
for D in 6 4 9 14; do
  MVBA_TILE_DIV=$D timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/sweep_$D.json 2> gpurun_out/sweep_$D.err
  tail -1 gpurun_out/sweep_$D.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('div', $D, round(d['value'],2), round(d['e2e']['value'],2), d['phases_s'])"
done
