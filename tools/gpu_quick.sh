#!/bin/bash
# quick GPU check: parity tests, bench, optional ncu capture ($1 = kernel regex), phase profile (if the profiling lib was built)
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],2), d['phases_s'], round(d['roofline']['launch_ms'],3), round(d['e2e']['value'],2))"
if [ -n "$1" ]; then
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:"$1" -c ${2:-1} -o gpurun_out/quick -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_quick.log 2>&1
fi
if [ -f multiview_b200/libmvba_prof.so ]; then cp multiview_b200/libmvba_prof.so multiview_b200/libmvba.so; timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>&1 | grep "tile p" | sort | uniq | head -3; fi
