#!/bin/bash
# final check of the round: parity tests, smoke, bench line, reference arm (no ncu)
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log; tail -2 gpurun_out/pytest_gpu.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log; tail -2 gpurun_out/smoke.log
timeout 300 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
tail -1 gpurun_out/bench.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],2), round(d['e2e']['value'],2), d['phases_s'], d['cpu_baseline'], d['roofline'])"
