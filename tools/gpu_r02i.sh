#!/bin/bash
# round 2, late visit: whole GPU suite (no -x: every failure is listed), smoke, one config-5 bench line
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/final_smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/final_smoke.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_config5.json 2> gpurun_out/bench_config5.err; echo "bench exit $?" >> gpurun_out/bench_config5.err
tail -25 gpurun_out/pytest_gpu.log; tail -2 gpurun_out/final_smoke.log; head -c 300 gpurun_out/bench_config5.json
# phase profile of the round-1 tile kernel on config 4 and 3 (developer build with clock64 instrumentation)
if [ -f multiview_b200/libmvba_prof.so ]; then
  cp multiview_b200/libmvba.so /tmp/libmvba_keep.so
  cp multiview_b200/libmvba_prof.so multiview_b200/libmvba.so
  for c in 4 3; do
    timeout 300 python bench.py --config $c --steps 3 --warmup 3 --no-cpu-baseline 2>&1 | grep "tile" | grep -v metric | sort | uniq -c | sort -rn | head -6 > gpurun_out/tile_prof_config$c.txt
  done
  cp /tmp/libmvba_keep.so multiview_b200/libmvba.so
  cat gpurun_out/tile_prof_config4.txt gpurun_out/tile_prof_config3.txt
fi
