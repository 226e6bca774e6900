#!/bin/bash
# end-of-round check on one GPU: smoke, default bench line (with cpu_baseline), reference arm, then the profile set
set -x
mkdir -p gpurun_out
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/final_smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/final_smoke.log
( time timeout 900 python bench.py > gpurun_out/final_bench_default.json 2> gpurun_out/final_bench_default.err ) 2> gpurun_out/final_bench_default.time
( time timeout 900 python bench.py --impl reference > gpurun_out/final_bench_reference.json 2> gpurun_out/final_bench_reference.err ) 2> gpurun_out/final_bench_reference.time
bash tools/gpu_profile.sh r02f
tail -3 gpurun_out/final_smoke.log; cat gpurun_out/final_bench_default.time gpurun_out/final_bench_reference.time; cut -c1-300 gpurun_out/final_bench_reference.json
