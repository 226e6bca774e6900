#!/bin/bash
# last visit of the round on one GPU: whole suite (-x, as the driver runs it), smoke, the profile set at the final commit
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -x -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/final_smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/final_smoke.log
bash tools/gpu_profile.sh r02f
tail -3 gpurun_out/pytest_gpu.log; tail -2 gpurun_out/final_smoke.log
