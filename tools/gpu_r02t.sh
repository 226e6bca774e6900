#!/bin/bash
# final check of round 2: the whole GPU suite exactly as the driver runs it, then smoke
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r02g_pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02g_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r02g_smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/r02g_smoke.log
tail -4 gpurun_out/r02g_pytest_gpu.log; tail -2 gpurun_out/r02g_smoke.log
