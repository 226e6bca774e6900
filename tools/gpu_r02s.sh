#!/bin/bash
# the rig_calibrator command lines on the GPU (tests/test_rig_pass.py incl. --no_rig)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_rig_pass.py -m gpu -q > gpurun_out/r02g_pytest_rig_pass.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02g_pytest_rig_pass.log
tail -30 gpurun_out/r02g_pytest_rig_pass.log
