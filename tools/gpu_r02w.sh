#!/bin/bash
# examples/fit_rpc on the GPU
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_fit_rpc_tool.py -m gpu -q > gpurun_out/r02g_pytest_fit_rpc_tool.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02g_pytest_fit_rpc_tool.log
tail -15 gpurun_out/r02g_pytest_fit_rpc_tool.log
