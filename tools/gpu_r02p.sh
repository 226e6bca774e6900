#!/bin/bash
# two GPUs, final code of round 2: the multi-rank parity tests and the config-5 bench line at N = 2
mkdir -p gpurun_out
nvidia-smi --query-gpu=name --format=csv > gpurun_out/r02g_n2_smi.txt
timeout 600 python -m pytest tests/test_gpu_multirank.py -m gpu -q > gpurun_out/r02g_pytest_multirank_2gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02g_pytest_multirank_2gpu.log
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r02g_scale_n2.json 2> gpurun_out/r02g_scale_n2.err
tail -4 gpurun_out/r02g_pytest_multirank_2gpu.log; cut -c1-300 gpurun_out/r02g_scale_n2.json
