#!/bin/bash
# two GPUs: the multi-rank parity tests, then the config-5 bench at N = 1 and 2 on the same box
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name --format=csv > gpurun_out/n2_smi.txt
timeout 900 python -m pytest tests/test_gpu_multirank.py -m gpu -q > gpurun_out/n2_pytest_multirank.log 2>&1; echo "pytest exit $?" >> gpurun_out/n2_pytest_multirank.log
MVBA_BENCH_DEBUG=1 timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/n2_bench_n1.json 2> gpurun_out/n2_bench_n1.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/n2_bench_n2.json 2> gpurun_out/n2_bench_n2.err
tail -6 gpurun_out/n2_pytest_multirank.log; grep "rank" gpurun_out/n2_bench_n1.err | head -3
for f in n2_bench_n1 n2_bench_n2; do python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/$f.json') if l.startswith('{')][-1])
    print('$f', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), {k:round(v*1e3/d['steps'],3) for k,v in d['phases_s'].items()}, d['solve'])
except Exception as e: print('$f', e)
PY
done
