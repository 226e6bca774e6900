#!/bin/bash
# N = 8 with image-ordered landmark sharding (MVBA_SHARD_LOCALITY=1)
MVBA_SHARD_LOCALITY=1 timeout 170 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/bench_n8_loc1.json 2> gpurun_out/bench_n8_loc1.err
echo "exit $?"; tail -1 gpurun_out/bench_n8_loc1.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['n_gpus'], round(d['value'],2), d['phases_s'], round(d['e2e']['value'],2), d['solve'])"
tail -3 gpurun_out/bench_n8_loc1.err
