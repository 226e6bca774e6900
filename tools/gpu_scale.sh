#!/bin/bash
# bench at N ranks of one box (driver launch line); $1 = N
N=$1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
echo "exit $?"; tail -1 gpurun_out/bench_n$N.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['n_gpus'], round(d['value'],2), d['phases_s'], round(d['e2e']['value'],2))"
tail -3 gpurun_out/bench_n$N.err
