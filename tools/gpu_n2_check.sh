#!/bin/bash
# N = 2 with both landmark-sharding modes (solve summaries must equal the N = 1 run's) + the RPC fit timing on GPU 0.
for L in 0 1; do
  MVBA_SHARD_LOCALITY=$L timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2953$L bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_n2_loc$L.json 2> gpurun_out/bench_n2_loc$L.err
  echo "locality $L exit $?"; tail -1 gpurun_out/bench_n2_loc$L.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['n_gpus'], round(d['value'],2), d['phases_s'], round(d['e2e']['value'],2), d['solve'])"
done
timeout 300 python tools/rpc_bench.py 400 4 120 > gpurun_out/rpc_bench.json 2> gpurun_out/rpc_bench.err; echo "rpc exit $?"; cat gpurun_out/rpc_bench.json; tail -3 gpurun_out/rpc_bench.err
