#!/bin/bash
for L in 1 0; do
MVBA_BENCH_DEBUG=1 MVBA_SHARD_LOCALITY=$L timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2955$L bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/dbg_n2_$L.json 2> gpurun_out/dbg_n2_$L.err
grep "rank" gpurun_out/dbg_n2_$L.err; tail -1 gpurun_out/dbg_n2_$L.json | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['e2e']['value'])"
done
