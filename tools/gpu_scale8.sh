#!/bin/bash
# config 5 at N = 1, 2, 4, 8 on one 8-GPU box (strong scaling; per-phase times in each line)
mkdir -p gpurun_out
timeout 400 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r02f_scale_n1.json 2> gpurun_out/r02f_scale_n1.err
for n in 2 4 8; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r02f_scale_n$n.json 2> gpurun_out/r02f_scale_n$n.err
done
for n in 1 2 4 8; do python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/r02f_scale_n$n.json') if l.startswith('{')][-1])
    print($n, round(d['value'],1), 'e2e', round(d['e2e']['value'],1), {k:round(v*1e3/d['steps'],3) for k,v in d['phases_s'].items()}, d['solve']['final_cost'], d['solve']['pcg_iterations'])
except Exception as e: print($n, e)
PY
done
