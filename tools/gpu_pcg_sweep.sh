#!/bin/bash
# PCG launch geometry sweep (config 5 and config 2): CTAs x threads of the single-launch PCG
mkdir -p gpurun_out
for cfg in 5 2; do
for t in 256 512; do
for g in 296 148 96 64 32; do
  MVBA_PCG_GRID=$g MVBA_PCG_THREADS=$t timeout 300 python bench.py --config $cfg --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
d=json.loads(sys.stdin.readline()); print('config $cfg threads $t grid $g linear ms/it', round(d['phases_s']['linear']*1e3/d['steps'],4), 'pcg its', d['solve']['pcg_iterations'], 'value', round(d['value'],1))"
done; done; done | tee gpurun_out/pcg_sweep.txt
