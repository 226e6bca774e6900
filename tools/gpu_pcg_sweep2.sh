#!/bin/bash
# PCG launch geometry of config 5 after the per-CTA border sums
for t in 512 256; do for g in 296 148 96 74 64; do
  MVBA_PCG_GRID=$g MVBA_PCG_THREADS=$t timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
d=json.loads(sys.stdin.readline()); print('threads $t grid $g linear ms/it', round(d['phases_s']['linear']*1e3/d['steps'],4), 'pcg its', d['solve']['pcg_iterations'], 'value', round(d['value'],1))"
done; done
