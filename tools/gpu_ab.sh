#!/bin/bash
# A/B of an environment switch ($1, e.g. MVBA_BACKSUB_KEEP_R) on the config-5 bench line, alternating, three repetitions
for rep in 1 2 3; do for val in 0 1; do
  env $1=$val python bench.py --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$1=$val', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), {k:round(v*1e3/d['steps'],3) for k,v in d['phases_s'].items()})"
done; done
