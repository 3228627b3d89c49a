#!/bin/bash
# A/B on C1 (cbox.obj 512x512, 64 spp, depth 8), one GPU
for rep in 1 2; do
  i=0
  for f in "$@"; do
    i=$((i+1))
    python bench.py --workload c1 --no-cpu-baseline --no-gpu-baseline --no-extra --steps 20 --warmup 5 $f 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('rep $rep variant $i [$f] value', round(d['value'],1), 'traced-only', round(d['culling']['value_traced_only'],1), 'ms', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value'],1))"
  done
done
