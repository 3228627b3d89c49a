#!/bin/bash
# A/B on the headline workload (C2, one GPU): tools/c2_ab.sh "<bench flags A>" "<bench flags B>" ...
for rep in 1 2; do
  i=0
  for f in "$@"; do
    i=$((i+1))
    python bench.py --no-cpu-baseline --no-gpu-baseline --no-extra --steps 5 --warmup 3 $f 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('rep $rep variant $i [$f] value', round(d['value'],1), 'traced-only', round(d['culling']['value_traced_only'],1), 'ms', round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],1), 'frac', round(d['roofline']['frac'],3))"
  done
done
