#!/bin/bash
# A/B on C3: tools/c3_ab.sh "<bench flags A>" "<bench flags B>" ... ; three alternating repetitions each
for rep in 1 2 3; do
  i=0
  for f in "$@"; do
    i=$((i+1))
    python bench.py --workload c3 --no-cpu-baseline --no-gpu-baseline --no-extra --steps 4 --warmup 2 $f 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('rep $rep variant $i [$f] value', round(d['value'],1), 'ms', round(d['ms_per_step'],2))"
  done
done
