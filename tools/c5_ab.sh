#!/bin/bash
# C5 (two-level, 305 x 327 680 triangles) on one GPU at 32 spp: variants of the instanced kernel.  usage: tools/c5_ab.sh "<flags A>" "<flags B>" ...
for rep in 1 2; do
  i=0
  for f in "$@"; do
    i=$((i+1))
    python bench.py --workload c5 --spp 32 --no-cpu-baseline --no-gpu-baseline --no-extra --steps 2 --warmup 1 $f 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('rep $rep variant $i [$f] value', round(d['value'],1), 'ms', round(d['ms_per_step'],2), 'build_ms', round(d['scene']['build_ms'],1))"
  done
done
