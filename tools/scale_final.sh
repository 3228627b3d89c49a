#!/bin/bash
# the driver's scaling run (N = 2, 4, 8; default bench line with its extras) on the final round-2 code
out=gpurun_out/r2scale; mkdir -p $out
for n in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29540 + n)) bench.py --gpus $n --steps 10 --warmup 3 > $out/bench_c2_n$n.json 2> $out/bench_c2_n$n.err
  echo n=$n rc $?
  python - $out/bench_c2_n$n.json <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print({k: d[k] for k in ("value", "ms_per_step", "n_gpus")}, "e2e", d["e2e"]["value"], "traced-only", d["culling"]["value_traced_only"], "check", d["multi_rank_check"]["ok"], d["multi_rank_check"]["pixels_differing"])
for k, v in d.get("extra", {}).items(): print("  extra", k, round(v["value"], 1), round(v["ms_per_step"], 2), v.get("multi_rank_check", {}).get("ok"))
PY
done
