out=gpurun_out/r2s; mkdir -p $out
for tag in default nocost; do
  flag=""; [ $tag = nocost ] && flag="--no-cost-order"
  for rng in compat philox; do
  python bench.py --workload c3 --no-cpu-baseline --no-gpu-baseline --no-extra --steps 4 --warmup 2 --rng $rng $flag > $out/$tag.$rng.json 2> $out/$tag.$rng.err
  python - $out/$tag.$rng.json $tag.$rng <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[2], "value", round(d["value"], 1), "ms", round(d["ms_per_step"], 2), "frac", round(d["roofline"]["frac"], 3), flush=True)
PY
  done
done
