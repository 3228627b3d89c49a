#!/bin/bash
# C3 experiment matrix (round 2): residency, leaf size and shading batch variants of k_path_persist on the 10.5 M-triangle scene.
# usage: tools/c3_matrix.sh <out-dir>
out=${1:-gpurun_out/c3_matrix}; mkdir -p $out
run() { tag=$1; shift; python bench.py --workload c3 --no-cpu-baseline --no-gpu-baseline --no-extra --steps 3 --warmup 2 "$@" > $out/$tag.json 2> $out/$tag.err; python - $out/$tag.json $tag <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[2], "value", round(d["value"], 1), "ms", round(d["ms_per_step"], 2), "frac", round(d["roofline"]["frac"], 3), "build_ms", round(d["scene"]["build_ms"], 1), "nodes", d["scene"]["n_nodes"], "sah", round(d["scene"]["sah_cost"], 3), flush=True)
except Exception as e:
    print(sys.argv[2], "FAILED", e)
PY
}
run default
run minb4 --min-blocks 4
run minb7 --min-blocks 7
run leaf1 --max-leaf 1
run leaf2 --max-leaf 2
run batch6 --shade-batch 6
run batch16 --shade-batch 16
run batch20 --shade-batch 20
run philox --rng philox
