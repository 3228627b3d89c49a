#!/bin/bash
# C4 at its full configuration (3840x2160, 4096 spp, depth 8) on 1 and 8 GPUs
out=gpurun_out/r2c4; mkdir -p $out
python bench.py --workload c4 --steps 2 --warmup 1 --no-extra --no-cpu-baseline --no-gpu-baseline > $out/bench_c4_n1.json 2> $out/bench_c4_n1.err; echo n1 rc $?
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus 8 --workload c4 --steps 3 --warmup 1 --no-extra --no-cpu-baseline --no-gpu-baseline > $out/bench_c4_n8.json 2> $out/bench_c4_n8.err; echo n8 rc $?
python - <<'PY'
import json
for n in (1, 8):
    d = json.loads(open(f"gpurun_out/r2c4/bench_c4_n{n}.json").read().strip().splitlines()[-1])
    print(n, {k: d[k] for k in ("value", "ms_per_step", "n_gpus")}, d["config"]["workload"], "traced-only", d["culling"]["value_traced_only"], d.get("multi_rank_check"))
PY
