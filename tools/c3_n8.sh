#!/bin/bash
# C3 (10.5 M triangles, 1080p, 256 spp) on 8 GPUs: tiles + reference RNG, sample ranges + counter RNG
out=gpurun_out/r2c3n8; mkdir -p $out
for rng in compat philox; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29571 bench.py --gpus 8 --workload c3 --rng $rng --steps 5 --warmup 2 --no-extra --no-cpu-baseline --no-gpu-baseline > $out/bench_c3_n8_$rng.json 2> $out/bench_c3_n8_$rng.err; echo $rng rc $?
done
python - <<'PY'
import json
for r in ("compat", "philox"):
    d = json.loads(open(f"gpurun_out/r2c3n8/bench_c3_n8_{r}.json").read().strip().splitlines()[-1])
    print(r, {k: d[k] for k in ("value", "ms_per_step", "n_gpus")}, "frac", d["roofline"]["frac"], d.get("multi_rank_check"))
PY
