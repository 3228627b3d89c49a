#!/bin/bash
# 8-GPU records of the final round-2 code: the default bench line (C2 + extras) and C5 at its full configuration
out=gpurun_out/r2n8; mkdir -p $out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533"
$T bench.py --gpus 8 --steps 10 --warmup 3 > $out/bench_c2_n8.json 2> $out/bench_c2_n8.err; echo c2 rc $?; tail -c 400 $out/bench_c2_n8.json
$T bench.py --gpus 8 --workload c5 --steps 2 --warmup 1 --no-check --no-extra --no-cpu-baseline --no-gpu-baseline > $out/bench_c5_n8.json 2> $out/bench_c5_n8.err; echo c5 rc $?; tail -c 300 $out/bench_c5_n8.json
