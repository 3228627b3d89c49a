out=gpurun_out/r2af; mkdir -p $out
B="python bench.py --steps 2 --warmup 3 --no-extra --no-cpu-baseline --no-gpu-baseline"
$B > $out/bench_plain.json 2> $out/bench_plain.err && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/launches_c2.csv $B > $out/ncu_launches.log 2>&1
echo launches rc $?
python tools/ncu_case.py c2 1920 1080 256 16 6 0 0 1 > $out/case_c2.log 2>&1 && cat $out/case_c2.log && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_path_flat -c 1 -f -o $out/c2_flat python tools/ncu_case.py c2 1920 1080 256 16 6 0 0 1 > $out/ncu_c2.log 2>&1
echo c2 rc $?
ls -la $out
