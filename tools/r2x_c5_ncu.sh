out=gpurun_out/r2x; mkdir -p $out
python tools/ncu_case.py c5 1920 1080 16 8 3 0 0 2 > $out/case_c5.log 2>&1 && cat $out/case_c5.log && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_path_persist -c 1 -f -o $out/c5_inst python tools/ncu_case.py c5 1920 1080 8 8 3 0 0 1 > $out/ncu_c5.log 2>&1
echo c5 rc $?; tail -3 $out/ncu_c5.log; ls -la $out
