#!/bin/bash
# compute-sanitizer evidence: memcheck, racecheck, initcheck, synccheck over tools/sanitize_case.py.
# Usage (GPU box): bash tools/sanitize.sh <out_dir>
out=${1:-gpurun_out/sanitize}
mkdir -p "$out"
for tool in memcheck racecheck initcheck synccheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 --log-file "$out/$tool.log" python tools/sanitize_case.py > "$out/$tool.stdout" 2>&1
  echo "$tool exit $?" >> "$out/summary.txt"
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY" "$out/$tool.log" >> "$out/summary.txt"
done
cat "$out/summary.txt"
