#!/bin/bash
# compute-sanitizer (memcheck, racecheck, synccheck) over tools/sanitize_case.py; summaries -> gpurun_out/r02_sanitizer_*.txt
mkdir -p gpurun_out
CS=/usr/local/cuda/bin/compute-sanitizer
for tool in memcheck racecheck synccheck; do
  timeout 900 $CS --tool $tool --print-limit 20 python tools/sanitize_case.py > gpurun_out/r02_sanitizer_$tool.txt 2>&1
  echo "exit $?" >> gpurun_out/r02_sanitizer_$tool.txt
  tail -4 gpurun_out/r02_sanitizer_$tool.txt
done
