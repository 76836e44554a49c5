#!/bin/bash
O=gpurun_out; mkdir -p $O
for tool in memcheck racecheck synccheck; do
  timeout 1500 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitizer_run.py > $O/sanitizer_$tool.txt 2>&1
  tail -6 $O/sanitizer_$tool.txt
done
