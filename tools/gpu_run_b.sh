#!/bin/bash
O=gpurun_out; mkdir -p $O
for m in 0 1 2; do timeout 600 python tools/sanitizer_run.py build/variants/lib_checks.so $m 64 2>&1 | grep -v Warn | tail -8; done | tee $O/checks_run.txt
timeout 600 python tools/sanitizer_run.py build/variants/lib_checks.so 2 4000 2>&1 | grep -v Warn | tail -8 | tee -a $O/checks_run.txt
