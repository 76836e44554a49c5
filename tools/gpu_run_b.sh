#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -15 > $O/pytest_gpu.txt; tail -3 $O/pytest_gpu.txt
timeout 600 python tools/sanitizer_run.py build/variants/lib_checks.so > $O/checks_run.txt 2>&1; tail -8 $O/checks_run.txt
python tools/solver_ab.py 4096,16384,65536 product 2>&1 | tee $O/ab.txt
