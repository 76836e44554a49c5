#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -5 > $O/pytest_gpu.txt; tail -2 $O/pytest_gpu.txt
python tools/solver_ab.py 4096,65536 product 2>&1 | tee $O/ab.txt
python tools/solver_ab.py 32768 cfg=4 product 2>&1 | tee -a $O/ab.txt
