#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -15 > $O/pytest_gpu.txt; tail -3 $O/pytest_gpu.txt
python tools/solver_ab.py 4096,8192 product@warm=0 product@warm=1 product@warm=0 product@warm=1 2>&1 | tee $O/ab.txt
python tools/solver_ab.py 16384,65536 product@warm=0 product@warm=1 2>&1 | tee -a $O/ab.txt
python tools/solver_ab.py 32768 cfg=4 product@warm=0 product@warm=1 2>&1 | tee -a $O/ab.txt
