#!/bin/bash
O=gpurun_out; mkdir -p $O
python tools/solver_ab.py 8192 cfg=2 product product@zero_copy=0 2>&1 | tee $O/ab_zc.txt
python tools/solver_ab.py 16384,65536 product product@zero_copy=0 2>&1 | tee -a $O/ab_zc.txt
python tools/solver_ab.py 262144 cfg=4 product product@zero_copy=0 2>&1 | tee -a $O/ab_zc.txt
