#!/bin/bash
O=gpurun_out; mkdir -p $O
python tools/solver_ab.py 4096,16384,65536 product build/variants/lib_mpark.so 2>&1 | tee $O/ab.txt
python tools/solver_ab.py 8192 cfg=2 product build/variants/lib_mpark.so 2>&1 | tee -a $O/ab.txt
python tools/solver_ab.py 16384 cfg=3 product build/variants/lib_mpark.so 2>&1 | tee -a $O/ab.txt
