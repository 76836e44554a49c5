#!/bin/bash
O=gpurun_out; mkdir -p $O
python tools/solver_ab.py 4096 product build/variants/lib_out.so product build/variants/lib_out.so 2>&1 | tee $O/ab_out.txt
python tools/solver_ab.py 65536 product build/variants/lib_out.so 2>&1 | tee -a $O/ab_out.txt
