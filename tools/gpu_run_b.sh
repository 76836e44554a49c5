#!/bin/bash
O=gpurun_out; mkdir -p $O
python tools/solver_ab.py 4096,65536 product build/variants/lib_fidle.so product build/variants/lib_fidle.so 2>&1 | tee $O/ab_fidle.txt
