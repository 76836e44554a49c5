#!/bin/bash
O=gpurun_out; mkdir -p $O
python tools/solver_ab.py 4096,16384,65536 product build/variants/lib_setup2.so product build/variants/lib_setup2.so 2>&1 | tee $O/ab_setup2.txt
