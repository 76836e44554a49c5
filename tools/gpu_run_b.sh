#!/bin/bash
O=gpurun_out; mkdir -p $O
python tools/solver_ab.py 4096,65536 build/variants/lib_base.so build/variants/lib_nobudget.so 2>&1 | tee $O/ab.txt
