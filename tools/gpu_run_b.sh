#!/bin/bash
O=gpurun_out; mkdir -p $O
python tools/solver_ab.py 4096 build/variants/lib_hostprof.so build/variants/lib_hostprof.so@graph=0 2>&1 | tee $O/ab.txt
