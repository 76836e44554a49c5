#!/bin/bash
O=gpurun_out; mkdir -p $O
python tools/solver_ab.py 16384,65536 product build/variants/lib_blk32.so build/variants/lib_blk16.so 2>&1 | tee $O/ab.txt
python tools/solver_ab.py 32768 cfg=4 product build/variants/lib_blk32.so 2>&1 | tee -a $O/ab.txt
