#!/bin/bash
O=gpurun_out; mkdir -p $O
python tools/solver_ab.py 4096,65536 product build/variants/lib_inl.so product build/variants/lib_inl.so 2>&1 | tee $O/ab_inl.txt
