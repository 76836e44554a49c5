#!/bin/bash
O=gpurun_out; mkdir -p $O
python tools/solver_ab.py 32768,65536 product@occ=2 build/variants/lib_p13.so@occ=2 build/variants/lib_p13t.so@occ=2 product@occ=2 build/variants/lib_p13.so@occ=2 2>&1 | tee $O/ab_p13.txt
python tools/solver_ab.py 16384 product@occ=0 build/variants/lib_p13.so@occ=2 2>&1 | tee -a $O/ab_p13.txt
