#!/bin/bash
O=gpurun_out; mkdir -p $O
python tools/solver_ab.py 4096 build/variants/lib_pipe_f0.so build/variants/lib_pin.so build/variants/lib_fmin.so build/variants/lib_pinfmin.so build/variants/lib_pipe_f0.so build/variants/lib_pin.so build/variants/lib_fmin.so build/variants/lib_pinfmin.so 2>&1 | tee $O/ab_pipe4.txt
