#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_configs.py tests/test_gpu_parity.py tests/test_gpu_reference_properties.py -m gpu -q -x 2>&1 | tail -3
python tools/solver_ab.py 4096,16384,65536 build/variants/lib_base.so product build/variants/lib_base.so product 2>&1 | tee $O/ab_rows3.txt
python tools/solver_ab.py 8192 cfg=2 build/variants/lib_base.so product 2>&1 | tee -a $O/ab_rows3.txt
python tools/solver_ab.py 16384 cfg=3 build/variants/lib_base.so product 2>&1 | tee -a $O/ab_rows3.txt
