#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_configs.py -m gpu -q -x 2>&1 | tail -3
python bench.py --config 3 --steps 100 --warmup 10 > $O/bench_c3_1gpu.json 2>/dev/null; head -c 250 $O/bench_c3_1gpu.json; echo
