#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -15 > $O/pytest_gpu.txt; tail -3 $O/pytest_gpu.txt
python tools/solver_ab.py 4096 product product@graph=0 2>&1 | tee $O/ab.txt
python bench.py --steps 300 --warmup 20 > $O/bench_4096.json 2>/dev/null; cat $O/bench_4096.json | head -c 300
