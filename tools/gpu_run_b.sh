#!/bin/bash
O=gpurun_out; mkdir -p $O
python bench.py --steps 100 --warmup 10 --envs 65536 --no-cpu-baseline > $O/bench_65536.json 2>/dev/null; head -c 250 $O/bench_65536.json; echo
python bench.py --config 4 --steps 60 --warmup 10 --no-cpu-baseline > $O/bench_c4_1gpu.json 2>/dev/null; head -c 250 $O/bench_c4_1gpu.json; echo
