#!/bin/bash
# round 2 GPU session: the whole -m gpu suite, smoke, bench lines and the box probe
O=gpurun_out
mkdir -p $O
{ python -c "import pybullet" 2>&1 | tail -1; python -m pip list 2>/dev/null | grep -i -E "bullet|gym" || echo "pip list: no bullet / gym packages"; nvidia-smi --query-gpu=name,driver_version --format=csv,noheader; lscpu | grep -E "Model name|^CPU\(s\)"; } > $O/box_probe.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q --durations=8 -s 2>&1 | tail -250 > $O/pytest_gpu.txt; tail -30 $O/pytest_gpu.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python bench.py --steps 300 --warmup 20 > $O/bench_4096.json 2> $O/bench_4096.err; cat $O/bench_4096.json
timeout 600 python bench.py --steps 300 --warmup 20 --stream default --no-cpu-baseline > $O/bench_4096_default_stream.json 2>/dev/null; cat $O/bench_4096_default_stream.json
