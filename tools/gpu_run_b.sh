#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_render.py -m gpu -q -x 2>&1 | tail -12
python tools/bench_render.py 4096 1024 20 2>/dev/null | tee $O/render_1024.json
python tools/bench_render.py 4096 1 200 2>/dev/null | tee $O/render_1.json
