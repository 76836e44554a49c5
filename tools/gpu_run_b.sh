#!/bin/bash
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_render.py -m gpu -q -x 2>&1 | tail -2
python tools/bench_render.py 4096 1024 20 2>/dev/null | tee $O/render_1024.json
python tools/bench_render.py 4096 1 200 2>/dev/null | tee $O/render_1.json
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -k regex:"render_raster" --launch-skip 3 -c 1 --csv --log-file $O/render_launches.csv python tools/bench_render.py 4096 1024 3 > /dev/null 2>&1
grep -v "^==" $O/render_launches.csv | cut -d, -f13- | tail -2
