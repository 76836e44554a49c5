#!/bin/bash
# One GPU session's worth of evidence into gpurun_out/ (then tools/save_profiles.py <tag> ...
# copies it into profiles/).  usage (through gpurun): bash tools/collect_evidence.sh <tag>
TAG=${1:-r1x}
O=gpurun_out
mkdir -p $O
K='setup_legpar_kernel|solve_warp_kernel|finish_kernel'
python -m pytest tests -m gpu -q 2>&1 | tail -3 > $O/pytest_gpu.txt; cat $O/pytest_gpu.txt
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
python bench.py --steps 300 --warmup 20 > $O/bench_4096.json 2> $O/bench_4096.err; cat $O/bench_4096.json
python bench.py --steps 100 --warmup 10 --envs 65536 --no-cpu-baseline > $O/bench_65536.json 2>/dev/null; cat $O/bench_65536.json
python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.json 2>/dev/null; cat $O/bench_ref.json
python tools/bench_render.py 4096 1024 20 > $O/render_1024.json 2>/dev/null; cat $O/render_1024.json
python tools/bench_render.py 4096 1 200 > $O/render_1.json 2>/dev/null; cat $O/render_1.json
# profiler passes last: nothing printed under ncu is a bench value
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"$K" --launch-skip 1509 -c 120 --csv \
  --log-file $O/launches_c.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
for k in setup_legpar_kernel solve_warp_kernel finish_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$k --launch-skip 520 -c 1 -f -o $O/${k}_$TAG \
    python bench.py --steps 10 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
done
for k in render_project_kernel render_raster_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$k --launch-skip 3 -c 1 -f -o $O/${k}_$TAG \
    python tools/bench_render.py 4096 1024 2 > /dev/null 2>&1
done
ls -la $O/*_$TAG.ncu-rep
