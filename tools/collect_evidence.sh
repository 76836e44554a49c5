#!/bin/bash
# One GPU session's worth of profiler evidence into gpurun_out/ (then tools/save_profiles.py <tag> ...
# copies it into profiles/).  usage (through gpurun): bash tools/collect_evidence.sh <tag>
# Every command below has exited 0 without ncu before (the bench / test sessions); nothing printed
# under ncu is a bench value.
TAG=${1:-r2}
O=gpurun_out
mkdir -p $O
K='setup_legpar_kernel|solve_warp_kernel|finish_kernel'
python bench.py --steps 300 --warmup 20 > $O/bench_4096.json 2> $O/bench_4096.err; head -c 300 $O/bench_4096.json; echo
python bench.py --steps 100 --warmup 10 --envs 65536 --no-cpu-baseline > $O/bench_65536.json 2>/dev/null; head -c 300 $O/bench_65536.json; echo
python bench.py --impl reference --steps 20 --warmup 5 > $O/bench_ref.json 2>/dev/null; head -c 300 $O/bench_ref.json; echo
# launch list of 40 steps inside the timed region (1500 settle launches + warm-up in front)
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"$K" --launch-skip 1509 -c 120 --csv \
  --log-file $O/launches_c.csv python bench.py --steps 40 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
for k in setup_legpar_kernel solve_warp_kernel finish_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$k --launch-skip 520 -c 1 -f -o $O/${k}_$TAG \
    python bench.py --steps 30 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
done
# the throughput-bound sizes: solver (14-block build) and setup at 65536 envs, solver at 32768
for n in 65536 32768; do
  for k in solve_warp_kernel setup_legpar_kernel; do
    ncu --set full --clock-control none --import-source on -k regex:$k --launch-skip 515 -c 1 -f -o $O/${k}_${TAG}_$n \
      python bench.py --steps 20 --warmup 3 --envs $n --no-cpu-baseline > /dev/null 2>&1
  done
done
ls -la $O/*.ncu-rep
