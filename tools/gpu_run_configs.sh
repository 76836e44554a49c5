#!/bin/bash
# one GPU: bench lines of configs 2, 3, 4 (configs[4]: 262144 envs on one GPU) + reference arms
O=gpurun_out; mkdir -p $O
for C in 2 3 4; do
  python bench.py --config $C --steps 100 --warmup 10 > $O/bench_c${C}_1gpu.json 2> $O/bench_c${C}_1gpu.err
  head -c 300 $O/bench_c${C}_1gpu.json; echo
  python bench.py --impl reference --config $C --steps 3 --warmup 1 > $O/bench_c${C}_ref.json 2>/dev/null
done
python bench.py --impl reference --steps 20 --warmup 5 > $O/bench_c1_ref.json 2>/dev/null; head -c 300 $O/bench_c1_ref.json
