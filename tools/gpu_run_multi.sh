#!/bin/bash
# usage (through gpurun --gpus N): bash tools/gpu_run_multi.sh N  -> bench lines of configs 1, 2, 4 on N GPUs
N=${1:-2}
O=gpurun_out; mkdir -p $O
for C in 1 2 4; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + C)) \
    bench.py --gpus $N --config $C --steps 100 --warmup 10 > $O/bench_c${C}_${N}gpu.json 2> $O/bench_c${C}_${N}gpu.err
  head -c 400 $O/bench_c${C}_${N}gpu.json; echo
done
