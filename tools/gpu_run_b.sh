#!/bin/bash
# ncu captures of the solver at the throughput-bound sizes (persistent build)
O=gpurun_out; mkdir -p $O
for n in 65536 32768; do
  ncu --set full --clock-control none --import-source on -k regex:solve_warp_kernel --launch-skip 515 -c 1 -f -o $O/solve_warp_kernel_r2_$n \
    python bench.py --steps 20 --warmup 3 --envs $n --no-cpu-baseline > /dev/null 2>&1
done
ls -la $O/*.ncu-rep
