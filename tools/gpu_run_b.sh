#!/bin/bash
O=gpurun_out; mkdir -p $O
python tools/solver_ab.py 4096,65536 product product@blob=build/solo8v2_r1_geometry.json 2>&1 | tee $O/ab.txt
