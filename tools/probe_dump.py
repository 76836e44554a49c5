"""Dump the per-env solver statistics rows of the bench workload to an .npy
(used with the SOLO_PROBE builds of the library, see solo_solve_warp.cuh)."""
import os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
n, samples, out = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3]
dev = torch.device('cuda', 0)
env = bench.make_env(n, dev)
gen = torch.Generator(device=dev).manual_seed(1234)
act = lambda: (torch.rand((n, 12), generator=gen, device=dev) * 2 - 1) * 2 * np.pi
for k in range(200):
  env.step(act())
rows = []
for k in range(samples):
  for _ in range(10):
    env.step(act())
  torch.cuda.synchronize()
  rows.append(env.client.solver_stats().cpu().numpy().copy())
np.save(out, np.concatenate(rows, axis=1))
