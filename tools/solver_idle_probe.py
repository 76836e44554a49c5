"""How often does the solver visit a promoted contact that carries no impulse at all?  Needs the
probe build (the "active contacts" statistics row then carries n_active + 1000 * percent idle):

  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -shared -Xcompiler -fPIC -DSOLO_IDLE_PROBE \\
       -o tools/_probe/libidle.so gym_solo_b200/csrc/solo_abi.cu
  python tools/solver_idle_probe.py
"""
import json, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gym_solo_b200 import _lib
_lib.LIB_PATH = os.path.join(ROOT, 'tools', '_probe', 'libidle.so')
_lib._stale = lambda: False
import bench
n = 4096
dev = torch.device('cuda', 0)
env = bench.make_env(n, dev)
gen = torch.Generator(device=dev).manual_seed(1234)
act = lambda: (torch.rand((n, 12), generator=gen, device=dev) * 2 - 1) * 2 * np.pi
for k in range(300):
  env.step(act())
rows = []
for k in range(10):
  for _ in range(5):
    env.step(act())
  torch.cuda.synchronize()
  rows.append(env.client.solver_stats().cpu().numpy().copy())
st = np.concatenate(rows, axis=1)
it, nc, a, cy = st
nact = a % 1000; pct = a // 1000
heavy = it >= 50
print('all envs: mean idle pct (weighted by it*nact)', float((pct * it * nact).sum() / max(1, (it * nact).sum())))
print('capped envs: n', int(heavy.sum()), 'mean nact', float(nact[heavy].mean()), 'mean idle pct', float(pct[heavy].mean()))
w = np.argsort(-cy)[:15]
print('worst:', [(int(it[i]), int(nc[i]), int(nact[i]), int(pct[i]), int(cy[i])) for i in w])
