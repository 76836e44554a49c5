"""Per-environment solver statistics of the bench workload: how the solve time of
one env.step is distributed over environments (iterations, contacts, contacts
that carried an impulse, SM cycles), i.e. what the tail of a launch looks like.

  python tools/solver_hist.py [n_envs] [n_sampled_steps]
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
samples = int(sys.argv[2]) if len(sys.argv) > 2 else 20
dev = torch.device('cuda', 0)
env = bench.make_env(n, dev)
gen = torch.Generator(device=dev).manual_seed(1234)


def act():
  return (torch.rand((n, 12), generator=gen, device=dev) * 2 - 1) * 2 * np.pi


for k in range(200):
  env.step(act())
rows = []
for k in range(samples):
  for _ in range(10):
    env.step(act())
  torch.cuda.synchronize()
  rows.append(env.client.solver_stats().cpu().numpy().copy())
st = np.concatenate(rows, axis=1)   # [4][samples * n]: iters, nc, active, cycles
it, nc, ac, cy = st[0], st[1], st[2], st[3]
out = {'n_envs': n, 'samples': samples}
out['cycles_percentiles'] = {str(p): float(np.percentile(cy, p)) for p in (50, 90, 99, 99.9, 100)}
out['per_launch_max_cycles_mean'] = float(np.mean([r[3].max() for r in rows]))
out['per_launch_sum_cycles_mean'] = float(np.mean([r[3].sum() for r in rows]))
out['iters_hist'] = {str(b): int(((it >= b) & (it < b2)).sum()) for b, b2 in
                     ((0, 3), (3, 5), (5, 9), (9, 17), (17, 33), (33, 50), (50, 51))}
out['nc_percentiles'] = {str(p): float(np.percentile(nc, p)) for p in (50, 90, 99, 100)}
out['active_percentiles'] = {str(p): float(np.percentile(ac, p)) for p in (50, 90, 99, 100)}
# least squares: cycles ~ a + it * (b + c * nc + d * active)
A = np.stack([np.ones_like(it), it, it * nc, it * ac], axis=1)
coef, *_ = np.linalg.lstsq(A, cy, rcond=None)
out['fit_cycles=a+it*(b+c*nc+d*active)'] = [float(x) for x in coef]
cap = it >= 50
out['capped'] = {'frac': float(cap.mean()), 'mean_nc': float(nc[cap].mean()) if cap.any() else 0,
                 'mean_active': float(ac[cap].mean()) if cap.any() else 0,
                 'mean_cycles': float(cy[cap].mean()) if cap.any() else 0,
                 'max_cycles': float(cy[cap].max()) if cap.any() else 0}
worst = np.argsort(-cy)[:10]
out['worst_iters_nc_active_cycles'] = [[int(it[i]), int(nc[i]), int(ac[i]), int(cy[i])] for i in worst]
print(json.dumps(out))
