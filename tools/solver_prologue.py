"""Where an AVERAGE environment's solver cycles go (probe build): record copy, contact listing, row
build, iterations.

  nvcc ... -DSOLO_PROLOGUE_PROBE -o build/variants/lib_pro.so solo_abi.cu
  python tools/solver_prologue.py build/variants/lib_pro.so [n_envs]
"""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gym_solo_b200 import _lib
_lib.LIB_PATH = os.path.abspath(sys.argv[1])
import bench

n = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
dev = torch.device('cuda', 0)
env = bench.make_env(1, n, dev)
K = bench.CONFIGS[1]
gen = torch.Generator(device=dev).manual_seed(1234)
acts = [(torch.rand((n, K['n_api']), generator=gen, device=dev) * 2 - 1) * float(K['bound']) for _ in range(64)]
for k in range(40): env.step(acts[k % 64])
acc = []
for k in range(40, 60):
  env.step(acts[k % 64])
  acc.append(env.client.solver_stats().double().cpu().numpy())
st = np.stack(acc)            # [launch, 4, n]
plain = st[:, 0] < 0
st[:, 0] = np.abs(st[:, 0])
print('records that came the plain way: %.1f %%; record copy, staged: %.0f cycles, plain: %.0f'
      % (100 * plain.mean(), st[:, 0][~plain].mean() if (~plain).any() else 0, st[:, 0][plain].mean() if plain.any() else 0))
tot = st.sum(1)
print('%d environments, 20 launches, cycles per environment (mean | median | p99):' % n)
for name, x in (('record -> shared memory', st[:, 0]), ('contact listing', st[:, 1]), ('row build', st[:, 2]),
                ('iterations', st[:, 3]), ('sum', tot)):
  print('  %-24s %8.0f | %8.0f | %8.0f   share of all cycles %.1f %%' % (name, x.mean(), np.median(x), np.percentile(x, 99),
                                                                       100 * x.sum() / tot.sum()))
