"""Where the slowest environments of a solve launch spend their cycles (probe build).

  nvcc ... -DSOLO_SECTION_PROBE [-DSOLO_V_...] -o build/variants/lib_sec.so solo_abi.cu
  python tools/solver_sections.py build/variants/lib_sec.so [n_envs]

The probe build overwrites the per-environment contact counts of solo_solver_stats with the cycles
spent in the normal rows (+ active contacts in the low three digits) and in the friction pairs
(+ screens / 10 in the low three digits).  Bench workload of configs[1], 40 warm-up steps.
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
rows = []
for k in range(40, 60):
  env.step(acts[k % 64])
  st = env.client.solver_stats().double().cpu().numpy()
  i = int(st[3].argmax())
  tot = st[3][i]
  cm, it = divmod(st[0][i], 100); cm *= 100
  cn, nu = divmod(st[1][i], 1000); cn *= 100
  cf, fu = divmod(st[2][i], 1000); cf *= 100
  rows.append((tot, it, cm, cn, cf, nu, fu))
  print('slowest env: %7.0f cycles, %2.0f iterations | motor rows %6.0f | normal rows %7.0f (%3.0f updates) | friction %7.0f (%3.0f pairs) | rest %6.0f'
        % (tot, it, cm, cn, nu, cf, fu, tot - cm - cn - cf))
r = np.array(rows)
# per-iteration cost against rows per iteration, over the slow environments of the last launch
it = st[0] % 100
sel = np.argsort(-st[3])[:400]
sel = sel[it[sel] >= 10]
x_n = (st[1][sel] % 1000) / it[sel]; y_n = (st[1][sel] // 1000) * 100 / it[sel]
x_f = (st[2][sel] % 1000) / it[sel]; y_f = (st[2][sel] // 1000) * 100 / it[sel]
for name, x, y in (('normal', x_n, y_n), ('friction', x_f, y_f)):
  A = np.stack([np.ones_like(x), x], axis=1)
  coef, *_ = np.linalg.lstsq(A, y, rcond=None)
  print('%s rows, %d environments with >= 10 iterations: cycles per iteration = %.0f + %.1f x (updates per iteration), rms %.0f'
        % (name, len(sel), coef[0], coef[1], np.sqrt(np.mean((A @ coef - y) ** 2))))
print('motor rows over those: %.0f cycles per iteration' % np.mean((st[0][sel] // 100) * 100 / it[sel]))

print('mean over %d launches: total %.0f; per iteration: motor %.0f, normal %.0f, friction %.0f; per normal update %.0f, per friction pair %.0f; outside the sections %.0f'
      % (len(r), r[:, 0].mean(), (r[:, 2] / r[:, 1]).mean(), (r[:, 3] / r[:, 1]).mean(), (r[:, 4] / r[:, 1]).mean(),
         (r[:, 3] / r[:, 5]).mean(), (r[:, 4] / r[:, 6]).mean(), (r[:, 0] - r[:, 2] - r[:, 3] - r[:, 4]).mean()))
