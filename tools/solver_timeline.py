"""When does every solver block of one launch start and end?  Needs the probe build of the
library (the per-env statistics rows carry %globaltimer instead of contact counts):

  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -shared -Xcompiler -fPIC -DSOLO_TIMELINE \
       -o tools/_probe/libtimeline.so gym_solo_b200/csrc/solo_abi.cu
  python tools/solver_timeline.py [n_envs]

Prints, for a few launches: the span of the launch, when the last block started, the longest
environments (start, duration, iterations, cycles) and how many blocks were running every 10 us.
"""
import json, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gym_solo_b200 import _lib
_lib.LIB_PATH = os.path.join(ROOT, 'tools', '_probe', 'libtimeline.so')
_lib._stale = lambda: False
import bench
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
dev = torch.device('cuda', 0)
env = bench.make_env(n, dev)
gen = torch.Generator(device=dev).manual_seed(1234)
act = lambda: (torch.rand((n, 12), generator=gen, device=dev) * 2 - 1) * 2 * np.pi
for k in range(300):
  env.step(act())
res = []
for k in range(8):
  for _ in range(5):
    env.step(act())
  torch.cuda.synchronize()
  st = env.client.solver_stats().cpu().numpy().copy()
  b = st[1].view(np.uint32).astype(np.int64); e = st[2].view(np.uint32).astype(np.int64)
  t0 = b.min(); b -= t0; e -= t0
  dur = e - b
  w = int(np.argmax(e))
  order = np.argsort(-dur)[:5]
  running = [int(((b <= t) & (e > t)).sum()) for t in range(0, int(e.max()), 10000)]
  res.append(dict(span_us=float(e.max()) / 1e3, last_start_us=float(b.max()) / 1e3,
                  start_pct_us=[float(np.percentile(b, p)) / 1e3 for p in (50, 90, 99)],
                  last_to_end=dict(env=w, start_us=float(b[w]) / 1e3, dur_us=float(dur[w]) / 1e3, iters=int(st[0][w]),
                                   cycles=int(st[3][w])),
                  longest=[dict(env=int(i), start_us=float(b[i]) / 1e3, dur_us=float(dur[i]) / 1e3, iters=int(st[0][i]),
                                cycles=int(st[3][i])) for i in order],
                  running_every_10us=running))
print(json.dumps(res))
