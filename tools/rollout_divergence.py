"""Short-horizon rollout divergence (north_star: "short-horizon rollout divergence is reported"):
the CUDA path (float32) and the double-precision oracle run the same random-action sequences from
the reset snapshot; per horizon the median / 90th percentile / maximum over the environments of the
state error groups.  Contact makes the map chaotic, so this is a report, not a tolerance.

  python tools/rollout_divergence.py [n_envs] [dt]  -> one JSON line
"""
import json, os, sys
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'tests'))
from helpers import START_TARGETS, model_and_config, oracle, quat_angle
from gym_solo_b200.client import BatchedClient

n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
dt = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0 / 240.0
blob, model, config = model_and_config(dt=dt)
c = BatchedClient(blob, model, config, n, 'cuda:0')
o = oracle(model, config)
snap = o.settle(START_TARGETS).astype(np.float32).astype(np.float64)
states = np.tile(snap, (n, 1))
c.set_state(states)
rng = np.random.RandomState(7)
horizons = (1, 10, 100, 1000)
out = {'n_envs': n, 'dt': dt, 'actions': 'U(-2pi, 2pi)^12, held for 20 steps', 'horizons': {}}
a = None
for k in range(1, horizons[-1] + 1):
  if (k - 1) % 20 == 0:
    a = rng.uniform(-2 * np.pi, 2 * np.pi, (n, 12)).astype(np.float32)
  c.simulate(torch.tensor(a, device='cuda:0'))
  states = o.step_batch(states, a.astype(np.float64), n_steps=1, n_threads=8)
  if k in horizons:
    g = c.get_state().double().cpu().numpy()
    err = dict(base_pos=np.abs(g[:, 0:3] - states[:, 0:3]).max(1), base_rot=quat_angle(g[:, 3:7], states[:, 3:7]),
               q=np.abs(g[:, 13:21] - states[:, 13:21]).max(1), qd=np.abs(g[:, 21:] - states[:, 21:]).max(1))
    out['horizons'][k] = {key: dict(median=float(np.median(v)), p90=float(np.percentile(v, 90)), max=float(v.max()))
                          for key, v in err.items()}
print(json.dumps(out))
