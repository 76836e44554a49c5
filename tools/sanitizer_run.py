"""Workload for a -DSOLO_CHECKS build of the library (every index into the solver warp's shared
arrays checked; compute-sanitizer itself is closed on this pool, profiles/r2_compute_sanitizer_closed.txt):
  python tools/sanitizer_run.py build/variants/lib_checks.so [solver build 0|1|2] [environments]
64 environments, reset (the
500 settle steps), 50 env.steps with random actions, then a belly-flop state with more than 32
contacts so that the two-rows-per-lane instantiation warp_solve<CH, 2> and its promotions run,
the torque / limit-row path, and one render call.

  compute-sanitizer --tool racecheck python tools/sanitizer_run.py
"""
import os, sys, warnings
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
if len(sys.argv) > 1:
  from gym_solo_b200 import _lib
  _lib.LIB_PATH = os.path.abspath(sys.argv[1])
import bench

from gym_solo_b200 import _abi
dev = torch.device('cuda', 0)
# optional second argument: solver build (SOLO_OPT_SOLVER_OCCUPANCY 0 / 1 / 2) and batch size
mode = int(sys.argv[2]) if len(sys.argv) > 2 else -1
n = int(sys.argv[3]) if len(sys.argv) > 3 else 64
env = bench.make_env(1, n, dev)
env.client.set_option(_abi.OPT_SOLVER_OCCUPANCY, mode)
print('solver build', mode, 'environments', n)
g = torch.Generator(device=dev).manual_seed(3)
for k in range(50):
  o, r, d, _ = env.step((torch.rand((n, 12), generator=g, device=dev) * 2 - 1) * 2 * np.pi)
torch.cuda.synchronize()
st = env.client.solver_stats().cpu().numpy()
print('steps done; promoted contacts max', int(st[2].max()), 'contacts max', int(st[1].max()))
# belly flop: folded robot dropped flat -> more than 32 support points inside the threshold
s = env.get_state().clone()
s[:, 2] = 0.03
s[:, 3:7] = torch.tensor([0., 0., 0., 1.], device=dev)
s[:, 7:13] = 0
env.set_state(s)
folded = torch.tensor([np.pi / 2, np.pi, 0, np.pi / 2, np.pi, 0, -np.pi / 2, -np.pi, 0, -np.pi / 2, -np.pi, 0],
                      dtype=torch.float32, device=dev)
worst = 0
for k in range(30):
  env.step(folded + (torch.rand((n, 12), generator=g, device=dev) * 2 - 1) * 0.3)
  worst = max(worst, int(env.client.solver_stats()[1].max()))
print('belly flop: contacts max', worst)
env.step_host(np.zeros((n, 12), dtype=np.float32))
img = env.render()
print('render', img.shape)
env.close()
# torque mode with joint-limit rows
env = bench.make_env(4, n, dev)
env.client.set_option(_abi.OPT_SOLVER_OCCUPANCY, mode)
for k in range(60):
  env.step(torch.full((n, 12), 2.0, device=dev))
torch.cuda.synchronize()
print('torque/limits: max |q|', float(env.get_state()[:, 13:21].abs().max()))
env.close()
print('sanitizer workload finished')
