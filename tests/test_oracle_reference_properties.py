"""The dynamics oracle (oracle/solo_oracle.c) against what the reference holds about the real
engine's behaviour (CPU; the CUDA path gets the same checks in test_gpu_reference_properties.py):

* examples/solo8_vanilla/interactive_pos_control.py:23 hard-codes the standing height
  0.33698 = 0.16 + 0.16 + foot hull 0.015993 + 1 mm margin - slop: a number out
  of pybullet itself;
* gym_solo/envs/test_solo8v2vanilla.py:77-104: after reset and 1000 zero-action steps (dt 1e-3)
  the base pose is at rest to 6 decimals, action 5.0 moves it; :141-163, :177-194 reset determinism.
"""
import numpy as np

from helpers import START_TARGETS, model_and_config, oracle

STANDING_HEIGHT = 0.33698
TOTAL_MASS = 2.17785


def _reset(o):
  s = o.initial_state()                 # (0, 0, 0.5), zero joints: solo8v2vanilla.py:151-155
  for _ in range(500):                  # solo8v2vanilla.py:124-136
    s = o.step(s, START_TARGETS)
  return s


def test_rest_state_response_and_standing_height():
  blob, model, config = model_and_config(dt=1e-3)
  o = oracle(model, config)
  s = _reset(o)
  home = s.copy()
  no_op = np.zeros(12)
  for _ in range(1000):
    s = o.step(s, no_op)
  pose = s[:7].copy()
  # zero targets = straight legs: the folded robot has stood up, to the examples' height
  assert abs(pose[2] - STANDING_HEIGHT) <= 1e-5, pose[2]
  for _ in range(10):
    s, d = o.step(s, no_op, diag=True)
  np.testing.assert_array_almost_equal(pose, s[:7])           # 6 decimals, like the reference
  assert abs(d['normal_force'].sum() - TOTAL_MASS * 9.81) <= 1e-3
  body_of = np.array(blob['cand_body'])
  assert set(np.unique(body_of[d['normal_force'] > 1e-3]).tolist()) == {2, 4, 6, 8}   # the four feet
  for _ in range(10):
    s = o.step(s, no_op + 5.0)
  assert (np.abs(s[:3] - pose[:3]) >= 1.5e-6).any() and (np.abs(s[3:7] - pose[3:7]) >= 1.5e-6).any()
  # reset is deterministic (test_reset, test_disjoint_environments)
  np.testing.assert_array_equal(_reset(o), home)


def test_standing_height_at_the_bench_time_step():
  blob, model, config = model_and_config(dt=1.0 / 240.0)
  o = oracle(model, config)
  s = o.initial_state(pos=(0., 0., 0.36))
  for _ in range(720):
    s, d = o.step(s, np.zeros(12), diag=True)
  assert abs(s[2] - STANDING_HEIGHT) <= 1e-5, s[2]
  assert abs(d['normal_force'].sum() - TOTAL_MASS * 9.81) <= 1e-3
