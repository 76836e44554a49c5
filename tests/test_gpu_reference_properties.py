"""What the reference itself holds about the dynamics, on the CUDA path (run with -m gpu).

pybullet is on neither box, so no post-step numbers of the real engine can be recorded; these
are the things the reference DOES pin, ported through the drop-in surface:

* the live-pybullet property tests of gym_solo/envs/test_solo8v2vanilla.py: rest-state
  stability to 6 decimals after 1000 zero-action steps and the response to action 5.0
  (:77-104), reset determinism with the termination reset counter (:141-163), two disjoint
  environments giving the same reset observation (:177-194) -- at the reference's own
  dt = 1e-3 and its default configuration;
* the standing height the examples hard-code, TorsoHeightReward(env.robot, 0.33698, ...)
  (examples/solo8_vanilla/interactive_pos_control.py:23): a number that came out of the real
  engine and checks the foot hull's support point, the 1 mm collision margin, the slop and
  the erp in one figure (0.16 + 0.16 + 0.015993 + 0.001 - slop), together
  with sum of normal forces = total mass (2.17785 kg) x 9.81;
* engine-independent invariants of the articulated-body step (free flight: no contact,
  motors off through torque mode): free fall leaves the joints alone, Bullet's body damping
  in closed form, conservation of linear / angular momentum and energy without damping;
* the PGS iteration counts of the solver warp against the double-precision oracle's.
"""
import warnings

import numpy as np
import pytest
import torch

from gym_solo_b200 import _abi
from gym_solo_b200 import testing
from helpers import model_and_config, oracle, state_errors

pytestmark = pytest.mark.gpu

STANDING_HEIGHT = 0.33698      # examples/solo8_vanilla/interactive_pos_control.py:23
TOTAL_MASS = 2.17785           # SURVEY App. A (sum of the URDF masses)


def _assert_almost_equal6(a, b):
  """np.testing.assert_array_almost_equal's default: |a - b| < 1.5e-6."""
  np.testing.assert_array_almost_equal(np.asarray(a), np.asarray(b), decimal=6)


def _not_almost_equal6(a, b):
  return bool((np.abs(np.asarray(a) - np.asarray(b)) >= 1.5e-6).any())


def _reference_env(n, **kw):
  """setUp of the reference's TestSolo8v2VanillaEnv: default config (dt = 1e-3)."""
  from gym_solo_b200.envs import Solo8VanillaEnv, Solo8VanillaConfig
  return Solo8VanillaEnv(config=Solo8VanillaConfig(), num_envs=n, device='cuda:0', **kw)


def _base_pose(env):
  pos, orn = env.client.getBasePositionAndOrientation(env.robot)
  return pos.double().cpu().numpy().copy(), orn.double().cpu().numpy().copy()


def test_actions_rest_state_and_response():
  """test_solo8v2vanilla.py:77-104 (test_actions), every environment of a small batch."""
  n = 4
  env = _reference_env(n)
  env.obs_factory.register_observation(testing.CompliantObs(None))
  env.termination_factory.register_termination(testing.DummyTermination(0, True))
  with warnings.catch_warnings():
    warnings.simplefilter('ignore', DeprecationWarning)
    env.reward_factory.register_reward(1, testing.SimpleReward())
  env.reset()
  no_op = torch.zeros((n, env.action_space.shape[0]), device='cuda:0')
  # "Let the robot stabilize first"
  for _ in range(1000):
    env.step(no_op)
  position, orientation = _base_pose(env)
  # the zero action is the straight-legged stand: the height the examples use as the target
  assert np.abs(position[:, 2] - STANDING_HEIGHT).max() <= 3e-5, position[:, 2]
  for _ in range(10):
    env.step(no_op)
  new_pos, new_or = _base_pose(env)
  _assert_almost_equal6(position, new_pos)
  _assert_almost_equal6(orientation, new_or)
  action = torch.full((n, env.action_space.shape[0]), 5., device='cuda:0')
  for _ in range(10):
    env.step(action)
  new_pos, new_or = _base_pose(env)
  assert _not_almost_equal6(position, new_pos)
  assert _not_almost_equal6(orientation, new_or)
  env.close()


def test_reset_restores_the_base_pose_and_counts():
  """test_solo8v2vanilla.py:141-163 (test_reset)."""
  n = 3
  env = _reference_env(n)
  env.obs_factory.register_observation(testing.CompliantObs(None))
  env.termination_factory.register_termination(testing.DummyTermination(0, True))
  with warnings.catch_warnings():
    warnings.simplefilter('ignore', DeprecationWarning)
    env.reward_factory.register_reward(1, testing.SimpleReward())
  env.reset()
  term = env.termination_factory._terminations[0]
  counter0 = term.reset_counter
  base_pos, base_or = _base_pose(env)
  action = torch.full((n, env.action_space.shape[0]), 5., device='cuda:0')
  for _ in range(100):
    env.step(action)
  assert term.reset_counter == counter0
  new_pos, new_or = _base_pose(env)
  assert _not_almost_equal6(base_pos, new_pos)
  assert _not_almost_equal6(base_or, new_or)
  env.reset()
  assert term.reset_counter == counter0 + 1
  new_pos, new_or = _base_pose(env)
  _assert_almost_equal6(base_pos, new_pos)
  _assert_almost_equal6(base_or, new_or)
  env.close()


def test_disjoint_environments_share_the_reset_observation():
  """test_solo8v2vanilla.py:177-194 (test_disjoint_environments)."""
  from gym_solo_b200.core import obs as solo_obs

  def make(n):
    env = _reference_env(n)
    env.obs_factory.register_observation(solo_obs.TorsoIMU(env.robot))
    env.obs_factory.register_observation(solo_obs.MotorEncoder(env.robot))
    with warnings.catch_warnings():
      warnings.simplefilter('ignore', DeprecationWarning)
      env.reward_factory.register_reward(1, testing.SimpleReward())
    env.termination_factory.register_termination(testing.DummyTermination(0, True))
    return env

  env1 = make(2)
  home_position = env1.reset().double().cpu().numpy().copy()
  g = torch.Generator(device='cuda:0').manual_seed(7)
  for _ in range(1000):
    env1.step((torch.rand((2, 12), generator=g, device='cuda:0') * 2 - 1) * 2 * np.pi)   # action_space.sample()
  env2 = make(5)
  home2 = env2.reset().double().cpu().numpy()
  _assert_almost_equal6(home_position[0], home_position[1])
  for i in range(5):
    _assert_almost_equal6(home_position[0], home2[i])
  env1.close(); env2.close()


@pytest.mark.parametrize('dt', [1e-3, 1.0 / 240.0])
def test_standing_height_and_weight(dt):
  """Upright robot, straight legs, zero targets: base z settles at the examples' 0.33698 and the
  feet carry the robot's weight."""
  from gym_solo_b200.client import BatchedClient
  blob, model, config = model_and_config(dt=dt)
  n = 8
  c = BatchedClient(blob, model, config, n, 'cuda:0')
  nf_dev, _ = c.enable_debug(True)
  c.set_state(c.initial_state((0., 0., 0.36), (0., 0., 0.)))
  zero = torch.zeros((12,), device='cuda:0')
  c.simulate(zero, n_steps=int(round(3.0 / dt)), broadcast=True)
  st = c.get_state().double().cpu().numpy()
  nf = nf_dev.t().double().cpu().numpy()          # of the last step
  assert np.abs(st[:, 2] - STANDING_HEIGHT).max() <= 3e-5, st[:, 2]
  assert np.abs(nf.sum(axis=1) - TOTAL_MASS * 9.81).max() <= 2e-3, nf.sum(axis=1)
  # on the four feet, nothing else: support points of the lower-leg bodies only
  body_of = np.array(blob['cand_body'])
  loaded = np.unique(body_of[np.nonzero(nf[0] > 1e-3)[0]])
  assert set(loaded.tolist()) == {2, 4, 6, 8}
  assert np.abs(st[:, 13:21]).max() <= 1e-3 and np.abs(st[:, 21:]).max() <= 1e-3
  # the oracle stands at the same height
  o = oracle(model, config)
  s = o.initial_state(pos=(0., 0., 0.36))
  for _ in range(int(round(3.0 / dt))):
    s = o.step(s, np.zeros(12))
  assert abs(s[2] - STANDING_HEIGHT) <= 1e-5
  assert np.abs(st[:, 2] - s[2]).max() <= 2e-5
  c.disconnect()


# ---- free flight: no contact, motors off (torque mode, zero torque) ------------------------
def _free_client(n, dt, gravity=(0., 0., 0.), damping=0.0):
  from gym_solo_b200.client import BatchedClient
  blob, model, config = model_and_config(dt=dt, control_mode=_abi.CONTROL_TORQUE, gravity=gravity,
                                         linear_damping=damping, angular_damping=damping)
  return BatchedClient(blob, model, config, n, 'cuda:0'), model, config


def _momenta(o, states):
  """Total linear momentum, angular momentum about the world origin and kinetic energy, from the
  oracle's dense mass matrix in (v_com0, omega0, qdot) coordinates: p = M v is conjugate to
  (base translation, base rotation about its COM, joints)."""
  out = []
  for s in states:
    nd = (len(s) - 13) // 2
    v = np.concatenate([s[7:10], s[10:13], s[13 + nd:]])
    M = o.mass_matrix(s)
    p = M @ v
    out.append(np.concatenate([p[0:3], p[3:6] + np.cross(s[0:3], p[0:3]), [0.5 * v @ p]]))
  return np.array(out)


def test_free_fall_leaves_the_joints_alone():
  """Gravity accelerates every body alike: a robot released at rest keeps q, qdot and omega at
  zero and v_z = -g t exactly (semi-implicit Euler), whatever the pose -- the bias forces and
  the mass matrix must cancel (no damping)."""
  n, dt, steps = 64, 1.0 / 240.0, 48
  c, model, config = _free_client(n, dt, gravity=(0., 0., -9.81))
  rng = np.random.RandomState(3)
  s0 = np.zeros((n, 29))
  s0[:, 2] = 10.0
  q = rng.normal(size=(n, 4)); s0[:, 3:7] = q / np.linalg.norm(q, axis=1, keepdims=True)
  s0[:, 13:21] = rng.uniform(-3, 3, (n, 8))
  c.set_state(s0)
  c.simulate(torch.zeros((12,), device='cuda:0'), n_steps=steps, broadcast=True)
  st = c.get_state().double().cpu().numpy()
  assert np.abs(st[:, 13:21] - s0[:, 13:21]).max() <= 2e-5
  assert np.abs(st[:, 21:]).max() <= 2e-4 and np.abs(st[:, 10:13]).max() <= 2e-4
  assert np.abs(st[:, 9] + 9.81 * dt * steps).max() <= 1e-4
  assert np.abs(st[:, 7:9]).max() <= 1e-5
  z_want = 10.0 - 9.81 * dt * dt * steps * (steps + 1) / 2
  assert np.abs(st[:, 2] - z_want).max() <= 1e-4
  c.disconnect()


def test_body_damping_closed_form():
  """Bullet's multibody damping f = -m v (k + k |v|) on every body: a robot
  translating without rotation decelerates as one body, v' = v (1 - dt k (1 + |v|)), joints at rest."""
  n, dt, steps, k = 32, 1.0 / 240.0, 60, 0.04
  c, model, config = _free_client(n, dt, damping=k)
  rng = np.random.RandomState(4)
  s0 = np.zeros((n, 29))
  s0[:, 2] = 10.0
  s0[:, 6] = 1.0
  s0[:, 13:21] = rng.uniform(-2, 2, (n, 8))
  s0[:, 7:10] = rng.uniform(-3, 3, (n, 3))
  c.set_state(s0)
  c.simulate(torch.zeros((12,), device='cuda:0'), n_steps=steps, broadcast=True)
  st = c.get_state().double().cpu().numpy()
  v = s0[:, 7:10].astype(np.float32).astype(np.float64)
  for _ in range(steps):
    v = v * (1 - dt * k * (1 + np.linalg.norm(v, axis=1, keepdims=True)))
  assert np.abs(st[:, 7:10] - v).max() <= 2e-5
  assert np.abs(st[:, 21:]).max() <= 2e-4 and np.abs(st[:, 10:13]).max() <= 2e-4
  c.disconnect()


def test_momentum_and_energy_without_damping():
  """No gravity, no damping, no contact, no motors: total linear momentum, angular momentum
  about the world origin and kinetic energy are constants of the motion.  The integrator is
  first order, so the drift is O(dt) per unit time; it must shrink with dt and match the
  oracle's drift (same discretisation) far more closely than it drifts (float32 host build of the
  core: 1e-6 relative against the oracle, drift 4e-3 at dt = 1e-3 and 1e-3 at dt = 2.5e-4)."""
  n, horizon = 48, 0.1
  rng = np.random.RandomState(5)
  s0 = np.zeros((n, 29))
  s0[:, 2] = 10.0
  q = rng.normal(size=(n, 4)); s0[:, 3:7] = q / np.linalg.norm(q, axis=1, keepdims=True)
  s0[:, 7:10] = rng.uniform(-1, 1, (n, 3))
  s0[:, 10:13] = rng.uniform(-2, 2, (n, 3))
  s0[:, 13:21] = rng.uniform(-3, 3, (n, 8))
  s0[:, 21:] = rng.uniform(-5, 5, (n, 8))
  s0 = s0.astype(np.float32).astype(np.float64)
  drift = {}
  for dt in (1e-3, 2.5e-4):
    steps = int(round(horizon / dt))
    c, model, config = _free_client(n, dt)
    o = oracle(model, config)
    m0 = _momenta(o, s0)
    c.set_state(s0)
    c.simulate(torch.zeros((12,), device='cuda:0'), n_steps=steps, broadcast=True)
    st = c.get_state().double().cpu().numpy()
    m1 = _momenta(o, st)
    want = o.step_batch(s0, np.zeros((n, 12)), n_steps=steps, n_threads=4)
    mo = _momenta(o, want)
    scale = np.abs(m0).max(axis=0) + 1e-9
    drift[dt] = (np.abs(m1 - m0) / np.maximum(np.abs(m0), 0.05 * scale)).max(axis=0)
    # the kernel drifts like the oracle (same semi-implicit scheme)
    assert (np.abs(m1 - mo) / scale).max() <= 1e-4, (dt, np.abs(m1 - mo).max(axis=0))
    err = state_errors(st, want, model.ndof)
    assert err['q'] <= 1e-4 and err['base_rot'] <= 1e-4 and err['qd'] <= 1e-3, err
    c.disconnect()
  print('relative drift of (p, L, E) over %.2f s:' % horizon, drift)
  # first-order drift (in these coordinates the total linear momentum depends on q too):
  # a 4x smaller time step must give a clearly smaller drift of all seven constants
  assert (drift[2.5e-4] <= 0.5 * drift[1e-3] + 1e-4).all()
  assert drift[1e-3].max() <= 2e-2 and drift[2.5e-4].max() <= 5e-3


def test_pgs_iteration_counts_match_the_oracle():
  """The solver warp's early exit (least-squares residual <= 1e-7, Bullet's default) against the
  double-precision oracle's on the rollout pool: iteration counts per environment."""
  from test_gpu_parity import make_client, rollout_states
  n = 256
  for dt in (1.0 / 240.0, 1e-3):
    c, model, config = make_client(n, dt=dt)
    states = rollout_states(model, config, n, seed=1)
    rng = np.random.RandomState(2)
    actions = rng.uniform(-2 * np.pi, 2 * np.pi, (n, 12))
    c.set_state(states)
    c.simulate(torch.tensor(actions, dtype=torch.float32, device='cuda:0'))
    got = c.solver_stats()[0].cpu().numpy().astype(int)
    o = oracle(model, config)
    want = np.zeros(n, dtype=int)
    for i in range(n):
      s32 = states[i].astype(np.float32).astype(np.float64)
      a32 = actions[i].astype(np.float32).astype(np.float64)
      _, d = o.step(s32, a32, diag=True)
      want[i] = d['iterations']
    diff = got - want
    vals, cnts = np.unique(diff, return_counts=True)
    print('dt %.5f  iterations kernel - oracle: %s   (oracle mean %.2f, kernel mean %.2f, capped %d / %d)'
          % (dt, dict(zip(vals.tolist(), cnts.tolist())), want.mean(), got.mean(),
             int((want == 50).sum()), int((got == 50).sum())))
    assert (np.abs(diff) <= 1).mean() >= 0.97, dict(zip(vals.tolist(), cnts.tolist()))
    assert got.sum() <= want.sum() * 1.02 + n     # no float32 tail of extra iterations
    c.disconnect()


# ---- the rest of gym_solo/envs/test_solo8v2vanilla.py, through the drop-in surface -------
@pytest.mark.parametrize('limit', [2 * np.pi, 1.5])
def test_action_space_follows_the_configured_limit(limit):
  """test_solo8v2vanilla.py:55-64: 12 joints (8 dof + 4 fixed ankles), +-max_motor_rotation."""
  from gym_solo_b200 import spaces
  from gym_solo_b200.envs import Solo8VanillaEnv, Solo8VanillaConfig
  config = Solo8VanillaConfig(motor_torque_limit=limit)
  config.max_motor_rotation = limit        # a class attribute in the reference too (configs.py:20)
  env = Solo8VanillaEnv(config=config, device='cuda:0')
  assert env.action_space == spaces.Box(-limit, limit, shape=(12,))
  assert env.action_space.sample().shape == (12,)
  env._close()


def test_normalized_and_invalid_action_space():
  """test_solo8v2vanilla.py:66-74."""
  from gym_solo_b200 import spaces
  env = _reference_env(1, normalize_actions=True)
  assert env.action_space == spaces.Box(-1, 1, shape=(12,))
  env._action_space = None
  with pytest.raises(ValueError):
    env.action_space
  env._close()


def test_step_with_nothing_registered_is_refused():
  """test_solo8v2vanilla.py:165-168: a step without rewards raises ValueError."""
  env = _reference_env(2)
  with pytest.raises(ValueError):
    env.step(np.zeros(12))
  env._close()


def test_step_simple_reward_and_result_shapes():
  """test_solo8v2vanilla.py:170-175: reward == 1 with SimpleReward, on every environment, and
  the (obs, reward, done, info) tuple of the reference with a leading num_envs axis."""
  env = _reference_env(3)
  env.reward_factory.register_reward(1, testing.SimpleReward())
  env.obs_factory.register_observation(testing.CompliantObs(None))
  env.termination_factory.register_termination(testing.DummyTermination(0, True))
  obs, reward, done, info = env.step(np.stack([env.action_space.sample() for _ in range(3)]))
  assert tuple(obs.shape) == (3, 2) and tuple(reward.shape) == (3,) and tuple(done.shape) == (3,)
  assert (reward.cpu().numpy() == 1).all()
  assert done.cpu().numpy().all()              # DummyTermination(0, True) is always done
  assert 'labels' in info and len(info['labels']) == 2
  env._close()


def test_realtime_sleeps_one_dt_per_step(monkeypatch):
  """test_solo8v2vanilla.py:37-47: realtime=True sleeps between steps (time.sleep patched out
  the way the reference's test does)."""
  import time
  slept = []
  monkeypatch.setattr(time, 'sleep', lambda s: slept.append(s))
  env = _reference_env(1, realtime=True)
  env.reward_factory.register_reward(1, testing.SimpleReward())
  env.obs_factory.register_observation(testing.CompliantObs(None))
  env.termination_factory.register_termination(testing.DummyTermination(0, True))
  env.step(env.action_space.sample())
  assert slept and slept[-1] == env.config.dt
  env._close()
