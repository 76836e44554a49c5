"""Generate tests/golden/epilogue_golden.npz by running the REFERENCE's own
Python classes (gym_solo/core/obs.py, rewards.py, termination.py and
Solo8VanillaEnv.step's action normalisation) on seeded states.

Runs only in the authoring container: /root/reference is imported with stub
`pybullet`, `pybullet_utils`, `pybullet_data`, `gym` and `deprecation` modules
(none is installed; the reference files themselves are unmodified).  The stub
client answers the BulletClient getters from a flat state vector; its
getEulerFromQuaternion is Bullet's published formula --
that one function is pybullet's, not the reference's, and is pinned separately
by the reference's known answer (test_obs_observations.py:67-88).

usage: python tools/gen_golden.py   (writes tests/golden/epilogue_golden.npz)
"""
import math
import os
import sys
import types

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = '/root/reference'


# ---- stub third-party modules -------------------------------------------------
class Box:
  def __init__(self, low, high, shape=None, dtype=np.float32):
    if shape is None:
      low = np.asarray(low); high = np.asarray(high)
      shape = low.shape
    self.shape = tuple(shape)
    self.low = np.broadcast_to(np.asarray(low, dtype=dtype), self.shape).copy()
    self.high = np.broadcast_to(np.asarray(high, dtype=dtype), self.shape).copy()


def install_stubs():
  gym = types.ModuleType('gym')
  gym.Env = object
  gym.Space = object
  spaces = types.ModuleType('gym.spaces')
  spaces.Box = Box
  spaces.Space = object
  space_mod = types.ModuleType('gym.spaces.space')
  spaces.space = space_mod
  gym.spaces = spaces
  envs = types.ModuleType('gym.envs')
  reg = types.ModuleType('gym.envs.registration')
  reg.register = lambda **kw: None
  envs.registration = reg
  gym.envs = envs
  sys.modules.update({'gym': gym, 'gym.spaces': spaces, 'gym.spaces.space': space_mod,
                      'gym.envs': envs, 'gym.envs.registration': reg})
  p = types.ModuleType('pybullet')
  p.POSITION_CONTROL, p.DIRECT, p.GUI, p.URDF_USE_INERTIA_FROM_FILE = 2, 2, 1, 2
  sys.modules['pybullet'] = p
  pu = types.ModuleType('pybullet_utils')
  bc = types.ModuleType('pybullet_utils.bullet_client')
  bc.BulletClient = object
  pu.bullet_client = bc
  sys.modules.update({'pybullet_utils': pu, 'pybullet_utils.bullet_client': bc})
  pbd = types.ModuleType('pybullet_data')
  pbd.getDataPath = lambda: ''
  sys.modules['pybullet_data'] = pbd
  dep = types.ModuleType('deprecation')
  dep.deprecated = lambda **kw: (lambda f: f)
  sys.modules['deprecation'] = dep


def euler_from_quat(q):
  x, y, z, w = q
  sarg = -2.0 * (x * z - w * y)
  if sarg <= -0.99999:
    return (0.0, -0.5 * math.pi, 2 * math.atan2(x, -y))
  if sarg >= 0.99999:
    return (0.0, 0.5 * math.pi, 2 * math.atan2(-x, y))
  return (math.atan2(2 * (y * z + w * x), w * w - x * x - y * y + z * z), math.asin(sarg),
          math.atan2(2 * (x * y + w * z), w * w + x * x - y * y - z * z))


API_NAMES = ['FL_HFE', 'FL_KFE', 'FL_ANKLE', 'FR_HFE', 'FR_KFE', 'FR_ANKLE',
             'HL_HFE', 'HL_KFE', 'HL_ANKLE', 'HR_HFE', 'HR_KFE', 'HR_ANKLE']
API_DOF = [0, 1, -1, 2, 3, -1, 4, 5, -1, 6, 7, -1]
NDOF = 8


class StubClient:
  """BulletClient getters answered from one state vector
  (pos 3, quat 4, lin 3, ang 3, q 8, qdot 8)."""

  def __init__(self):
    self.s = None
    self.motor_calls = []

  def getBasePositionAndOrientation(self, body):
    return tuple(self.s[0:3]), tuple(self.s[3:7])

  def getEulerFromQuaternion(self, q):
    return euler_from_quat(q)

  def getBaseVelocity(self, body):
    return tuple(self.s[7:10]), tuple(self.s[10:13])

  def getNumJoints(self, body):
    return 12

  def getJointInfo(self, body, j):
    return (j, API_NAMES[j].encode(), 4 if API_DOF[j] < 0 else 0, -1, -1, 0, 0., 0., -10., 10.,
            1000., 1000., b'')

  def getJointState(self, body, j):
    d = API_DOF[j]
    if d < 0:
      return (0.0, 0.0, (0.,) * 6, 0.0)
    return (self.s[13 + d], self.s[13 + NDOF + d], (0.,) * 6, 0.0)

  def setJointMotorControlArray(self, *a, **kw):
    self.motor_calls.append(np.array(kw['targetPositions'], dtype=np.float64))

  def stepSimulation(self):
    pass


def random_states(rng, n):
  s = np.zeros((n, 13 + 2 * NDOF))
  s[:, 0:3] = rng.uniform(-1, 1, (n, 3)) * [2, 2, 0.6] + [0, 0, 0.3]
  q = rng.normal(size=(n, 4))
  s[:, 3:7] = q / np.linalg.norm(q, axis=1, keepdims=True)
  s[:, 7:10] = rng.normal(size=(n, 3)) * 3
  s[:, 10:13] = rng.normal(size=(n, 3)) * 6
  s[:, 13:13 + NDOF] = rng.uniform(-7, 7, (n, NDOF))
  s[:, 13 + NDOF:] = rng.normal(size=(n, NDOF)) * 8
  # a few special poses: level, gimbal lock both ways, large velocities (clipping)
  s[0, 3:7] = [0, 0, 0, 1]; s[0, 7:13] = 0; s[0, 13 + NDOF:] = 0; s[0, 2] = 0.33698
  s[1, 3:7] = [0, math.sin(math.pi / 4), 0, math.cos(math.pi / 4)]
  s[2, 3:7] = [0, -math.sin(math.pi / 4), 0, math.cos(math.pi / 4)]
  s[3, 7:13] = [20, -20, 16, 12, -12, 11]
  s[4, 3:7] = [0, 0, .707, .707]
  return s


def main():
  install_stubs()
  sys.path.insert(0, REF)
  from gym_solo.core import obs as robs, rewards as rrew, termination as rterm
  from gym_solo.envs.solo8v2vanilla import Solo8VanillaEnv, Solo8VanillaConfig

  rng = np.random.RandomState(20261104)
  states = random_states(rng, 96)
  client = StubClient()
  out = {'states': states}

  def per_state(fn):
    rows = []
    for s in states:
      client.s = s
      rows.append(np.asarray(fn(), dtype=np.float64))
    return np.stack(rows)

  # ---- observations --------------------------------------------------------------
  obs_cases = {
      'imu_rad': lambda: robs.TorsoIMU(1),
      'imu_deg': lambda: robs.TorsoIMU(1, degrees=True, max_lin_velocity=5, max_angular_velocity=180),
      'enc_rad': lambda: robs.MotorEncoder(1),
      'enc_deg': lambda: robs.MotorEncoder(1, degrees=True),
      'enc_clip': lambda: robs.MotorEncoder(1, max_rotation=2.5),
      'enc_deg_clip': lambda: robs.MotorEncoder(1, degrees=True, max_rotation=90.),
  }
  for name, mk in obs_cases.items():
    o = mk(); o.client = client
    out['obs_' + name] = per_state(o.compute)
    out['obs_' + name + '_low'] = np.asarray(o.observation_space.low, dtype=np.float64)
    out['obs_' + name + '_high'] = np.asarray(o.observation_space.high, dtype=np.float64)
  # factory: concatenation + normalisation (obs.py:130-159)
  for norm in (False, True):
    for combo in (('imu_rad', 'enc_rad'), ('imu_deg', 'enc_clip'), ('enc_deg_clip', 'imu_rad')):
      client.s = states[0]
      f = robs.ObservationFactory(client, normalize=norm)
      for c in combo:
        f.register_observation(obs_cases[c]())
      key = 'factory_%s_%s_%s' % (combo[0], combo[1], 'norm' if norm else 'raw')
      out[key] = per_state(lambda: f.get_obs()[0])
      client.s = states[0]
      out[key + '_labels'] = np.array(f.get_obs()[1])
      sp = f.get_observation_space()
      out[key + '_low'] = np.asarray(sp.low, dtype=np.float64)
      out[key + '_high'] = np.asarray(sp.high, dtype=np.float64)

  # ---- rewards -------------------------------------------------------------------
  def with_client(r):
    r.client = client
    return r
  rew_cases = {
      'upright': lambda: with_client(rrew.UprightReward(1)),
      'flat_default': lambda: with_client(rrew.FlatTorsoReward(1)),
      'flat_wide': lambda: with_client(rrew.FlatTorsoReward(1, hard_margin=.1, soft_margin=np.pi)),
      'flat_step': lambda: with_client(rrew.FlatTorsoReward(1, hard_margin=1.0, soft_margin=0.)),
      'small_default': lambda: with_client(rrew.SmallControlReward(1)),
      'small_10': lambda: with_client(rrew.SmallControlReward(1, margin=10)),
      'hspeed_0': lambda: with_client(rrew.HorizontalMoveSpeedReward(1, 0, hard_margin=.5, soft_margin=3)),
      'hspeed_2': lambda: with_client(rrew.HorizontalMoveSpeedReward(1, 2.0)),
      'height_home': lambda: with_client(rrew.TorsoHeightReward(1, 0.33698, 0.025, 0.15)),
      'height_default': lambda: with_client(rrew.TorsoHeightReward(1, 0.3)),
  }
  for name, mk in rew_cases.items():
    r = mk()
    out['rew_' + name] = per_state(r.compute)

  def home_pos():
    # examples/solo8_vanilla/interactive_pos_control.py:22-35, verbatim structure
    flat = rrew.FlatTorsoReward(1, hard_margin=.1, soft_margin=np.pi)
    height = rrew.TorsoHeightReward(1, 0.33698, 0.025, 0.15)
    small_control = rrew.SmallControlReward(1, margin=10)
    no_move = rrew.HorizontalMoveSpeedReward(1, 0, hard_margin=.5, soft_margin=3)
    stand = rrew.AdditiveReward()
    stand.client = client
    stand.add_term(0.5, flat)
    stand.add_term(0.5, height)
    home = rrew.MultiplicitiveReward(1, stand, small_control, no_move)
    return home
  f = rrew.RewardFactory(client)
  f.register_reward(1, home_pos())
  out['rew_home_pos'] = per_state(f.get_reward)
  f = rrew.RewardFactory(client)
  f.register_reward(1, rrew.UprightReward(1))
  f.register_reward(1, home_pos())
  out['rew_upright_plus_home'] = per_state(f.get_reward)
  f = rrew.RewardFactory(client)
  f.register_reward(0.25, rrew.UprightReward(1))
  f.register_reward(-2.0, rrew.SmallControlReward(1, margin=10))
  f.register_reward(3.0, rrew.MultiplicitiveReward(0.5, rrew.FlatTorsoReward(1), rrew.TorsoHeightReward(1, 0.3)))
  out['rew_mixed'] = per_state(f.get_reward)

  xs = np.linspace(-3, 3, 61)
  out['gauss_x'] = xs
  out['gauss_a'] = np.array([rrew.gaussian(float(x), bounds=(-.5, .5), margin=1.) for x in xs])
  out['gauss_b'] = np.array([rrew.gaussian(float(x), bounds=(0., 0.), margin=.3, margin_value=.5) for x in xs])
  out['gauss_c'] = np.array([rrew.gaussian(float(x), bounds=(1., 2.), margin=0.) for x in xs])
  out['linear_a'] = np.array([rrew.linear(float(x), 1., 1.5) for x in xs])
  out['linear_b'] = np.array([rrew.linear(float(x), 0., -2., symmetric=True) for x in xs])

  # ---- terminations ----------------------------------------------------------------
  tf = rterm.TerminationFactory()
  tf.register_termination(rterm.TimeBasedTermination(5), rterm.PerpetualTermination(),
                          rterm.TimeBasedTermination(3))
  seq = []
  for i in range(12):
    seq.append(tf.is_terminated())
    if i == 7:
      tf.reset()
  out['term_seq'] = np.array(seq)
  out['term_counters_end'] = np.array([tf._terminations[0].step_delta, tf._terminations[2].step_delta])

  # ---- action normalisation (solo8v2vanilla.py:84-90) --------------------------------
  env = Solo8VanillaEnv.__new__(Solo8VanillaEnv)
  env._normalize, env._realtime = True, False
  env.config = Solo8VanillaConfig()
  env.client = client
  env.robot = 1
  env._action_space = Box(-env.config.max_motor_rotation, env.config.max_motor_rotation, shape=(12,))
  env.obs_factory = robs.ObservationFactory(client)
  env.obs_factory.register_observation(robs.TorsoIMU(1))
  env.reward_factory = rrew.RewardFactory(client)
  env.reward_factory.register_reward(1, rrew.UprightReward(1))
  env.termination_factory = rterm.TerminationFactory()
  env.termination_factory.register_termination(rterm.PerpetualTermination())
  acts = rng.uniform(-1, 1, (16, 12))
  client.motor_calls = []
  client.s = states[0]
  for a in acts:
    env.step(a)
  out['norm_actions_in'] = acts
  out['norm_actions_out'] = np.stack(client.motor_calls)

  path = os.path.join(ROOT, 'tests', 'golden', 'epilogue_golden.npz')
  np.savez_compressed(path, **out)
  print('wrote', path, len(out), 'arrays')


if __name__ == '__main__':
  main()
