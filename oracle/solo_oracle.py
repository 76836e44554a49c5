"""CPU oracle for the Solo 8 env-step path.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this.  Two halves:

* physics: ctypes over oracle/libsolo_oracle.so (solo_oracle.c, double
  precision).  PARITY UNPINNED against pybullet (see that file's header).
* observations / rewards / terminations: numpy float64 restatement of the
  reference's pure-Python code, function by function, batched over a leading
  env axis.  PINNED: tests/golden/epilogue_golden.npz was produced by the
  reference's own classes (tools/gen_golden.py imports /root/reference with
  stub pybullet/gym modules) and tests/test_oracle_golden.py checks this file
  against it, plus the reference tests' known answers.
"""
import ctypes as C
import math
import os
import subprocess

import numpy as np

from gym_solo_b200 import _abi

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


class OracleDiag(C.Structure):
  _fields_ = [('n_contacts', C.c_int32), ('iterations', C.c_int32),
              ('normal_force', C.c_double * _abi.MAX_CAND),
              ('motor_torque', C.c_double * _abi.MAX_DOF)]


def build(force=False):
  so = os.path.join(_HERE, 'libsolo_oracle.so')
  src = os.path.join(_HERE, 'solo_oracle.c')
  hdr = os.path.join(_HERE, '..', 'include', 'solo8_b200.h')
  if (force or not os.path.exists(so)
      or os.path.getmtime(so) < max(os.path.getmtime(src), os.path.getmtime(hdr))):
    subprocess.check_call(['make', '-C', _HERE, '-B', 'libsolo_oracle.so'],
                          stdout=subprocess.DEVNULL)
  return so


def lib():
  global _LIB
  if _LIB is None:
    so = os.path.join(_HERE, 'libsolo_oracle.so')
    if not os.path.exists(so):
      build()
    _LIB = C.CDLL(so)
    _LIB.oracle_step.restype = C.c_int
    _LIB.oracle_step_batch.restype = C.c_int
    assert _LIB.oracle_sizeof_model() == C.sizeof(_abi.SoloModel)
    assert _LIB.oracle_sizeof_config() == C.sizeof(_abi.SoloConfig)
    assert _LIB.oracle_sizeof_diag() == C.sizeof(OracleDiag)
  return _LIB


def _dp(a):
  return a.ctypes.data_as(C.POINTER(C.c_double))


class PhysicsOracle:
  """Double-precision Bullet-style step for a batch of independent envs."""

  def __init__(self, model, config):
    self.model, self.config = model, config
    self.ndof, self.n_api = model.ndof, model.n_api
    self.sdim = _abi.state_dim(model.ndof)

  def initial_state(self, pos=(0., 0., .5), euler=(0., 0., 0.)):
    """State right after loadURDF (solo8v2vanilla.py:151-155): start pose,
    zero joint angles, zero velocity."""
    s = np.zeros(self.sdim)
    s[0:3] = pos
    s[3:7] = quat_from_euler(euler)
    return s

  def step(self, state, action, diag=False):
    """state [sdim] (copied), action [n_api] -> new state (and diag dict)."""
    st = np.ascontiguousarray(state, dtype=np.float64).copy()
    a = np.ascontiguousarray(action, dtype=np.float64)
    assert st.shape == (self.sdim,) and a.shape == (self.n_api,)
    d = OracleDiag()
    rc = lib().oracle_step(C.byref(self.model), C.byref(self.config), _dp(st), _dp(a),
                           C.byref(d))
    if rc:
      raise RuntimeError('oracle_step failed rc=%d' % rc)
    if diag:
      return st, dict(n_contacts=d.n_contacts, iterations=d.iterations,
                      normal_force=np.array(d.normal_force[:self.model.n_cand]),
                      motor_torque=np.array(d.motor_torque[:self.ndof]))
    return st

  def step_batch(self, states, actions, n_steps=1, n_threads=1):
    """states [n, sdim], actions [n, n_api]; in place on a copy."""
    st = np.ascontiguousarray(states, dtype=np.float64).copy()
    a = np.ascontiguousarray(actions, dtype=np.float64)
    assert st.ndim == 2 and st.shape[1] == self.sdim and a.shape == (st.shape[0], self.n_api)
    rc = lib().oracle_step_batch(C.byref(self.model), C.byref(self.config), _dp(st), _dp(a),
                                 st.shape[0], n_steps, n_threads)
    if rc:
      raise RuntimeError('oracle_step_batch failed rc=%d' % rc)
    return st

  def forward_dynamics(self, state, tau):
    st = np.ascontiguousarray(state, dtype=np.float64)
    t = np.ascontiguousarray(tau, dtype=np.float64)
    acc = np.zeros(6 + self.ndof)
    rc = lib().oracle_forward_dynamics(C.byref(self.model), C.byref(self.config), _dp(st), _dp(t),
                                       _dp(acc))
    assert rc == 0
    return acc

  def mass_matrix(self, state):
    st = np.ascontiguousarray(state, dtype=np.float64)
    n = 6 + self.ndof
    M = np.zeros((n, n))
    lib().oracle_mass_matrix(C.byref(self.model), _dp(st), _dp(M))
    return M

  def support_points(self, state):
    st = np.ascontiguousarray(state, dtype=np.float64)
    out = np.zeros((self.model.n_cand, 3))
    lib().oracle_support_points(C.byref(self.model), _dp(st), _dp(out))
    return out

  def settle(self, start_targets_api, n_steps=500, pos=(0., 0., .5), euler=(0., 0., 0.)):
    """reset()'s 500 POSITION_CONTROL steps toward starting_joint_pos
    (solo8v2vanilla.py:124-136).  Always position control, one substep."""
    s = self.initial_state(pos, euler)
    cfg = _abi.SoloConfig.from_buffer_copy(self.config)
    cfg.control_mode = _abi.CONTROL_POSITION
    cfg.num_substeps = 1
    a = np.ascontiguousarray(start_targets_api, dtype=np.float64)
    for _ in range(n_steps):
      rc = lib().oracle_step(C.byref(self.model), C.byref(cfg), _dp(s), _dp(a), None)
      assert rc == 0
    return s


# ---- pybullet getters restated (recollection) ---------------

def quat_from_euler(e):
  """pybullet.getQuaternionFromEuler (solo8v2vanilla.py:153): XYZ fixed-axis
  roll/pitch/yaw -> (x, y, z, w)."""
  r, p, y = (0.5 * float(v) for v in e)
  cr, sr, cp, sp, cy, sy = math.cos(r), math.sin(r), math.cos(p), math.sin(p), math.cos(y), math.sin(y)
  return np.array([sr * cp * cy - cr * sp * sy, cr * sp * cy + sr * cp * sy,
                   cr * cp * sy - sr * sp * cy, cr * cp * cy + sr * sp * sy])


def euler_from_quat(q):
  """pybullet.getEulerFromQuaternion (obs.py:271, rewards.py:233,265), batched
  [..., 4] (x, y, z, w) -> [..., 3] (roll, pitch, yaw).  The quaternion is NOT
  normalised first (golden: test_obs_observations.py:67-88 feeds (0,0,.707,.707)
  and expects yaw = pi/2)."""
  q = np.asarray(q, dtype=np.float64)
  x, y, z, w = q[..., 0], q[..., 1], q[..., 2], q[..., 3]
  sarg = -2.0 * (x * z - w * y)
  sqx, sqy, sqz, sqw = x * x, y * y, z * z, w * w
  roll = np.arctan2(2 * (y * z + w * x), sqw - sqx - sqy + sqz)
  pitch = np.arcsin(np.clip(sarg, -1.0, 1.0))
  yaw = np.arctan2(2 * (x * y + w * z), sqw + sqx - sqy - sqz)
  hi, lo = sarg >= 0.99999, sarg <= -0.99999
  roll = np.where(hi | lo, 0.0, roll)
  pitch = np.where(hi, 0.5 * np.pi, np.where(lo, -0.5 * np.pi, pitch))
  yaw = np.where(hi, 2 * np.arctan2(-x, y), np.where(lo, 2 * np.arctan2(x, -y), yaw))
  return np.stack([roll, pitch, yaw], axis=-1)


class StateView:
  """Batched stand-in for the BulletClient getters the plug-ins call."""

  def __init__(self, states, model):
    s = np.asarray(states, dtype=np.float64)
    self.s = s[None] if s.ndim == 1 else s
    self.ndof, self.n_api = model.ndof, model.n_api
    self.api_dof = [model.api_dof[i] for i in range(model.n_api)]
    self.api_lower = np.array([model.api_lower[i] for i in range(model.n_api)])
    self.api_upper = np.array([model.api_upper[i] for i in range(model.n_api)])

  @property
  def pos(self): return self.s[:, 0:3]
  @property
  def quat(self): return self.s[:, 3:7]
  @property
  def lin(self): return self.s[:, 7:10]
  @property
  def ang(self): return self.s[:, 10:13]

  def joint_pos(self):
    """getJointState(i)[0] for all API joints; fixed joints read 0 (App. B5)."""
    out = np.zeros((self.s.shape[0], self.n_api))
    for j, d in enumerate(self.api_dof):
      if d >= 0:
        out[:, j] = self.s[:, _abi.STATE_BASE + d]
    return out

  def joint_vel(self):
    out = np.zeros((self.s.shape[0], self.n_api))
    for j, d in enumerate(self.api_dof):
      if d >= 0:
        out[:, j] = self.s[:, _abi.STATE_BASE + self.ndof + d]
    return out


# ---- observations (gym_solo/core/obs.py) ---------------------------------------

def torso_imu(view, degrees=False, max_lin_velocity=15, max_angular_velocity=10.):
  """TorsoIMU.compute (obs.py:260-282) with the bounds of
  TorsoIMU.observation_space (obs.py:230-258)."""
  amax = 180. if degrees else np.pi
  low = np.array([-amax] * 3 + [-max_lin_velocity] * 3 + [-max_angular_velocity] * 3, dtype=np.float64)
  high = -low
  orien = euler_from_quat(view.quat)
  v_lin, v_ang = view.lin, view.ang
  if degrees:
    orien = np.degrees(orien)
    v_ang = np.degrees(v_ang)
  raw = np.concatenate([orien, v_lin, v_ang], axis=-1)
  return np.clip(raw, low, high), low, high


def motor_encoder(view, degrees=False, max_rotation=None):
  """MotorEncoder.compute / observation_space (obs.py:310-334, 347-363)."""
  vals = view.joint_pos()
  if degrees:
    vals = np.degrees(vals)
  if max_rotation:
    vals = np.clip(vals, -max_rotation, max_rotation)
    low = np.full(view.n_api, -max_rotation, dtype=np.float64)
    high = -low
  else:
    low, high = view.api_lower.copy(), view.api_upper.copy()
    if degrees:
      low, high = np.degrees(low), np.degrees(high)
  return vals, low, high


def normalize_obs(values, low, high):
  """ObservationFactory.get_obs normalisation (obs.py:148-152); gym Box stores
  low/high as float32."""
  low = low.astype(np.float32)
  high = high.astype(np.float32)
  return ((2 * (values - low)) / (high - low)) - 1


# ---- rewards (gym_solo/core/rewards.py) ----------------------------------------

def gaussian(x, bounds=(0., 0.), margin=0., margin_value=.1):
  """rewards.py:381-431 (tolerance -> gaussian), batched."""
  lower, upper = bounds
  if lower > upper:
    raise ValueError('Lower bound is greater than upper bound')
  if margin < 0:
    raise ValueError('Margin must be non-negative')
  if not 0 < margin_value <= 1:
    raise ValueError('Margin value must be valued in (0, 1]')
  x = np.asarray(x, dtype=np.float64)
  within = np.logical_and(lower <= x, x <= upper)
  if margin == 0:
    return np.where(within, 1., 0.)
  scale = np.sqrt(-2 * np.log(margin_value))
  sigmas = np.where(x < lower, lower - x, x - upper) / margin
  return np.where(within, 1., np.exp(-0.5 * (sigmas * scale) ** 2))


def linear(x, target, span, symmetric=False):
  """rewards.py:434-461 (scalar)."""
  if span == 0:
    return 1. if x == target else 0.
  x_delta = x - target
  if abs(x_delta) > abs(span):
    return 0.
  ratio = x_delta / span
  if not symmetric and ratio < 0:
    return 0
  return 1 - abs(ratio)


def upright_reward(view):
  """UprightReward.compute (rewards.py:221-234)."""
  fully_upright = -np.pi / 2
  y = euler_from_quat(view.quat)[:, 1]
  return fully_upright * y / fully_upright ** 2


def flat_torso_reward(view, hard_margin=.1, soft_margin=.1):
  """FlatTorsoReward.compute (rewards.py:254-269)."""
  e = euler_from_quat(view.quat)
  rmse = np.sqrt(e[:, 0] ** 2 + e[:, 1] ** 2)
  return gaussian(rmse, bounds=(-hard_margin, hard_margin), margin=soft_margin)


def small_control_reward(view, margin=1.):
  """SmallControlReward.compute (rewards.py:289-301): mean |qdot| over ALL API
  joints, fixed ones included."""
  return gaussian(np.mean(np.abs(view.joint_vel()), axis=1), margin=margin)


def horizontal_move_speed_reward(view, target_speed, hard_margin=.1, soft_margin=.1):
  """HorizontalMoveSpeedReward.compute (rewards.py:326-338)."""
  speed = np.sqrt(view.lin[:, 0] ** 2 + view.lin[:, 1] ** 2)
  return gaussian(speed, bounds=(target_speed - hard_margin, target_speed + hard_margin),
                  margin=soft_margin)


def torso_height_reward(view, target_height, hard_margin=.1, soft_margin=.1):
  """TorsoHeightReward.compute (rewards.py:362-373)."""
  return gaussian(view.pos[:, 2], bounds=(target_height - hard_margin, target_height + hard_margin),
                  margin=soft_margin)


def home_position_reward(view):
  """The hand-built "home_pos" composite of
  examples/solo8_vanilla/interactive_pos_control.py:22-35 (BASELINE.json's
  HomePositionReward)."""
  flat = flat_torso_reward(view, hard_margin=.1, soft_margin=np.pi)
  height = torso_height_reward(view, 0.33698, 0.025, 0.15)
  small = small_control_reward(view, margin=10)
  still = horizontal_move_speed_reward(view, 0, hard_margin=.5, soft_margin=3)
  stand = 0.5 * flat + 0.5 * height
  return 1 * (stand * small * still)


# ---- terminations (gym_solo/core/termination.py) --------------------------------

class TimeBased:
  """TimeBasedTermination (termination.py:59-83), batched counters."""

  def __init__(self, max_step_delta, n):
    self.max_step_delta = max_step_delta
    self.step_delta = np.zeros(n, dtype=np.int64)

  def reset(self, mask=None):
    if mask is None:
      self.step_delta[:] = 0
    else:
      self.step_delta[mask] = 0

  def is_terminated(self, active=None):
    if active is None:
      self.step_delta += 1
    else:
      self.step_delta[active] += 1
    return self.step_delta > self.max_step_delta
