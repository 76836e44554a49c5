"""Observations and the observation factory -- batched drop-in for
gym_solo/core/obs.py (Observation :16-92, ObservationFactory :95-190, TorsoIMU
:193-282, MotorEncoder :285-363).

Same classes, constructor arguments, labels, spaces and ValueErrors; values are
[num_envs, d] float32 tensors on the GPU.  The built-in observations are not
evaluated in Python on the step path: they compile to SoloObsTerm descriptors
(include/solo8_b200.h) that the fused step kernel evaluates.  `compute()` on a
built-in launches the observe kernel with a one-term program.  User subclasses
run as ordinary torch code against the BatchedClient getters.
"""
from abc import ABC, abstractmethod
from typing import List, Tuple

import numpy as np
import torch

from gym_solo_b200 import _abi
from gym_solo_b200 import solo_types
from gym_solo_b200 import spaces
from gym_solo_b200.core import _epoch


class Observation(ABC):
  """An observation for a body in the simulation (obs.py:16-92)."""
  _client = None

  @abstractmethod
  def __init__(self, body_id: int):
    pass

  @property
  @abstractmethod
  def observation_space(self) -> spaces.Space:
    pass

  @property
  @abstractmethod
  def labels(self) -> List[str]:
    pass

  @abstractmethod
  def compute(self) -> solo_types.obs:
    pass

  @property
  def client(self):
    if not self._client:
      raise ValueError('PyBullet client needs to be set')
    return self._client

  @client.setter
  def client(self, client):
    self._client = client

  # Built-ins override this and return a SoloObsTerm; None = evaluate in Python.
  def _fused_term(self, normalize: bool):
    return None


def _fill(arr, values):
  for i, v in enumerate(values):
    arr[i] = float(v)


def _run_single_term(obs, term, dim):
  """Evaluate one descriptor on the current state with the observe kernel."""
  client = obs.client
  prog = _abi.SoloProgram()
  prog.n_obs = 1
  prog.obs[0] = term
  client.set_program(prog)
  out = torch.empty((client.n, dim), dtype=torch.float32, device=client.device)
  client.observe(out)
  return out


class ObservationFactory:
  def __init__(self, client, normalize: bool = False):
    self._client = client
    self._observations = []
    self._obs_space = None
    self._normalize = normalize

  def register_observation(self, obs: Observation):
    """obs.py:109-128, with `.size` read as the per-env width."""
    obs.client = self._client

    lbl_len = len(obs.labels)
    obs_space_len = len(obs.observation_space.low)
    value = obs.compute()
    obs_len = value.shape[-1] if hasattr(value, 'shape') and len(value.shape) else 1

    if lbl_len != obs_space_len:
      raise ValueError('Labels have length {} != obs space len {}'.format(
        lbl_len, obs_space_len))
    if lbl_len != obs_len:
      raise ValueError('Labels have length {} != obs len {}'.format(
        lbl_len, obs_len))

    self._observations.append(obs)
    _epoch.bump()

  def get_obs(self) -> Tuple[solo_types.obs, List[str]]:
    """obs.py:130-159 (the non-fused path: one kernel / torch expression per
    registered observation)."""
    if not self._observations:
      raise ValueError('Need to register at least one observation instance')

    all_obs = []
    all_labels = []
    for obs in self._observations:
      all_labels.append(obs.labels)
      values = obs.compute()
      if self._normalize:
        a = torch.as_tensor(values)
        low = torch.as_tensor(obs.observation_space.low, device=a.device)
        hi = torch.as_tensor(obs.observation_space.high, device=a.device)
        values = ((2 * (a - low)) / (hi - low)) - 1
      all_obs.append(values)

    observations = torch.cat([torch.as_tensor(v).float() for v in all_obs], dim=-1)
    labels = [l for lbl in all_labels for l in lbl]
    return observations, labels

  def get_observation_space(self, generate=False) -> spaces.Box:
    """obs.py:161-190; cached unless `generate`."""
    if not self._observations:
      raise ValueError('Can\'t generate an empty observation space')
    if self._obs_space and not generate:
      return self._obs_space

    lower, upper = [], []
    for obs in self._observations:
      lower.extend(obs.observation_space.low)
      upper.extend(obs.observation_space.high)

    if self._normalize:
      self._obs_space = spaces.Box(low=-1, high=1, shape=(len(lower),))
    else:
      self._obs_space = spaces.Box(low=np.array(lower), high=np.array(upper))
    return self._obs_space

  # ---- fused path ---------------------------------------------------------------
  def labels(self) -> List[str]:
    return [l for obs in self._observations for l in obs.labels]

  def fused_terms(self):
    """SoloObsTerm list for the step kernel, or None if any registered
    observation is a user subclass (then the env evaluates get_obs())."""
    if not self._observations or len(self._observations) > _abi.MAX_OBS_TERMS:
      return None
    terms = [o._fused_term(self._normalize) for o in self._observations]
    return None if any(t is None for t in terms) else terms



class TorsoIMU(_epoch.Tracked, Observation):
  """Orientation and velocities of the torso (obs.py:193-282)."""
  labels: List[str] = ['θx', 'θy', 'θz', 'vx', 'vy', 'vz', 'wx', 'wy', 'wz']

  def __init__(self, body_id: int, degrees: bool = False,
               max_lin_velocity: float = 15,
               max_angular_velocity: float = 10.):
    self.robot = body_id
    self._degrees = degrees
    self._max_lin = max_lin_velocity
    self._max_ang = max_angular_velocity

    self._low = None
    self._high = None
    self.observation_space  # Populate the bounds (obs.py:228)

  @property
  def observation_space(self) -> spaces.Box:
    angle_min = -180. if self._degrees else -np.pi
    angle_max = 180. if self._degrees else np.pi

    lower = [angle_min, angle_min, angle_min,
             -self._max_lin, -self._max_lin, -self._max_lin,
             -self._max_ang, -self._max_ang, -self._max_ang]
    upper = [angle_max, angle_max, angle_max,
             self._max_lin, self._max_lin, self._max_lin,
             self._max_ang, self._max_ang, self._max_ang]

    # clip bounds are cached on first access (obs.py:254-256)
    if not (self._low and self._high):
      self._low = lower
      self._high = upper

    return spaces.Box(low=np.array(lower), high=np.array(upper))

  def _fused_term(self, normalize: bool):
    t = _abi.SoloObsTerm()
    t.kind = _abi.OBS_TORSO_IMU
    t.degrees = 1 if self._degrees else 0
    t.clip = 1
    t.normalize = 1 if normalize else 0
    _fill(t.clip_lo, self._low)
    _fill(t.clip_hi, self._high)
    space = self.observation_space
    _fill(t.norm_lo, space.low)
    _fill(t.norm_hi, space.high)
    return t

  def compute(self) -> solo_types.obs:
    """obs.py:260-282, on the GPU: [num_envs, 9]."""
    return _run_single_term(self, self._fused_term(False), 9)


class MotorEncoder(_epoch.Tracked, Observation):
  """Position of all the joints, fixed ones included (obs.py:285-363)."""

  def __init__(self, body_id: int, degrees: bool = False,
               max_rotation: float = None):
    self.robot = body_id
    self._degrees = degrees
    self._max_rot = max_rotation

  @property
  def _num_joints(self):
    return self.client.getNumJoints(self.robot)

  @property
  def observation_space(self) -> spaces.Space:
    if self._max_rot:
      return spaces.Box(low=-self._max_rot, high=self._max_rot,
                        shape=(self._num_joints, ))

    lower, upper = [], []
    for joint in range(self._num_joints):
      joint_info = self.client.getJointInfo(self.robot, joint)
      lower.append(joint_info[8])
      upper.append(joint_info[9])

    lower = np.array(lower)
    upper = np.array(upper)

    if self._degrees:
      lower = np.degrees(lower)
      upper = np.degrees(upper)

    return spaces.Box(low=lower, high=upper)

  @property
  def labels(self) -> List[str]:
    return [self.client.getJointInfo(self.robot, joint)[1].decode('UTF-8')
            for joint in range(self._num_joints)]

  def _fused_term(self, normalize: bool):
    t = _abi.SoloObsTerm()
    t.kind = _abi.OBS_MOTOR_ENCODER
    t.degrees = 1 if self._degrees else 0
    t.clip = 1 if self._max_rot else 0
    t.normalize = 1 if normalize else 0
    n = self._num_joints
    if self._max_rot:
      _fill(t.clip_lo, [-self._max_rot] * n)
      _fill(t.clip_hi, [self._max_rot] * n)
    space = self.observation_space
    _fill(t.norm_lo, space.low)
    _fill(t.norm_hi, space.high)
    return t

  def compute(self) -> solo_types.obs:
    """obs.py:347-363, on the GPU: [num_envs, num_joints]."""
    return _run_single_term(self, self._fused_term(False), self._num_joints)
