"""Rewards and the reward factory -- batched drop-in for
gym_solo/core/rewards.py (Reward :15-55, RewardFactory :64-118, AdditiveReward
:121-160, MultiplicitiveReward :163-199, UprightReward :202-234,
FlatTorsoReward :237-269, SmallControlReward :272-301,
HorizontalMoveSpeedReward :304-338, TorsoHeightReward :341-373,
tolerance/gaussian :376-431, linear :434-461).

A reward value is a [num_envs] float32 tensor.  Trees made of the built-in
classes compile to a postfix SoloRewardOp program (include/solo8_b200.h) that
the fused step kernel evaluates; a tree containing a user subclass is evaluated
here with torch, the user's compute() reading the BatchedClient getters.
"""
from abc import ABC, abstractmethod
from dataclasses import dataclass
from typing import Tuple, List
import functools
import math
import warnings

import numpy as np
import torch

from gym_solo_b200 import _abi
from gym_solo_b200 import solo_types
from gym_solo_b200.core import _epoch


class Reward(ABC):
  """A reward for a body in the simulation (rewards.py:15-55)."""
  _client = None

  @abstractmethod
  def compute(self) -> solo_types.reward:
    pass

  @property
  def client(self):
    if self._client is None:
      raise ValueError('PyBullet client needs to be set')
    return self._client

  @client.setter
  def client(self, client):
    self._client = client

  def _emit(self, ops: list) -> bool:
    """Append this reward's postfix ops; False = not fusable (user subclass)."""
    return False


@dataclass
class _WeightedReward:
  reward: Reward
  weight: float


def _op(code, n=0, p=(), coef=()):
  o = _abi.SoloRewardOp()
  o.op, o.n = code, n
  for i, v in enumerate(p):
    o.p[i] = float(v)
  for i, v in enumerate(coef):
    o.coef[i] = float(v)
  return o


def _weighted_program(terms: List[_WeightedReward], ops: list) -> bool:
  # emit in chunks of at most REWARD_STACK leaves, each chunk reduced by one ADD
  # carrying its weights; chunk results are then summed with unit weights
  chunks = [terms[i:i + _abi.REWARD_STACK] for i in range(0, len(terms), _abi.REWARD_STACK)]
  if len(chunks) > _abi.REWARD_STACK:
    return False
  for ch in chunks:
    for wr in ch:
      if not wr.reward._emit(ops):
        return False
    ops.append(_op(_abi.REW_ADD, len(ch), coef=[wr.weight for wr in ch]))
  if len(chunks) > 1:
    ops.append(_op(_abi.REW_ADD, len(chunks), coef=[1.0] * len(chunks)))
  return True


def _run_program(reward: Reward, ops: list):
  """Evaluate a postfix program on the current state with the reward kernel."""
  client = reward.client
  prog = _abi.SoloProgram()
  if len(ops) > _abi.MAX_REWARD_OPS:
    raise ValueError('reward tree too large for the fused program')
  prog.n_rew = len(ops)
  for i, o in enumerate(ops):
    prog.rew[i] = o
  client.set_program(prog)
  out = torch.empty((client.n,), dtype=torch.float32, device=client.device)
  client.reward(out)
  return out


class RewardFactory:
  """Linear combination of the registered rewards (rewards.py:64-118)."""

  def __init__(self, client):
    self._client = client
    self._rewards: List[_WeightedReward] = []

  def register_reward(self, weight: float, reward: Reward):
    # the reference wraps this in @deprecation.deprecated (rewards.py:87-90)
    warnings.warn('register_reward is deprecated. This will probaby be removed for '
                  'AdditiveReward down the line. For best future-proofing pratices, create '
                  'an AdditiveReward and register it with weight 1',
                  DeprecationWarning, stacklevel=2)
    reward.client = self._client
    self._rewards.append(_WeightedReward(reward=reward, weight=weight))
    _epoch.bump()

  def get_reward(self):
    if not self._rewards:
      raise ValueError('Need to register at least one reward instance')
    return sum(wr.weight * wr.reward.compute() for wr in self._rewards)

  def fused_ops(self):
    """Postfix program for the step kernel, or None (user subclass in the tree,
    or the tree exceeds the kernel's table)."""
    if not self._rewards:
      return None
    ops = []
    if not _weighted_program(self._rewards, ops) or len(ops) > _abi.MAX_REWARD_OPS:
      return None
    return ops



class AdditiveReward(_epoch.Tracked, Reward):
  """c1 r1 + c2 r2 + ... (rewards.py:121-160)."""

  def __init__(self):
    self._terms: List[_WeightedReward] = []

  def add_term(self, coefficient: float, reward: Reward):
    # reads self.client first, like the reference: raises ValueError unless the
    # caller set it (rewards.py:143)
    reward.client = self.client
    self._terms.append(_WeightedReward(reward=reward, weight=coefficient))
    _epoch.bump()

  def compute(self):
    if not self._terms:
      raise ValueError('Need to register at least one term')
    return sum(wr.weight * wr.reward.compute() for wr in self._terms)

  def _emit(self, ops):
    if not self._terms:
      raise ValueError('Need to register at least one term')
    return _weighted_program(self._terms, ops)


class MultiplicitiveReward(_epoch.Tracked, Reward):
  """c * r1 * r2 * ... (rewards.py:163-199)."""

  def __init__(self, coefficient: float, *terms: Reward):
    self._coeff = coefficient
    self._terms = terms

  def compute(self):
    if not self._terms:
      raise ValueError('Need to register at least one term')
    return self._coeff * functools.reduce(lambda a, b: a * b,
                                          [t.compute() for t in self._terms])

  def _emit(self, ops):
    if not self._terms:
      raise ValueError('Need to register at least one term')
    if len(self._terms) > _abi.REWARD_STACK:
      return False
    for t in self._terms:
      if not t._emit(ops):
        return False
    ops.append(_op(_abi.REW_MUL, len(self._terms), p=[self._coeff]))
    return True

  @property
  def client(self):
    return self._client

  @client.setter
  def client(self, client):
    self._client = client
    for t in self._terms:
      t.client = self._client


class _Leaf(_epoch.Tracked, Reward):
  """Built-in state reward: one op of the fused program."""

  def _leaf_op(self):
    raise NotImplementedError

  def _emit(self, ops):
    ops.append(self._leaf_op())
    return True

  def compute(self):
    return _run_program(self, [self._leaf_op()])


class UprightReward(_Leaf):
  """-2 pitch / pi in [-1, 1] (rewards.py:202-234)."""
  _fully_upright = -np.pi / 2

  def __init__(self, robot_id: int):
    self._robot_id = robot_id

  def _leaf_op(self):
    return _op(_abi.REW_UPRIGHT)


class FlatTorsoReward(_Leaf):
  """rewards.py:237-269."""

  def __init__(self, robot_id: int, hard_margin: float = .1,
               soft_margin: float = 0.1):
    self._robot_id = robot_id
    self._hard_margin = hard_margin
    self._soft_margin = soft_margin

  def _leaf_op(self):
    _check_gaussian((-self._hard_margin, self._hard_margin), self._soft_margin, .1)
    return _op(_abi.REW_FLAT_TORSO, p=[self._hard_margin, self._soft_margin])


class SmallControlReward(_Leaf):
  """rewards.py:272-301: mean |qdot| over ALL joints, fixed ones included."""

  def __init__(self, robot_id: int, margin: float = 1.):
    self._robot_id = robot_id
    self._margin = margin

  def _leaf_op(self):
    _check_gaussian((0., 0.), self._margin, .1)
    return _op(_abi.REW_SMALL_CONTROL, p=[0., self._margin])


class HorizontalMoveSpeedReward(_Leaf):
  """rewards.py:304-338."""

  def __init__(self, robot_id: int, target_speed: int, hard_margin: float = .1,
               soft_margin: float = 0.1):
    self._robot_id = robot_id
    self._target_speed = target_speed
    self._hard_margin = hard_margin
    self._soft_margin = soft_margin

  def _leaf_op(self):
    b = (self._target_speed - self._hard_margin, self._target_speed + self._hard_margin)
    _check_gaussian(b, self._soft_margin, .1)
    return _op(_abi.REW_HSPEED, p=[b[0], b[1], self._soft_margin])


class TorsoHeightReward(_Leaf):
  """rewards.py:341-373."""

  def __init__(self, robot_id: int, target_height: int, hard_margin: float = .1,
               soft_margin: float = 0.1):
    self._robot_id = robot_id
    self._target_height = target_height
    self._hard_margin = hard_margin
    self._soft_margin = soft_margin

  def _leaf_op(self):
    b = (self._target_height - self._hard_margin, self._target_height + self._hard_margin)
    _check_gaussian(b, self._soft_margin, .1)
    return _op(_abi.REW_HEIGHT, p=[b[0], b[1], self._soft_margin])


class HomePositionReward(MultiplicitiveReward):
  """BASELINE.json's "HomePositionReward".  The reference snapshot has no such
  class; this is the hand-built composite of
  examples/solo8_vanilla/interactive_pos_control.py:22-35 with its constants:
    1 * (0.5 Flat(.1, pi) + 0.5 Height(0.33698, .025, .15))
      * SmallControl(margin 10) * HorizontalSpeed(0, .5, 3)."""

  def __init__(self, robot_id: int, client=None):
    flat = FlatTorsoReward(robot_id, hard_margin=.1, soft_margin=np.pi)
    height = TorsoHeightReward(robot_id, 0.33698, 0.025, 0.15)
    small_control = SmallControlReward(robot_id, margin=10)
    no_move = HorizontalMoveSpeedReward(robot_id, 0, hard_margin=.5, soft_margin=3)
    stand = AdditiveReward()
    stand._terms = [_WeightedReward(flat, 0.5), _WeightedReward(height, 0.5)]
    self._stand = stand
    super().__init__(1, stand, small_control, no_move)
    if client is not None:
      self.client = client

  @MultiplicitiveReward.client.setter
  def client(self, client):
    self._client = client
    for t in self._terms:
      t.client = client
    for wr in self._stand._terms:
      wr.reward.client = client


def _check_gaussian(bounds, margin, margin_value):
  """The argument checks of gaussian() (rewards.py:406-417)."""
  lower, upper = bounds
  if lower > upper:
    raise ValueError('Lower bound ({}) is greater than upper bound ({})'.format(
                     lower, upper))
  if margin < 0:
    raise ValueError('Margin must be non-negative: {}'.format(margin))
  if not 0 < margin_value <= 1:
    raise ValueError('Margin value must be valued in (0, 1]: {}'.format(
      margin_value))


def tolerance(*args, **kwargs):
  """rewards.py:376-381."""
  return gaussian(*args, **kwargs)


def gaussian(x, bounds: Tuple[float, float] = (0., 0.),
             margin: float = 0., margin_value: float = .1):
  """rewards.py:384-431 for Python floats, numpy arrays and torch tensors (for
  user rewards; the built-ins evaluate it in the kernel, solo_core.h
  gaussian_tol)."""
  _check_gaussian(bounds, margin, margin_value)
  lower, upper = bounds
  if torch.is_tensor(x):
    within = (lower <= x) & (x <= upper)
    one = torch.ones_like(x)
    if margin == 0:
      return torch.where(within, one, torch.zeros_like(x))
    scale = math.sqrt(-2 * math.log(margin_value))
    sigmas = torch.where(x < lower, lower - x, x - upper) / margin
    return torch.where(within, one, torch.exp(-0.5 * (sigmas * scale) ** 2))

  within_bounds = np.logical_and(lower <= x, x <= upper)
  if margin == 0:
    value = np.where(within_bounds, 1., 0.)
  else:
    scale = np.sqrt(-2 * np.log(margin_value))
    sigmas = np.where(x < lower, lower - x, x - upper) / margin
    values = np.exp(-0.5 * (sigmas * scale) ** 2)
    value = np.where(within_bounds, 1., values)
  return float(value) if np.isscalar(x) else value


def linear(x: float, target: float, span: float, symmetric=False) -> float:
  """rewards.py:434-461."""
  if span == 0:
    return 1. if x == target else 0.
  x_delta = x - target
  if abs(x_delta) > abs(span):
    return 0.
  ratio = x_delta / span
  if not symmetric and ratio < 0:
    return 0
  return 1 - abs(ratio)
