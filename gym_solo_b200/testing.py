"""Test doubles (gym_solo/testing.py:11-80), batched where they touch an env."""
import numpy as np
import torch

from gym_solo_b200.core import obs
from gym_solo_b200.core import rewards
from gym_solo_b200.core import termination
from gym_solo_b200 import spaces


class CompliantObs(obs.Observation):
  """Always observes [1, 2] (one row per environment once a client is set)."""
  observation_space = spaces.Box(low=np.array([0., 0.]),
                                 high=np.array([3., 3.]))
  labels = ['1', '2']

  def __init__(self, body_id: int):
    pass

  def compute(self):
    n = getattr(self._client, 'n', None)
    dev = getattr(self._client, 'device', 'cpu')
    v = torch.tensor([1., 2.], device=dev if n else 'cpu')
    return v.expand(n, 2) if n else v


class SimpleReward(rewards.Reward):
  """A reward which will always return 1."""
  def compute(self):
    return 1


class ReflectiveReward(rewards.Reward):
  """A reward which returns a configurable fixed value."""
  def __init__(self, return_value: float):
    self._return_value = return_value

  def compute(self):
    return self._return_value


class DummyTermination(termination.Termination):
  def __init__(self, body_id: int, termination_var: bool):
    self.body_id = body_id
    self.termination_var = termination_var
    self.reset_counter = 0
    self.reset()

  def reset(self):
    self.reset_counter += 1

  def is_terminated(self) -> bool:
    return self.termination_var
