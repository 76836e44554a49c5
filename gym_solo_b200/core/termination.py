"""Terminations and the termination factory -- batched drop-in for
gym_solo/core/termination.py (Termination :4-16, TerminationFactory :19-56,
TimeBasedTermination :59-83, PerpetualTermination :86-98).

`is_terminated()` returns a [num_envs] bool tensor (a plain bool before the
termination is attached to an env).  On the fused step path the built-ins are
evaluated in the kernel from int32 per-env counters owned by the client.
"""
from abc import ABC, abstractmethod

import torch

from gym_solo_b200 import _abi
from gym_solo_b200.core import _epoch


class Termination(ABC):
  @abstractmethod
  def reset(self):
    """Resets the state of the termination condition"""
    pass

  @abstractmethod
  def is_terminated(self) -> bool:
    pass

  # Built-ins return a SoloTermTerm; None = evaluate in Python.
  def _fused_term(self):
    return None

  def _attach(self, client, slot):
    pass


class TerminationFactory:
  def __init__(self):
    self._terminations = []
    self._use_or = True
    self._client = None

  def register_termination(self, *terminations):
    self._terminations.extend(terminations)
    _epoch.bump()
    if self._client is not None:
      self._attach_all()

  def _set_client(self, client):
    self._client = client
    self._attach_all()

  def _attach_all(self):
    for slot, t in enumerate(self._terminations):
      if slot < _abi.MAX_TERM_TERMS:
        t._attach(self._client, slot)

  def is_terminated(self):
    """termination.py:38-50: OR with short circuit -- per environment: a
    TimeBasedTermination after one that fired is not ticked for that env."""
    if not self._terminations:
      raise ValueError('Need to register at least one termination instance')

    done = None
    for termination in self._terminations:
      if isinstance(termination, TimeBasedTermination) and done is not None and torch.is_tensor(done):
        r = termination.is_terminated(active=~done)
      else:
        r = termination.is_terminated()
      if done is None:
        done = r
      else:
        done = done | r if (torch.is_tensor(done) or torch.is_tensor(r)) else (done or r)
      if not torch.is_tensor(done) and done:
        return True
    return done

  def reset(self, mask=None):
    """termination.py:52-56; `mask` ([num_envs] bool) is batched-only."""
    for termination in self._terminations:
      if mask is not None and isinstance(termination, TimeBasedTermination):
        termination.reset(mask)
      else:
        termination.reset()

  def fused_terms(self):
    if not self._terminations or len(self._terminations) > _abi.MAX_TERM_TERMS:
      return None
    terms = [t._fused_term() for t in self._terminations]
    return None if any(t is None for t in terms) else terms



class TimeBasedTermination(_epoch.Tracked, Termination):
  """Terminates once step_delta exceeds max_step_delta, i.e. on call
  max_step_delta + 1 (termination.py:59-83)."""

  def __init__(self, max_step_delta: int):
    self.max_step_delta = max_step_delta
    self._client = None
    self._slot = None
    self._count = 0
    self.reset()

  def _attach(self, client, slot):
    if self._client is not client or self._slot != slot:
      self._client, self._slot = client, slot
      client.counters[slot].fill_(int(self._count) if not torch.is_tensor(self._count) else 0)

  @property
  def step_delta(self):
    """int before the termination is attached to an env, then the [num_envs]
    int32 device counter the step kernel ticks."""
    if self._client is not None:
      return self._client.counters[self._slot]
    return self._count

  @step_delta.setter
  def step_delta(self, v):
    if self._client is not None:
      self._client.counters[self._slot].copy_(torch.as_tensor(v, device=self._client.device).expand(self._client.n))
    else:
      self._count = v

  def reset(self, mask=None):
    if self._client is not None:
      if mask is None:
        self._client.counters[self._slot].zero_()
      else:
        self._client.counters[self._slot].masked_fill_(torch.as_tensor(mask, device=self._client.device).bool(), 0)
    else:
      self._count = 0

  def is_terminated(self, active=None):
    if self._client is not None:
      c = self._client.counters[self._slot]
      if active is None:
        c += 1
      else:
        c += active.to(c.dtype)
      return c > self.max_step_delta
    self._count += 1
    return self._count > self.max_step_delta

  def _fused_term(self):
    t = _abi.SoloTermTerm()
    t.kind, t.max_step_delta = _abi.TERM_TIME, int(self.max_step_delta)
    return t


class PerpetualTermination(_epoch.Tracked, Termination):
  """Never ends (termination.py:86-98)."""

  def __init__(self):
    self._client = None

  def _attach(self, client, slot):
    self._client = client

  def reset(self):
    pass

  def is_terminated(self):
    if self._client is not None:
      return torch.zeros((self._client.n,), dtype=torch.bool, device=self._client.device)
    return False

  def _fused_term(self):
    t = _abi.SoloTermTerm()
    t.kind, t.max_step_delta = _abi.TERM_PERPETUAL, 0
    return t
