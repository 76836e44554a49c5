"""Minimal stand-in for gym.spaces.Box (gym is not a dependency of this package).

Keeps what the reference reads: float32 `low` / `high` arrays, `shape`,
`sample()`, `contains()` (gym_solo/core/obs.py:116-117, 150-151, 183-189;
gym_solo/envs/solo8v2vanilla.py:66-70, 170-172).
"""
import numpy as np


class Space:
  pass


class Box(Space):
  def __init__(self, low, high, shape=None, dtype=np.float32):
    self.dtype = np.dtype(dtype)
    if shape is None:
      low = np.asarray(low)
      high = np.asarray(high)
      if low.shape != high.shape:
        raise ValueError('box dimension mismatch')
      shape = low.shape
    self.shape = tuple(shape)
    self.low = np.broadcast_to(np.asarray(low, dtype=self.dtype), self.shape).copy()
    self.high = np.broadcast_to(np.asarray(high, dtype=self.dtype), self.shape).copy()
    self._rng = np.random.RandomState()

  def seed(self, seed=None):
    self._rng = np.random.RandomState(seed)

  def sample(self):
    return self._rng.uniform(self.low, self.high, size=self.shape).astype(self.dtype)

  def contains(self, x):
    x = np.asarray(x)
    return x.shape == self.shape and bool(np.all(x >= self.low) and np.all(x <= self.high))

  def __eq__(self, other):
    return (isinstance(other, Box) and self.shape == other.shape
            and np.allclose(self.low, other.low) and np.allclose(self.high, other.high))

  def __repr__(self):
    return 'Box(%s, %s, %s, %s)' % (self.low.min(), self.high.max(), self.shape, self.dtype)
