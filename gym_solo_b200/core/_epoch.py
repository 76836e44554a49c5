"""Program epoch: one process-wide counter bumped whenever anything that is baked into a
fused step program may have changed -- a factory registration, AdditiveReward.add_term, or an
attribute of a built-in observation / reward / termination (margins, max_rotation, degrees,
max_step_delta ...).  Solo8VanillaEnv compares one integer per step and recompiles its
descriptor tables when it moved, so the kernel never keeps evaluating a stale program while
the Python get_obs() / get_reward() path sees the new values."""

value = 0


def bump():
  global value
  value += 1


class Tracked:
  """Mixin of the built-in plug-ins: every attribute assignment moves the epoch."""

  def __setattr__(self, name, v):
    object.__setattr__(self, name, v)
    bump()
