"""gym_solo_b200 -- B200-native batched Solo 8 environments with gym_solo's surface.

The reference registers gym ids (gym_solo/__init__.py:3-11); gym is not a
dependency here, so `make` plays that role for the one id on the step path.
"""

__version__ = '0.1'


def make(env_id, **kwargs):
  if env_id != 'solo8vanilla-v0':
    raise ValueError('unknown environment id %r' % (env_id,))
  from gym_solo_b200.envs import Solo8VanillaEnv
  return Solo8VanillaEnv(**kwargs)
