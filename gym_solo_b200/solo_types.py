"""Type aliases (gym_solo/solo_types.py:1-11), batched: an observation is a
[num_envs, d] tensor, a reward a [num_envs] tensor."""
import torch

# A state observation
obs = torch.Tensor

# A reward after a step
reward = torch.Tensor

# Return value for "no-op" functions when monkey-patching
no_op = 'NO_OP'
