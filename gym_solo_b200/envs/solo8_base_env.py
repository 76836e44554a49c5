"""Solo8BaseEnv -- batched drop-in for gym_solo/envs/solo8_base_env.py:18-161.

Same constructor contract: a client is created, gravity / timestep come from
the config, bodies are loaded, three empty factories are made and reset() is
called once with init_call=True.  `self.client` is a BatchedClient (num_envs
environments on one B200) instead of a BulletClient.
"""
from abc import ABC, abstractmethod
from typing import Any, Dict, List, Tuple
import random

import numpy as np

from gym_solo_b200.core import termination as terms
from gym_solo_b200.core import configs
from gym_solo_b200.core import obs
from gym_solo_b200.core import rewards
from gym_solo_b200 import solo_types
from gym_solo_b200 import spaces


class Solo8BaseEnv(ABC):
  """Solo 8 abstract base environment."""
  metadata = {'render.modes': ['rgb_array']}

  def __init__(self, config: configs.Solo8BaseConfig, use_gui: bool,
               normalize_observations: bool = False, num_envs: int = 1,
               device='cuda:0'):
    self.config = config
    self.num_envs = int(num_envs)
    self.device = device
    if use_gui:
      # p.GUI needs a graphics stack: unavailable on this box (solo8_base_env.py:34-35)
      raise NotImplementedError('use_gui=True needs the pybullet GUI; only DIRECT-mode '
                                'behaviour is provided')
    if not self.config.dt:
      # setRealTimeSimulation(1) (solo8_base_env.py:42-43) is a wall-clock GUI mode
      raise NotImplementedError('dt=None (real-time simulation) is not provided')

    self.client = self._create_client()
    self.client_configuration()

    self.plane = 0
    self.load_bodies()

    self.obs_factory = obs.ObservationFactory(self.client,
                                              normalize=normalize_observations)
    self.reward_factory = rewards.RewardFactory(self.client)
    self.termination_factory = terms.TerminationFactory()
    self.termination_factory._set_client(self.client)

    self.reset(init_call=True)

  @abstractmethod
  def _create_client(self):
    """BulletClient(connection_mode=DIRECT) + setGravity +
    setPhysicsEngineParameter (solo8_base_env.py:34-43)."""
    pass

  @abstractmethod
  def load_bodies(self):
    pass

  @property
  @abstractmethod
  def action_space(self) -> spaces.Space:
    pass

  @abstractmethod
  def reset(self, init_call: bool = False):
    pass

  @abstractmethod
  def step(self, action) -> Tuple[solo_types.obs, Any, Any, Dict[Any, Any]]:
    pass

  @property
  def observation_space(self) -> spaces.Space:
    return self.obs_factory.get_observation_space()

  def render(self, mode='rgb_array', envs=None):
    """solo8_base_env.py:116-142: the camera of `config` (render_fov, render_pos,
    render_cam_distance, render_yaw / pitch / roll, near 0.01, far 100, Z up)
    through computeProjectionMatrixFOV, computeViewMatrixFromYawPitchRoll and
    getCameraImage.  The image comes from the library's CUDA rasteriser (no
    graphics API): link hulls, flat shading, checkerboard ground plane.

    envs=None renders environment 0 and returns the reference's (h, w, 4) uint8
    array; envs=(begin, count) returns a uint8 device tensor [count, h, w, 4]."""
    if mode != 'rgb_array':
      raise ValueError('only the rgb_array render mode exists (no display on this box)')
    proj_matrix = self.client.computeProjectionMatrixFOV(
        self.config.render_fov, self.config.render_aspect, 0.01, 100)
    view_matrix = self.client.computeViewMatrixFromYawPitchRoll(
        self.config.render_pos, self.config.render_cam_distance,
        self.config.render_yaw, self.config.render_pitch,
        self.config.render_roll, 2)
    begin, count = (0, 1) if envs is None else (int(envs[0]), int(envs[1]))
    w, h, rgb, _, _ = self.client.getCameraImage(
        self.config.render_width, self.config.render_height, view_matrix,
        proj_matrix, env_begin=begin, env_count=count, depth=False, segmentation=False)
    if envs is None:
      # 4 channels for RGBA
      return np.reshape(rgb[0].cpu().numpy(), (h, w, 4))
    return rgb

  def client_configuration(self):
    """Overridable hook (solo8_base_env.py:144-148)."""
    pass

  def _close(self):
    """Soft shutdown the environment."""
    self.client.disconnect()

  def close(self):
    self._close()

  def _seed(self, seed: int) -> None:
    np.random.seed(seed)
    random.seed(seed)

  def seed(self, seed: int) -> None:
    self._seed(seed)
