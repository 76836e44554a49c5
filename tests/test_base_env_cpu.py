"""gym_solo/envs/test_solo8_base_env.py ported: the abstract base environment's constructor
contract, on the CPU (the client is whatever `_create_client` returns, here a stand-in)."""
import importlib
import random

import numpy as np
import pytest
import torch

from gym_solo_b200 import _abi, spaces, testing
from gym_solo_b200.core import configs, obs, rewards, termination as terms
from gym_solo_b200.envs.solo8_base_env import Solo8BaseEnv


class FakeClient:
  def __init__(self, n=2):
    self.n = n
    self.device = torch.device('cpu')
    self.counters = torch.zeros((_abi.MAX_TERM_TERMS, n), dtype=torch.int32)
    self.disconnected = False

  def disconnect(self):
    self.disconnected = True


class SimpleSoloEnv(Solo8BaseEnv):
  """test_solo8_base_env.py:19-39."""
  action_space = spaces.Box(3, 4, shape=(3, 4))

  def __init__(self, *args, **kwargs):
    self.reset_call = None
    self.client_call = 0
    self.load_bodies_call = None
    self.created = 0
    super().__init__(*args, **kwargs)

  def _create_client(self):
    self.created += 1
    return FakeClient(self.num_envs)

  def reset(self, init_call: bool):
    self.reset_call = init_call

  def load_bodies(self):
    self.load_bodies_call = True

  def step(self):
    pass

  def client_configuration(self):
    self.client_call += 1


def test_abstract_init():
  # test_solo8_base_env.py:50-52
  with pytest.raises(TypeError):
    Solo8BaseEnv()


def test_init():
  # test_solo8_base_env.py:54-80
  config = configs.Solo8BaseConfig()
  env = SimpleSoloEnv(config, False)
  assert env.config == config
  assert env.created == 1 and isinstance(env.client, FakeClient)
  assert isinstance(env.obs_factory, obs.ObservationFactory)
  assert isinstance(env.reward_factory, rewards.RewardFactory)
  assert isinstance(env.termination_factory, terms.TerminationFactory)
  assert env.load_bodies_call
  assert env.reset_call            # reset(init_call=True) once, at the end of __init__
  env._close()
  assert env.client.disconnected


def test_gui_and_realtime_are_not_provided():
  """test_solo8_base_env.py:82-91, 137-147: p.GUI and setRealTimeSimulation(1) need a display /
  wall-clock stepping: unavailable on this box, refused loudly instead of ignored."""
  with pytest.raises(NotImplementedError):
    SimpleSoloEnv(configs.Solo8BaseConfig(), True)
  with pytest.raises(NotImplementedError):
    SimpleSoloEnv(configs.Solo8BaseConfig(dt=None), False)


def test_seed():
  # test_solo8_base_env.py:93-117
  env = SimpleSoloEnv(configs.Solo8BaseConfig(), False)
  seed = 69
  env._seed(seed)
  numpy_control = float(np.random.rand(1)[0])
  random_control = random.random()
  assert numpy_control != float(np.random.rand(1)[0])
  assert random_control != random.random()
  env.seed(seed)                    # the new-gym alias
  assert numpy_control == float(np.random.rand(1)[0])
  assert random_control == random.random()


def test_obs_space():
  # test_solo8_base_env.py:119-122
  env = SimpleSoloEnv(configs.Solo8BaseConfig(), False)
  env.obs_factory.register_observation(testing.CompliantObs(None))
  a, b = env.observation_space, env.obs_factory.get_observation_space()
  assert a is b
  np.testing.assert_array_equal(a.low, [0., 0.])
  np.testing.assert_array_equal(a.high, [3., 3.])


def test_client_configuration_hook_runs_once():
  # test_solo8_base_env.py:149-151
  env = SimpleSoloEnv(configs.Solo8BaseConfig(), False)
  assert env.client_call == 1


def test_config_defaults_are_the_references():
  # gym_solo/core/configs.py:8-38; test_solo8_base_env.py:125-135 (dt is what the engine steps with)
  c = configs.Solo8BaseConfig()
  assert c.dt == 1e-3 and c.motor_torque_limit == 2 and c.gravity == (0., 0., -9.81)
  assert c.robot_start_pos == (0., 0., 0.5) and c.robot_start_orientation_euler == (0., 0., 0.)
  assert c.max_motor_rotation == 2 * np.pi and c.linear_damping == .04 and c.angular_damping == .04
  assert c.restitution == 0. and c.lateral_friction == 0.5
  assert (c.render_width, c.render_height, c.render_fov) == (369, 369, 80)
  assert configs.Solo8BaseConfig(dt=20).dt == 20
