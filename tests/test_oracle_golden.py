"""Pins the numpy half of the oracle (oracle/solo_oracle.py: observations,
rewards, terminations, action normalisation) against vectors produced by the
REFERENCE's own classes (tools/gen_golden.py -> tests/golden/epilogue_golden.npz)
and against the known answers in the reference's tests."""
import math
import os

import numpy as np
import pytest

from helpers import GOLDEN, model_and_config
from oracle import solo_oracle as so

G = np.load(os.path.join(GOLDEN, 'epilogue_golden.npz'))
TOL = dict(rtol=1e-12, atol=1e-12)


@pytest.fixture(scope='module')
def view():
  _, model, _ = model_and_config()
  return so.StateView(G['states'], model)


OBS_CASES = {
    'imu_rad': lambda v: so.torso_imu(v),
    'imu_deg': lambda v: so.torso_imu(v, degrees=True, max_lin_velocity=5, max_angular_velocity=180),
    'enc_rad': lambda v: so.motor_encoder(v),
    'enc_deg': lambda v: so.motor_encoder(v, degrees=True),
    'enc_clip': lambda v: so.motor_encoder(v, max_rotation=2.5),
    'enc_deg_clip': lambda v: so.motor_encoder(v, degrees=True, max_rotation=90.),
}


@pytest.mark.parametrize('name', sorted(OBS_CASES))
def test_observation_matches_reference(view, name):
  vals, low, high = OBS_CASES[name](view)
  np.testing.assert_allclose(vals, G['obs_' + name], **TOL)
  # gym Box bounds are float32 in the reference
  np.testing.assert_allclose(low.astype(np.float32), G['obs_%s_low' % name], rtol=0, atol=0)
  np.testing.assert_allclose(high.astype(np.float32), G['obs_%s_high' % name], rtol=0, atol=0)


@pytest.mark.parametrize('norm', [False, True])
@pytest.mark.parametrize('combo', [('imu_rad', 'enc_rad'), ('imu_deg', 'enc_clip'), ('enc_deg_clip', 'imu_rad')])
def test_factory_concat_and_normalisation(view, combo, norm):
  parts = []
  for c in combo:
    vals, low, high = OBS_CASES[c](view)
    parts.append(so.normalize_obs(vals, low, high) if norm else vals)
  got = np.concatenate(parts, axis=1)
  key = 'factory_%s_%s_%s' % (combo[0], combo[1], 'norm' if norm else 'raw')
  np.testing.assert_allclose(got, G[key], **TOL)


REW_CASES = {
    'upright': lambda v: so.upright_reward(v),
    'flat_default': lambda v: so.flat_torso_reward(v),
    'flat_wide': lambda v: so.flat_torso_reward(v, .1, np.pi),
    'flat_step': lambda v: so.flat_torso_reward(v, 1.0, 0.),
    'small_default': lambda v: so.small_control_reward(v),
    'small_10': lambda v: so.small_control_reward(v, 10),
    'hspeed_0': lambda v: so.horizontal_move_speed_reward(v, 0, .5, 3),
    'hspeed_2': lambda v: so.horizontal_move_speed_reward(v, 2.0),
    'height_home': lambda v: so.torso_height_reward(v, 0.33698, 0.025, 0.15),
    'height_default': lambda v: so.torso_height_reward(v, 0.3),
    'home_pos': lambda v: so.home_position_reward(v),
    'upright_plus_home': lambda v: so.upright_reward(v) + so.home_position_reward(v),
    'mixed': lambda v: (0.25 * so.upright_reward(v) - 2.0 * so.small_control_reward(v, 10)
                        + 3.0 * (0.5 * so.flat_torso_reward(v) * so.torso_height_reward(v, 0.3))),
}


@pytest.mark.parametrize('name', sorted(REW_CASES))
def test_reward_matches_reference(view, name):
  np.testing.assert_allclose(REW_CASES[name](view), G['rew_' + name], **TOL)


def test_gaussian_and_linear_match_reference():
  xs = G['gauss_x']
  np.testing.assert_allclose(so.gaussian(xs, (-.5, .5), 1.), G['gauss_a'], **TOL)
  np.testing.assert_allclose(so.gaussian(xs, (0., 0.), .3, .5), G['gauss_b'], **TOL)
  np.testing.assert_allclose(so.gaussian(xs, (1., 2.), 0.), G['gauss_c'], **TOL)
  np.testing.assert_allclose([so.linear(float(x), 1., 1.5) for x in xs], G['linear_a'], **TOL)
  np.testing.assert_allclose([so.linear(float(x), 0., -2., symmetric=True) for x in xs], G['linear_b'], **TOL)


def test_gaussian_argument_errors():
  # rewards.py:409-417; test_rewards.py:301-318
  with pytest.raises(ValueError):
    so.gaussian(0., bounds=(1, 0))
  with pytest.raises(ValueError):
    so.gaussian(0., margin=-1)
  with pytest.raises(ValueError):
    so.gaussian(0., margin=1, margin_value=0)
  with pytest.raises(ValueError):
    so.gaussian(0., margin=1, margin_value=1.5)


def test_termination_sequence_matches_reference():
  """TimeBased(5) | Perpetual | TimeBased(3) with short circuit, reset at call 8."""
  a, c = so.TimeBased(5, 1), so.TimeBased(3, 1)
  seq = []
  for i in range(12):
    da = a.is_terminated()
    dc = c.is_terminated(active=~da)     # short circuit: not ticked once `a` fired
    seq.append(bool(da[0] or dc[0]))
    if i == 7:
      a.reset(); c.reset()
  assert seq == list(G['term_seq'])
  assert [int(a.step_delta[0]), int(c.step_delta[0])] == list(G['term_counters_end'])


def test_action_normalisation_scale_is_the_float32_bound():
  """solo8v2vanilla.py:84-85: a * self._action_space.high, a float32 Box bound.
  The scale Solo8VanillaEnv._scale() hands to the kernel is spaces.Box(...).high[0]; the
  reference's outputs (golden) are its inputs times exactly that number.  The kernel side
  (action_scale in setup_legpar_kernel) is tested on the GPU: test_gpu_configs.py."""
  from gym_solo_b200 import spaces
  from gym_solo_b200.core.configs import Solo8BaseConfig
  high = spaces.Box(-Solo8BaseConfig.max_motor_rotation, Solo8BaseConfig.max_motor_rotation, shape=(12,)).high
  assert high.dtype == np.float32 and float(high[0]) == 6.2831854820251465
  np.testing.assert_array_equal(G['norm_actions_in'] * float(high[0]), G['norm_actions_out'])


# ---- known answers quoted from the reference's tests -------------------------------

def test_known_answer_euler_yaw_quarter_turn():
  # test_obs_observations.py:67-88: quat (0,0,.707,.707) -> (0, 0, pi/2), un-normalised
  e = so.euler_from_quat(np.array([0, 0, .707, .707]))
  np.testing.assert_allclose(e, [0, 0, np.pi / 2], atol=1e-12)


def test_known_answer_upright_reward():
  # test_rewards.py:74-92 (exact): pitch +90deg -> -1, -90deg -> +1, level -> 0
  _, model, _ = model_and_config()
  s = np.zeros((3, 29))
  s[0, 3:7] = so.quat_from_euler((0, np.pi / 2, 0))
  s[1, 3:7] = so.quat_from_euler((0, -np.pi / 2, 0))
  s[2, 3:7] = so.quat_from_euler((0, 0, 0))
  r = so.upright_reward(so.StateView(s, model))
  assert r[0] == -1.0 and r[1] == 1.0 and r[2] == 0.0


def test_known_answer_encoder_labels_and_fixed_joints():
  # test_obs_observations.py:237-239 label order; fixed ankles read 0.0
  blob, model, _ = model_and_config()
  assert blob['api_names'] == ['FL_HFE', 'FL_KFE', 'FL_ANKLE', 'FR_HFE', 'FR_KFE', 'FR_ANKLE',
                               'HL_HFE', 'HL_KFE', 'HL_ANKLE', 'HR_HFE', 'HR_KFE', 'HR_ANKLE']
  s = np.arange(29, dtype=np.float64)[None]
  v = so.StateView(s, model).joint_pos()[0]
  assert list(v[[2, 5, 8, 11]]) == [0, 0, 0, 0]
  assert list(v[[0, 1, 3, 4, 6, 7, 9, 10]]) == list(s[0, 13:21])
