"""Host logic of the drop-in Python surface, on CPU (no CUDA needed): the
factories, compositions, error behaviour and the compilation of built-ins into
the C-ABI descriptor tables.  Ports of the reference's pure-logic tests:
gym_solo/core/test_obs_factory.py, test_rewards.py, test_termination_factory.py,
test_termination_conditions.py (test doubles from gym_solo/testing.py)."""
import warnings

import numpy as np
import pytest
import torch

from gym_solo_b200 import _abi, spaces, testing
from gym_solo_b200.core import obs, rewards, termination as terms


class FakeClient:
  """Just enough of BatchedClient for the factories: batch size, device and the
  per-env termination counters."""

  def __init__(self, n=3):
    self.n = n
    self.device = torch.device('cpu')
    self.counters = torch.zeros((_abi.MAX_TERM_TERMS, n), dtype=torch.int32)


# ---- observation factory (test_obs_factory.py:19-176) -----------------------------
def test_obs_factory_empty_raises():
  of = obs.ObservationFactory(FakeClient())
  with pytest.raises(ValueError):
    of.get_obs()
  with pytest.raises(ValueError):
    of.get_observation_space()


def test_obs_factory_register_and_concat():
  c = FakeClient(4)
  of = obs.ObservationFactory(c)
  of.register_observation(testing.CompliantObs(None))
  of.register_observation(testing.CompliantObs(None))
  values, labels = of.get_obs()
  assert labels == ['1', '2', '1', '2']
  assert values.shape == (4, 4)
  np.testing.assert_array_equal(values.numpy(), np.tile([1., 2., 1., 2.], (4, 1)))
  sp = of.get_observation_space()
  np.testing.assert_array_equal(sp.low, [0, 0, 0, 0])
  np.testing.assert_array_equal(sp.high, [3, 3, 3, 3])
  # user subclass -> no fused program: the env evaluates get_obs() in Python
  assert of.fused_terms() is None


def test_obs_factory_normalisation_uses_box_bounds():
  # test_obs_factory.py:165-176: 2 (v - low) / (high - low) - 1 with low 0, high 3
  of = obs.ObservationFactory(FakeClient(2), normalize=True)
  of.register_observation(testing.CompliantObs(None))
  values, _ = of.get_obs()
  np.testing.assert_allclose(values.numpy(), np.tile([2 * 1 / 3 - 1, 2 * 2 / 3 - 1], (2, 1)), rtol=1e-6)
  sp = of.get_observation_space()
  np.testing.assert_array_equal(sp.low, [-1, -1])
  np.testing.assert_array_equal(sp.high, [1, 1])


def test_obs_factory_length_mismatches_raise():
  # obs.py:121-126
  class BadLabels(testing.CompliantObs):
    labels = ['1', '2', '3']

  class BadValues(testing.CompliantObs):
    def compute(self):
      return torch.zeros((3, 5))

  of = obs.ObservationFactory(FakeClient())
  with pytest.raises(ValueError):
    of.register_observation(BadLabels(None))
  with pytest.raises(ValueError):
    of.register_observation(BadValues(None))


def test_obs_space_is_cached_until_regenerated():
  # test_obs_factory.py:48-54, 156-163
  of = obs.ObservationFactory(FakeClient())
  of.register_observation(testing.CompliantObs(None))
  first = of.get_observation_space()
  of.register_observation(testing.CompliantObs(None))
  assert of.get_observation_space() is first
  assert len(of.get_observation_space(generate=True).low) == 4


def test_observation_without_client_raises():
  o = testing.CompliantObs(None)
  with pytest.raises(ValueError):
    _ = o.client


def test_builtin_observations_compile_to_descriptors():
  imu = obs.TorsoIMU(1, degrees=True, max_lin_velocity=5, max_angular_velocity=180)
  t = imu._fused_term(False)
  assert t.kind == _abi.OBS_TORSO_IMU and t.degrees == 1 and t.normalize == 0
  assert list(t.clip_hi)[:9] == [180.] * 3 + [5.] * 3 + [180.] * 3
  assert list(t.clip_lo)[:9] == [-180.] * 3 + [-5.] * 3 + [-180.] * 3
  assert imu.labels == ['θx', 'θy', 'θz', 'vx', 'vy', 'vz', 'wx', 'wy', 'wz']
  rad = obs.TorsoIMU(1)._fused_term(True)
  assert rad.normalize == 1
  assert rad.clip_hi[0] == np.float32(np.pi) and rad.clip_hi[3] == 15. and rad.clip_hi[6] == 10.


# ---- rewards (test_rewards.py:14-164) -----------------------------------------------
def _factory(client=None):
  return rewards.RewardFactory(client or FakeClient())


def test_reward_factory_empty_raises():
  with pytest.raises(ValueError):
    _factory().get_reward()


def test_register_reward_warns_deprecated_and_sets_client():
  c = FakeClient()
  rf = _factory(c)
  r = testing.SimpleReward()
  with pytest.warns(DeprecationWarning):
    rf.register_reward(1, r)
  assert r.client is c


@pytest.mark.parametrize('pairs,want', [([(1, 1)], 1), ([(1, 1), (1, 1)], 2), ([(.5, 2), (3, -1)], -2.),
                                         ([(0, 7), (2, .25)], .5)])
def test_reward_factory_linear_combination(pairs, want):
  # test_rewards.py:37-49
  rf = _factory()
  with warnings.catch_warnings():
    warnings.simplefilter('ignore', DeprecationWarning)
    for w, v in pairs:
      rf.register_reward(w, testing.ReflectiveReward(v))
  assert rf.get_reward() == pytest.approx(want)


def test_additive_reward():
  # test_rewards.py:95-129
  r = rewards.AdditiveReward()
  with pytest.raises(ValueError):
    r.compute()
  with pytest.raises(ValueError):      # add_term reads self.client first
    r.add_term(1, testing.SimpleReward())
  c = FakeClient()
  r.client = c
  a, b = testing.ReflectiveReward(2.), testing.ReflectiveReward(-1.)
  r.add_term(.5, a)
  r.add_term(3., b)
  assert a.client is c and b.client is c
  assert r.compute() == pytest.approx(.5 * 2 - 3.)


def test_multiplicitive_reward():
  # test_rewards.py:132-164
  with pytest.raises(ValueError):
    rewards.MultiplicitiveReward(1).compute()
  a, b = testing.ReflectiveReward(2.), testing.ReflectiveReward(-3.)
  r = rewards.MultiplicitiveReward(.5, a, b)
  assert r.client is None               # non-raising getter
  c = FakeClient()
  r.client = c
  assert a.client is c and b.client is c
  assert r.compute() == pytest.approx(.5 * 2 * -3)


def test_builtin_rewards_compile_to_postfix_program():
  c = FakeClient()
  rf = _factory(c)
  home = rewards.HomePositionReward(1)
  with warnings.catch_warnings():
    warnings.simplefilter('ignore', DeprecationWarning)
    rf.register_reward(1, rewards.UprightReward(1))
    rf.register_reward(1, home)
  ops = rf.fused_ops()
  assert ops is not None
  codes = [o.op for o in ops]
  # upright | (flat, height -> add) small hspeed -> mul | -> add (factory sum)
  assert codes == [_abi.REW_UPRIGHT, _abi.REW_FLAT_TORSO, _abi.REW_HEIGHT, _abi.REW_ADD,
                   _abi.REW_SMALL_CONTROL, _abi.REW_HSPEED, _abi.REW_MUL, _abi.REW_ADD]
  add = ops[3]
  assert add.n == 2 and list(add.coef)[:2] == [.5, .5]
  assert ops[2].p[0] == pytest.approx(0.33698 - 0.025) and ops[2].p[1] == pytest.approx(0.33698 + 0.025)
  assert ops[-1].n == 2 and list(ops[-1].coef)[:2] == [1., 1.]
  # a user subclass anywhere in the tree -> Python evaluation
  with warnings.catch_warnings():
    warnings.simplefilter('ignore', DeprecationWarning)
    rf.register_reward(1, testing.SimpleReward())
  assert rf.fused_ops() is None


def test_gaussian_and_linear_known_answers():
  # test_rewards.py:300-354
  assert rewards.gaussian(0., bounds=(-1, 1), margin=1) == 1.0
  assert rewards.gaussian(2., bounds=(0, 1), margin=1., margin_value=.1) == pytest.approx(.1)
  assert rewards.gaussian(1.5, bounds=(0, 1), margin=0.) == 0.0
  for bad in (dict(bounds=(1, 0)), dict(margin=-1), dict(margin=1, margin_value=0),
              dict(margin=1, margin_value=1.5)):
    with pytest.raises(ValueError):
      rewards.gaussian(0., **bad)
  assert rewards.linear(1., 1., 2.) == pytest.approx(1.)
  assert rewards.linear(2., 1., 2.) == pytest.approx(.5)
  assert rewards.linear(0., 1., 2.) == 0 and rewards.linear(0., 1., 2., symmetric=True) == pytest.approx(.5)


# ---- terminations (test_termination_factory.py, test_termination_conditions.py) -----
def test_termination_factory_empty_raises():
  with pytest.raises(ValueError):
    terms.TerminationFactory().is_terminated()


def test_termination_factory_or_and_reset():
  tf = terms.TerminationFactory()
  a, b = testing.DummyTermination(0, False), testing.DummyTermination(0, False)
  tf.register_termination(a, b)
  assert not tf.is_terminated()
  b.termination_var = True
  assert tf.is_terminated()
  before = (a.reset_counter, b.reset_counter)
  tf.reset()
  assert (a.reset_counter, b.reset_counter) == (before[0] + 1, before[1] + 1)


def test_time_based_termination_scalar_semantics():
  # test_termination_conditions.py:4-38: done on call max_step_delta + 1
  t = terms.TimeBasedTermination(3)
  assert [bool(t.is_terminated()) for _ in range(5)] == [False, False, False, True, True]
  t.reset()
  assert t.step_delta == 0 and not t.is_terminated()
  assert not terms.PerpetualTermination().is_terminated()


def test_time_based_termination_batched_counters_and_short_circuit():
  c = FakeClient(3)
  tf = terms.TerminationFactory()
  a, b = terms.TimeBasedTermination(1), terms.TimeBasedTermination(5)
  tf.register_termination(a, terms.PerpetualTermination(), b)
  tf._set_client(c)
  assert not bool(tf.is_terminated().any())
  c.counters[0, 1] = 5                       # env 1 is far past a's limit
  done = tf.is_terminated()
  assert done.tolist() == [True, True, True]   # a: 2 > 1 everywhere
  # short circuit per env: b was ticked on the first call only
  assert c.counters[2].tolist() == [1, 1, 1]
  tf.reset(mask=torch.tensor([True, False, False]))
  assert c.counters[0].tolist() == [0, 6, 2] and c.counters[2].tolist() == [0, 1, 1]
  ft = tf.fused_terms()
  assert [t.kind for t in ft] == [_abi.TERM_TIME, _abi.TERM_PERPETUAL, _abi.TERM_TIME]
  assert ft[0].max_step_delta == 1 and ft[2].max_step_delta == 5


def test_spaces_box_is_float32_like_gym():
  b = spaces.Box(-2 * np.pi, 2 * np.pi, shape=(12,))
  assert b.low.dtype == np.float32 and b.shape == (12,)
  assert float(b.high[0]) == 6.2831854820251465      # float32(2 pi), SURVEY D5


# ---- program epoch: what the env's fused-plan cache is keyed on (core/_epoch.py) --------------
def test_program_epoch_moves_on_every_change_that_a_fused_program_bakes_in():
  from gym_solo_b200.core import _epoch
  c = FakeClient()

  def moved(fn):
    before = _epoch.value
    fn()
    return _epoch.value != before

  of = obs.ObservationFactory(c)
  assert moved(lambda: of.register_observation(testing.CompliantObs(None)))
  rf = rewards.RewardFactory(c)
  flat = rewards.FlatTorsoReward(1, hard_margin=.1, soft_margin=1.)
  with warnings.catch_warnings():
    warnings.simplefilter('ignore', DeprecationWarning)
    assert moved(lambda: rf.register_reward(1, flat))
  add = rewards.AdditiveReward()
  add.client = c
  assert moved(lambda: add.add_term(0.5, rewards.UprightReward(1)))      # add_term after the first step
  assert moved(lambda: setattr(flat, '_soft_margin', 2.))                # a leaf's margins
  tb = terms.TimeBasedTermination(10)
  tf = terms.TerminationFactory()
  assert moved(lambda: tf.register_termination(tb))
  assert moved(lambda: setattr(tb, 'max_step_delta', 20))
  enc = obs.MotorEncoder(1, max_rotation=2.)
  assert moved(lambda: setattr(enc, '_max_rot', 3.))
  # reading does not move it
  assert not moved(lambda: (flat._soft_margin, tb.max_step_delta, rf.fused_ops()))
  # the mutated values are what the next compiled program carries
  ops = rf.fused_ops()
  assert any(abs(o.p[1] - 2.0) < 1e-12 for o in ops if o.op == _abi.REW_FLAT_TORSO)
