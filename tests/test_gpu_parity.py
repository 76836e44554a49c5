"""GPU parity (run with -m gpu on a B200): the CUDA path, called through the
C-ABI (gym_solo_b200.client.BatchedClient is a thin ctypes wrapper over
include/solo8_b200.h), against the double-precision oracle on the same seeded
states and actions.

Stated tolerances (BASELINE.json north_star; float32 kernel vs float64 oracle):
  single step at dt = 1/240 (and 1e-3): joint position <= 1e-4 rad, joint
  velocity <= 1e-3 rad/s, base position <= 1e-5 m, base rotation <= 1e-4 rad,
  base linear velocity <= 1e-3 m/s, base angular velocity <= 2e-3 rad/s, normal
  contact force <= 0.05 N + 1e-3 relative.
Epilogue on identical states against the reference-generated goldens (float32 kernel vs the
reference's float64 numpy): observations <= 1e-5 in radians / metres per second, <= 2e-4 in
degrees (1e-6 of the 180 degree range), <= 1e-5 normalised; rewards <= 1e-5; UprightReward's
three known answers (test_rewards.py:74-92, exact in the reference) <= 1e-6.
"""
import ctypes as C
import os

import numpy as np
import pytest
import torch

from gym_solo_b200 import _abi
from helpers import GOLDEN, START_TARGETS, model_and_config, oracle, state_errors
from oracle import solo_oracle as so

pytestmark = pytest.mark.gpu

TOL = dict(base_pos=1e-5, base_rot=1e-4, base_lin=1e-3, base_ang=2e-3, q=1e-4, qd=1e-3)
OBS_TOL, OBS_TOL_DEG, REW_TOL = 1e-5, 2e-4, 1e-5


def make_client(n, **cfg):
  from gym_solo_b200.client import BatchedClient
  blob, model, config = model_and_config(**cfg)
  return BatchedClient(blob, model, config, n, 'cuda:0'), model, config


def rollout_states(model, config, n_states, seed=0, stride=7):
  """Oracle states sampled along a settle + random-action rollout (free fall,
  impact, lying flat, flailing): the pool parity is measured on."""
  o = oracle(model, config)
  rng = np.random.RandomState(seed)
  s = o.initial_state()
  out = []
  k = 0
  a = START_TARGETS
  while len(out) < n_states:
    if k >= 300 and k % 20 == 0:
      a = rng.uniform(-2 * np.pi, 2 * np.pi, 12)
    s = o.step(s, a)
    if k % stride == 0:
      out.append(s.copy())
    k += 1
  return np.stack(out)


@pytest.mark.parametrize('dt', [1.0 / 240.0, 1e-3])
def test_single_step_state_and_contact_force_parity(dt):
  n = 256
  c, model, config = make_client(n, dt=dt)
  states = rollout_states(model, config, n, seed=1)
  rng = np.random.RandomState(2)
  actions = rng.uniform(-2 * np.pi, 2 * np.pi, (n, 12))
  nf_dev, mt_dev = c.enable_debug(True)
  c.set_state(states)
  c.simulate(torch.tensor(actions, dtype=torch.float32, device='cuda:0'))
  got = c.get_state().double().cpu().numpy()
  nf = nf_dev.t().double().cpu().numpy()
  mt = mt_dev.t().double().cpu().numpy()
  o = oracle(model, config)
  want = np.zeros_like(states)
  nf_want = np.zeros_like(nf)
  mt_want = np.zeros_like(mt)
  for i in range(n):
    # the kernel sees float32 inputs: give the oracle the same rounded numbers
    s32 = states[i].astype(np.float32).astype(np.float64)
    a32 = actions[i].astype(np.float32).astype(np.float64)
    want[i], d = o.step(s32, a32, diag=True)
    nf_want[i], mt_want[i] = d['normal_force'], d['motor_torque']
  err = state_errors(got, want, model.ndof)
  for k, tol in TOL.items():
    assert err[k] <= tol, (k, err)
  assert np.all(np.abs(nf - nf_want) <= 0.05 + 1e-3 * np.abs(nf_want)), np.abs(nf - nf_want).max()
  assert np.abs(mt - mt_want).max() <= 2e-3
  assert nf_want.max() > 1.0   # the pool really contains contact
  c.disconnect()


def test_settle_snapshot_and_rollout_divergence():
  """reset()'s 500 settle steps on the GPU vs the oracle; divergence after
  1/10/100/1000 steps is reported (float32 vs float64 through contact is
  chaotic, so long horizons are bounded loosely)."""
  c, model, config = make_client(1, dt=1.0 / 240.0)
  o = oracle(model, config)
  s = o.initial_state()
  c.set_state(s)
  a = torch.tensor(START_TARGETS, dtype=torch.float32, device='cuda:0')
  a64 = START_TARGETS.astype(np.float32).astype(np.float64)
  report = {}
  k = 0
  for horizon in (1, 10, 100, 500):
    while k < horizon:
      s = o.step(s, a64)
      k += 1
    c.set_state(o.initial_state())
    c.simulate(a, n_steps=horizon, broadcast=True)
    report[horizon] = state_errors(c.get_state().double().cpu().numpy(), s, model.ndof)
  print('rollout divergence (settle):', report)
  assert report[1]['q'] <= 1e-6 and report[10]['q'] <= 1e-5
  assert report[100]['base_pos'] <= 1e-3
  assert report[500]['base_pos'] <= 5e-3 and report[500]['q'] <= 5e-2
  c.disconnect()


def _program(obs_terms=(), rew_ops=(), term_terms=(), auto_reset=False):
  p = _abi.SoloProgram()
  p.n_obs, p.n_rew, p.n_term = len(obs_terms), len(rew_ops), len(term_terms)
  p.auto_reset = 1 if auto_reset else 0
  for i, t in enumerate(obs_terms):
    p.obs[i] = t
  for i, t in enumerate(rew_ops):
    p.rew[i] = t
  for i, t in enumerate(term_terms):
    p.term[i] = t
  return p


def test_epilogue_against_reference_golden():
  """Observations / rewards of the fused program on the golden states vs the
  values the REFERENCE's classes produced (tests/golden)."""
  from gym_solo_b200.core import obs as bobs, rewards as brew
  G = np.load(os.path.join(GOLDEN, 'epilogue_golden.npz'))
  states = G['states']
  n = states.shape[0]
  c, model, config = make_client(n)
  c.set_state(states)

  def obs_case(mk, normalize=False):
    o = mk()
    o.client = c
    return o._fused_term(normalize), len(o.labels)

  cases = {
      'imu_rad': lambda: bobs.TorsoIMU(1),
      'imu_deg': lambda: bobs.TorsoIMU(1, degrees=True, max_lin_velocity=5, max_angular_velocity=180),
      'enc_rad': lambda: bobs.MotorEncoder(1),
      'enc_deg': lambda: bobs.MotorEncoder(1, degrees=True),
      'enc_clip': lambda: bobs.MotorEncoder(1, max_rotation=2.5),
      'enc_deg_clip': lambda: bobs.MotorEncoder(1, degrees=True, max_rotation=90.),
  }
  for name, mk in cases.items():
    o = mk(); o.client = c
    got = o.compute().double().cpu().numpy()
    want = G['obs_' + name]
    tol = OBS_TOL_DEG if 'deg' in name else OBS_TOL
    assert np.abs(got - want).max() <= tol, (name, np.abs(got - want).max())
  for norm in (False, True):
    for combo in (('imu_rad', 'enc_rad'), ('imu_deg', 'enc_clip'), ('enc_deg_clip', 'imu_rad')):
      f = bobs.ObservationFactory(c, normalize=norm)
      for k in combo:
        f.register_observation(cases[k]())
      key = 'factory_%s_%s_%s' % (combo[0], combo[1], 'norm' if norm else 'raw')
      # fused program path
      prog = _program(obs_terms=f.fused_terms())
      c.set_program(prog)
      out = torch.empty((n, c.obs_dim()), dtype=torch.float32, device='cuda:0')
      c.observe(out)
      got = out.double().cpu().numpy()
      want = G[key]
      tol = OBS_TOL if norm else OBS_TOL_DEG
      assert got.shape == want.shape
      assert np.abs(got - want).max() <= tol, (key, np.abs(got - want).max())
      assert f.labels() == list(G[key + '_labels'])
      # python (per-observation) path gives the same numbers
      got2 = f.get_obs()[0].double().cpu().numpy()
      assert np.abs(got2 - want).max() <= tol

  def home():
    return brew.HomePositionReward(1)
  rew_cases = {
      'upright': lambda: [(1, brew.UprightReward(1))],
      'flat_default': lambda: [(1, brew.FlatTorsoReward(1))],
      'flat_wide': lambda: [(1, brew.FlatTorsoReward(1, .1, np.pi))],
      'flat_step': lambda: [(1, brew.FlatTorsoReward(1, 1.0, 0.))],
      'small_default': lambda: [(1, brew.SmallControlReward(1))],
      'small_10': lambda: [(1, brew.SmallControlReward(1, margin=10))],
      'hspeed_0': lambda: [(1, brew.HorizontalMoveSpeedReward(1, 0, .5, 3))],
      'hspeed_2': lambda: [(1, brew.HorizontalMoveSpeedReward(1, 2.0))],
      'height_home': lambda: [(1, brew.TorsoHeightReward(1, 0.33698, 0.025, 0.15))],
      'height_default': lambda: [(1, brew.TorsoHeightReward(1, 0.3))],
      'home_pos': lambda: [(1, home())],
      'upright_plus_home': lambda: [(1, brew.UprightReward(1)), (1, home())],
      'mixed': lambda: [(0.25, brew.UprightReward(1)), (-2.0, brew.SmallControlReward(1, margin=10)),
                        (3.0, brew.MultiplicitiveReward(0.5, brew.FlatTorsoReward(1),
                                                        brew.TorsoHeightReward(1, 0.3)))],
  }
  import warnings
  for name, mk in rew_cases.items():
    f = brew.RewardFactory(c)
    with warnings.catch_warnings():
      warnings.simplefilter('ignore', DeprecationWarning)
      for w, r in mk():
        f.register_reward(w, r)
    ops = f.fused_ops()
    assert ops is not None
    c.set_program(_program(rew_ops=ops))
    out = torch.empty((n,), dtype=torch.float32, device='cuda:0')
    c.reward(out)
    got = out.double().cpu().numpy()
    want = G['rew_' + name]
    assert np.abs(got - want).max() <= REW_TOL, (name, np.abs(got - want).max())
    got2 = f.get_reward().double().cpu().numpy()    # python composition of kernel leaves
    assert np.abs(got2 - want).max() <= REW_TOL, (name, 'python path')
  c.disconnect()


def test_upright_reward_known_answers():
  """test_rewards.py:74-92 asserts these exactly (float64): pitch +90 deg -> -1.0, -90 deg -> +1.0,
  level -> 0.0.  The float32 kernel: <= 1e-6."""
  from gym_solo_b200.core import rewards as brew
  c, model, config = make_client(3)
  s = np.zeros((3, 29))
  s[0, 3:7] = so.quat_from_euler((0, np.pi / 2, 0))
  s[1, 3:7] = so.quat_from_euler((0, -np.pi / 2, 0))
  s[2, 3:7] = so.quat_from_euler((0, 0, 0))
  c.set_state(s)
  r = brew.UprightReward(1)
  r.client = c
  got = r.compute().double().cpu().numpy()
  assert np.abs(got - np.array([-1.0, 1.0, 0.0])).max() <= 1e-6, got
  assert got[2] == 0.0
  c.disconnect()


def test_termination_sequence_on_the_kernel():
  """The golden `term_seq` program of the reference's factory -- TimeBased(5) | Perpetual |
  TimeBased(3), short circuit, reset after call 8 (tools/gen_golden.py) -- through solo_step:
  done flags and the two counters of eval_terminations in finish_kernel."""
  from gym_solo_b200.core import obs as bobs, rewards as brew, termination as bterm
  G = np.load(os.path.join(GOLDEN, 'epilogue_golden.npz'))
  n = 5
  c, model, config = make_client(n, dt=1.0 / 240.0)
  c.set_state(c.initial_state((0, 0, .5), (0, 0, 0)))
  c.set_snapshot(c.initial_state((0, 0, .5), (0, 0, 0)))
  imu = bobs.TorsoIMU(1); imu.client = c
  terms = [bterm.TimeBasedTermination(5), bterm.PerpetualTermination(), bterm.TimeBasedTermination(3)]
  ops = []
  assert brew.UprightReward(1)._emit(ops)
  prog = _program(obs_terms=[imu._fused_term(False)], rew_ops=ops,
                  term_terms=[t._fused_term() for t in terms])
  c.set_program(prog)
  obs = torch.empty((n, 9), device='cuda:0'); rew = torch.empty((n,), device='cuda:0')
  done = torch.empty((n,), dtype=torch.uint8, device='cuda:0')
  a = torch.zeros((n, 12), device='cuda:0')
  seq = []
  for i in range(12):
    c.step(a, 1.0, obs, rew, done)
    d = done.cpu().numpy()
    assert (d == d[0]).all()
    seq.append(bool(d[0]))
    if i == 7:
      c.reset()          # termination_factory.reset(): every counter back to 0
  assert seq == [bool(x) for x in G['term_seq']], (seq, list(G['term_seq']))
  counters = c.counters.cpu().numpy()
  assert [int(counters[0, 0]), int(counters[2, 0])] == [int(x) for x in G['term_counters_end']]
  assert (counters[1] == 0).all()
  c.disconnect()


@pytest.mark.parametrize('n', [8192, 16384, 32768])
def test_large_batches_against_the_oracle(n):
  """The batch sizes of BASELINE configs[2] / [3] / [4] per GPU (64-thread blocks above 9472
  environments; pick_block in solo_abi.cu): 256 environments sampled across the batch, single
  step with contact forces.  (a) The result of an environment does not depend on the batch it
  sits in: bit-identical to the same (state, action) pairs stepped as a 256-environment batch
  (32-thread blocks).  (b) Against the oracle: the common tolerance where the oracle's PGS
  converged.  Environments stopped by the 50-iteration cap are mid-way through a solve, and
  some of those states are discontinuous in their inputs (a contact or a friction cone flips):
  the ORACLE's own answer moves by up to 0.6 rad/s when its float64 input is perturbed by one
  float32 ulp (measured on this sample).  For them the bound is 5 x the tolerance plus twice
  the oracle's own sensitivity to such a perturbation."""
  c, model, config = make_client(n, dt=1.0 / 240.0)
  pool = rollout_states(model, config, 256, seed=3)
  rng = np.random.RandomState(n)
  which = rng.randint(0, 256, n)
  states = pool[which]
  actions = rng.uniform(-2 * np.pi, 2 * np.pi, (n, 12))
  nf_dev, _ = c.enable_debug(True)
  c.set_state(states)
  c.simulate(torch.tensor(actions, dtype=torch.float32, device='cuda:0'))
  sample = np.unique(np.concatenate([[0, 1, 31, 32, 63, 64, n - 65, n - 64, n - 33, n - 32, n - 1],
                                     rng.randint(0, n, 245)]))
  idx = torch.tensor(sample, device='cuda:0')
  got_t = c.get_state()[idx]
  got = got_t.double().cpu().numpy()
  nf = nf_dev.t()[idx].double().cpu().numpy()
  assert torch.isfinite(c.get_state()).all() and int(c.diag[2]) == 0
  c.disconnect()
  # (a) the same pairs as a small batch
  small, _, _ = make_client(len(sample), dt=1.0 / 240.0)
  small.set_state(states[sample])
  small.simulate(torch.tensor(actions[sample], dtype=torch.float32, device='cuda:0'))
  assert torch.equal(small.get_state(), got_t)
  small.disconnect()
  # (b) the oracle
  o = oracle(model, config)
  want = np.zeros_like(got)
  nf_want = np.zeros_like(nf)
  capped = np.zeros(len(sample), dtype=bool)
  sens = np.zeros((len(sample), states.shape[1]))
  prng = np.random.RandomState(0)
  for k, i in enumerate(sample):
    s32 = states[i].astype(np.float32).astype(np.float64)
    a32 = actions[i].astype(np.float32).astype(np.float64)
    want[k], d = o.step(s32, a32, diag=True)
    nf_want[k] = d['normal_force']
    capped[k] = d['iterations'] >= config.solver_iterations
    if capped[k]:
      for _ in range(8):       # the oracle's own sensitivity to a one-ulp (float32) input perturbation
        sp = s32 * (1 + 6e-8 * prng.standard_normal(s32.shape[0]))
        sp[3:7] /= np.linalg.norm(sp[3:7])
        sens[k] = np.maximum(sens[k], np.abs(o.step(sp, a32) - want[k]))
  err = state_errors(got[~capped], want[~capped], model.ndof)
  for k, tol in TOL.items():
    assert err[k] <= tol, (n, k, err)
  for k in np.nonzero(capped)[0]:
    e = state_errors(got[k], want[k], model.ndof)
    allow = 2 * sens[k].max()
    print('n = %d env %d (iteration cap): errors %s, oracle sensitivity %.3g' % (n, sample[k], e, sens[k].max()))
    for key, tol in TOL.items():
      assert e[key] <= 5 * tol + allow, (n, sample[k], key, e, sens[k].max())
  ok = np.abs(nf - nf_want) <= 0.05 + 1e-3 * np.abs(nf_want)
  assert ok[~capped].all(), np.abs(nf - nf_want)[~capped].max()
