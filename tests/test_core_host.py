"""The CUDA kernels' per-environment core (gym_solo_b200/csrc/solo_phases.h:
setup_step -> solve_rows -> finish_step) compiled for the host (tests/hostsim)
against the oracle.  No GPU needed: this pins the kernel ARITHMETIC on CPU; the
`-m gpu` tests pin the kernels themselves."""
import numpy as np
import pytest

import helpers as H
from gym_solo_b200 import _abi


@pytest.fixture(scope='module')
def world():
  blob, model, config = H.model_and_config(dt=1 / 240.)
  orc = H.oracle(model, config)
  snap = orc.settle(H.START_TARGETS)
  return model, config, orc, snap


def test_float64_core_tracks_oracle_over_contact_rollout(world):
  """Same algorithm, different formulation (whitened rows vs dense M^-1 J^T):
  in double precision the two must agree to rounding over a contact-rich rollout."""
  model, config, orc, snap = world
  rng = np.random.RandomState(3)
  a, b = snap.copy(), snap.copy()
  worst_free, worst_step = 0.0, 0.0
  its = []
  for k in range(300):
    act = rng.uniform(-2 * np.pi, 2 * np.pi, 12)
    c, d = H.hostsim_step(model, config, a, act, use_double=True)    # same state in
    b, _ = H.hostsim_step(model, config, b, act, use_double=True)    # free running
    a = orc.step(a, act)
    its.append(d['iterations'])
    worst_step = max(worst_step, np.abs(a - c).max())
    worst_free = max(worst_free, np.abs(a - b).max())
  assert worst_step < 1e-10, worst_step
  assert worst_free < 1e-5, worst_free
  assert max(its) > 10          # the rollout exercises long PGS solves too


def test_float32_core_single_step_within_north_star_tolerance(world):
  """BASELINE.json north_star: <= 1e-4 rad and <= 1e-3 rad/s per step at dt = 1/240."""
  model, config, orc, snap = world
  rng = np.random.RandomState(4)
  s = snap.copy()
  worst = dict(q=0., qd=0., base_pos=0., base_rot=0., base_lin=0., base_ang=0.)
  nf_err = 0.0
  for k in range(200):
    act = rng.uniform(-2 * np.pi, 2 * np.pi, 12)
    s32 = s.astype(np.float32).astype(np.float64)
    a32 = act.astype(np.float32).astype(np.float64)
    want, wd = orc.step(s32, a32, diag=True)
    got, d = H.hostsim_step(model, config, s32, a32, use_double=False)
    e = H.state_errors(got, want, model.ndof)
    for key in worst:
      worst[key] = max(worst[key], e[key])
    assert d['n_contacts'] == wd['n_contacts']
    nf_err = max(nf_err, np.abs(d['normal_force'] - wd['normal_force'][:model.n_cand]).max())
    s = want
  assert worst['q'] <= 1e-4 and worst['qd'] <= 1e-3, worst
  assert worst['base_pos'] <= 1e-5 and worst['base_rot'] <= 1e-4, worst
  assert worst['base_lin'] <= 1e-3 and worst['base_ang'] <= 2e-3, worst
  assert nf_err <= 0.5, nf_err      # N, contact normal force per support point


def test_torque_mode_and_free_flight(world):
  model, _, _, snap = world
  _, _, cfg = H.model_and_config(dt=1 / 240., control_mode=1)
  orc = H.oracle(model, cfg)
  s = snap.copy()
  s[2] += 0.5          # airborne: no contacts, no motor rows -> zero PGS iterations
  act = np.zeros(12)
  act[[0, 1, 3, 4]] = [0.5, -0.3, 1.0, 2.5]     # 2.5 > limit 2 -> clamped
  limited = 0
  for _ in range(50):       # (the robot lands after ~55 steps)
    want = orc.step(s, act)
    got, d = H.hostsim_step(model, cfg, s, act, use_double=True)
    assert d['n_contacts'] == 0
    if np.abs(s[13:21]).max() < 10.0:
      assert d['iterations'] == 0       # no rows at all
    else:
      # the undamped legs reach the URDF limit (+-10 rad, solo.urdf:47) within a few dozen steps:
      # joint-limit rows, the only rows of this solve
      assert d['iterations'] >= 1
      limited += 1
    np.testing.assert_allclose(got, want, atol=1e-10)
    s = want
  assert limited > 10
  assert np.abs(s[13:21]).max() < 10.5   # held at the limit, not running away at 100 rad/s


@pytest.mark.parametrize('name', ['solo8_mmr_v4', 'solo8_urdf_v3'])
def test_mmr_v4_model_float64_core_tracks_oracle(name):
  """BASELINE configs[3] model (solo8_URDF_v4: 10 dof, hind chains of 3 joints ->
  the CH = 3 instantiation of the core) and the unreferenced v3 URDF the model
  compiler also takes."""
  blob, model, config = H.model_and_config(name, dt=1 / 240.)
  assert model.ndof == 10 and list(model.chain_len)[:4] == [3, 3, 2, 2]
  orc = H.oracle(model, config)
  a = orc.settle(np.zeros(10))
  rng = np.random.RandomState(7)
  worst, its = 0.0, []
  for k in range(200):
    act = rng.uniform(-2 * np.pi, 2 * np.pi, 10)
    c, d = H.hostsim_step(model, config, a, act, use_double=True)
    a = orc.step(a, act)
    its.append(d['iterations'])
    worst = max(worst, np.abs(a - c).max())
  assert worst < 1e-9, worst
  assert max(its) > 10


def test_mmr_v4_float32_core_and_where_its_error_comes_from():
  """The few-gram distal links of the MMR v4 model in float32 (the arithmetic of the CUDA path):
  leg blocks formed about the joints' own origins keep the per-step joint-rate error at the
  1e-3 level on resting multi-contact states (round 1: 2e-2); the splice runs show that the
  mass-matrix factors are no longer the source (any float32 input of the non-converged
  28-contact solve gives the same ~1e-3)."""
  import ctypes as C
  blob, model, config = H.model_and_config('solo8_mmr_v4', dt=1 / 240.)
  orc = H.oracle(model, config)
  lib = H.hostsim()
  rng = np.random.RandomState(12)
  s = orc.initial_state()
  pool = []
  for k in range(330):
    s = orc.step(s, np.zeros(10) if k < 200 or k % 15 else rng.uniform(-2 * np.pi, 2 * np.pi, 10))
    if k % 5 == 0:
      pool.append(s.copy())
  worst, worst_fac = 0.0, 0.0
  for s in pool:
    s32 = s.astype(np.float32).astype(np.float64)
    a32 = rng.uniform(-2 * np.pi, 2 * np.pi, 10).astype(np.float32).astype(np.float64)
    want = orc.step(s32, a32)
    got, _ = H.hostsim_step(model, config, s32, a32, use_double=False)
    worst = max(worst, np.abs(got - want)[23:].max())
    st = s32.copy()
    rc = lib.hostsim_step_splice(C.byref(model), C.byref(config), 2, st.ctypes.data_as(C.POINTER(C.c_double)),
                                 a32.ctypes.data_as(C.POINTER(C.c_double)))
    if rc == 0:
      worst_fac = max(worst_fac, np.abs(st - want)[23:].max())
  assert worst <= 4e-3, worst
  assert worst_fac <= 5e-4, worst_fac       # float32 factors (Ls, Bh, E) in a float64 step
