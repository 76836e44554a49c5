"""Shared test plumbing: model / config fixtures, the oracle, the host build of
the kernel core (tests/hostsim) and tolerance helpers."""
import ctypes as C
import os

import numpy as np

from gym_solo_b200 import _abi
from oracle import solo_oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, 'tests', 'golden')

START_TARGETS = np.array([np.pi / 2, np.pi, 0, np.pi / 2, np.pi, 0,
                          -np.pi / 2, -np.pi, 0, -np.pi / 2, -np.pi, 0])   # solo8v2vanilla.py:21-34


def model_and_config(name='solo8v2', **cfg):
  blob = _abi.load_model_blob(name)
  model = _abi.make_model(blob, link_friction=0.5)    # configs.py:24 via changeDynamics
  config = _abi.make_config(**cfg)
  return blob, model, config


def oracle(model, config):
  return solo_oracle.PhysicsOracle(model, config)


_HOSTSIM = None


def hostsim():
  global _HOSTSIM
  if _HOSTSIM is None:
    lib = C.CDLL(os.path.join(ROOT, 'tests', 'hostsim', 'libhostsim.so'))
    assert lib.hostsim_sizeof_model() == C.sizeof(_abi.SoloModel)
    assert lib.hostsim_sizeof_program() == C.sizeof(_abi.SoloProgram)
    _HOSTSIM = lib
  return _HOSTSIM


def _dp(a):
  return a.ctypes.data_as(C.POINTER(C.c_double))


def hostsim_step(model, config, state, action, use_double=False):
  """One env.step worth of physics through the kernel core compiled for the host."""
  st = np.ascontiguousarray(state, dtype=np.float64).copy()
  a = np.ascontiguousarray(action, dtype=np.float64)
  diag = (C.c_int32 * 2)()
  nf = np.zeros(_abi.MAX_CAND)
  mt = np.zeros(_abi.MAX_DOF)
  rc = hostsim().hostsim_step(C.byref(model), C.byref(config), 1 if use_double else 0, _dp(st), _dp(a),
                              diag, _dp(nf), _dp(mt))
  assert rc == 0
  return st, dict(n_contacts=diag[0], iterations=diag[1], normal_force=nf[:model.n_cand],
                  motor_torque=mt[:model.ndof])


def hostsim_epilogue(model, program, state, use_double=False):
  obs = np.zeros(_abi.MAX_OBS_TERMS * _abi.MAX_API)
  rew = C.c_double(0)
  st = np.ascontiguousarray(state, dtype=np.float64)
  n = hostsim().hostsim_epilogue(C.byref(model), C.byref(program), 1 if use_double else 0, _dp(st), _dp(obs),
                                 C.byref(rew))
  assert n >= 0, n
  return obs[:n].copy(), rew.value


def quat_angle(qa, qb):
  """Rotation angle between two (x,y,z,w) quaternions, batched."""
  d = np.abs(np.sum(np.asarray(qa) * np.asarray(qb), axis=-1))
  d = d / (np.linalg.norm(qa, axis=-1) * np.linalg.norm(qb, axis=-1))
  return 2 * np.arccos(np.clip(d, 0, 1))


def state_errors(a, b, ndof):
  """max-abs errors per group between two [.., sdim] state arrays."""
  a, b = np.atleast_2d(a), np.atleast_2d(b)
  sb = _abi.STATE_BASE
  return dict(base_pos=np.abs(a[:, 0:3] - b[:, 0:3]).max(),
              base_rot=quat_angle(a[:, 3:7], b[:, 3:7]).max(),
              base_lin=np.abs(a[:, 7:10] - b[:, 7:10]).max(),
              base_ang=np.abs(a[:, 10:13] - b[:, 10:13]).max(),
              q=np.abs(a[:, sb:sb + ndof] - b[:, sb:sb + ndof]).max(),
              qd=np.abs(a[:, sb + ndof:] - b[:, sb + ndof:]).max())
