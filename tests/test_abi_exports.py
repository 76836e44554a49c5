"""The C-ABI library loads without a GPU and exports every symbol that
include/solo8_b200.h declares; with no CUDA device solo_create fails loudly
(no CPU fallback).  No compute calls here."""
import ctypes as C
import os
import re

import pytest

from gym_solo_b200 import _abi, _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
  src = open(os.path.join(ROOT, 'include', 'solo8_b200.h')).read()
  src = re.sub(r'/\*.*?\*/', '', src, flags=re.S)
  return sorted(set(re.findall(r'\b(solo_[a-z_0-9]+)\s*\(', src)))


@pytest.fixture(scope='module')
def lib():
  if not os.path.exists(_lib.LIB_PATH):
    _lib.build()
  return _lib.load()


def test_every_declared_symbol_is_exported(lib):
  names = _declared()
  assert len(names) >= 20
  missing = [n for n in names if not hasattr(lib, n)]
  assert not missing, missing
  # and every symbol the Python layer binds is declared in the header
  assert sorted(_lib.EXPORTS) == names


def test_struct_sizes_match_header_layout():
  # the compiled library checks sizeof(SoloProgram) etc. with static_asserts;
  # here: the ctypes mirrors are self-consistent with the documented limits
  assert C.sizeof(_abi.SoloObsTerm) == 16 + 4 * 4 * _abi.MAX_API
  assert C.sizeof(_abi.SoloRewardOp) == 8 + 16 + 4 * 8
  assert C.sizeof(_abi.SoloTermTerm) == 8


def test_create_fails_loudly_without_a_gpu(lib):
  import torch
  if torch.cuda.is_available():
    pytest.skip('a CUDA device is present')
  blob = _abi.load_model_blob('solo8v2')
  model = _abi.make_model(blob, link_friction=0.5)
  cfg = _abi.make_config(dt=1 / 240.)
  h = C.c_void_p()
  rc = lib.solo_create(C.byref(model), C.byref(cfg), 16, 0, C.byref(h))
  assert rc == -3 and not h.value          # SOLO_ENOGPU
  assert b'no CPU fallback' in lib.solo_last_error(None)
