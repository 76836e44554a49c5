"""Loader / builder of the C-ABI shared library (include/solo8_b200.h).

The product path is this library and nothing else: if it is missing or there
is no CUDA device every entry below raises -- there is no CPU fallback.
"""
import ctypes as C
import os
import subprocess

from gym_solo_b200 import _abi

_PKG = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_PKG)
LIB_PATH = os.path.join(_PKG, 'libsolo8_b200.so')
SOURCES = [os.path.join(_PKG, 'csrc', f) for f in
           ('solo_abi.cu', 'solo_kernels.cuh', 'solo_core.h', 'solo_phases.h', 'solo_setup.h', 'solo_solve.h',
            'solo_finish.h', 'solo_solve_warp.cuh', 'solo_contact.cuh', 'solo_convert.h', 'solo_render.cuh', 'solo_setup_legpar.cuh')]
HEADER = os.path.join(_ROOT, 'include', 'solo8_b200.h')
NVCC_FLAGS = ['-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-O3',
              '-std=c++17', '-shared', '-Xcompiler', '-fPIC']

_lib = None


def _stale():
  if not os.path.exists(LIB_PATH):
    return True
  t = os.path.getmtime(LIB_PATH)
  return any(os.path.exists(s) and os.path.getmtime(s) > t for s in SOURCES + [HEADER])


def build(force=False, verbose=False):
  """nvcc -gencode arch=compute_100a,code=sm_100a ... -> libsolo8_b200.so (in tree)."""
  if not force and not _stale():
    return LIB_PATH
  nvcc = os.environ.get('NVCC', 'nvcc')
  cmd = [nvcc] + NVCC_FLAGS + (['-Xptxas', '-v'] if verbose else []) + ['-o', LIB_PATH, SOURCES[0]]
  subprocess.check_call(cmd, cwd=os.path.join(_PKG, 'csrc'))
  return LIB_PATH


EXPORTS = {
    # name: (restype, argtypes)
    'solo_version': (C.c_char_p, []),
    'solo_last_error': (C.c_char_p, [C.c_void_p]),
    'solo_create': (C.c_int, [C.POINTER(_abi.SoloModel), C.POINTER(_abi.SoloConfig), C.c_int32,
                              C.c_int32, C.POINTER(C.c_void_p)]),
    'solo_destroy': (C.c_int, [C.c_void_p]),
    'solo_set_program': (C.c_int, [C.c_void_p, C.POINTER(_abi.SoloProgram)]),
    'solo_obs_dim': (C.c_int, [C.c_void_p]),
    'solo_scratch_bytes': (C.c_int64, [C.c_void_p]),
    'solo_bind': (C.c_int, [C.c_void_p] * 6),
    'solo_snapshot_layout': (C.c_int, [C.c_void_p, C.c_int32]),
    'solo_bind_diag': (C.c_int, [C.c_void_p, C.c_void_p]),
    'solo_bind_link_friction': (C.c_int, [C.c_void_p, C.c_void_p]),
    'solo_bind_base_mass_scale': (C.c_int, [C.c_void_p, C.c_void_p]),
    'solo_bind_debug': (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    'solo_bind_staging': (C.c_int, [C.c_void_p] * 5),
    'solo_simulate': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]),
    'solo_step': (C.c_int, [C.c_void_p, C.c_void_p, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p,
                            C.c_void_p, C.c_void_p]),
    'solo_observe': (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    'solo_reward': (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    'solo_reset': (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    'solo_step_host': (C.c_int, [C.c_void_p, C.c_void_p, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p,
                                 C.c_void_p]),
    'solo_set_option': (C.c_int, [C.c_void_p, C.c_int32, C.c_int32]),
    'solo_launch_count': (C.c_int64, [C.c_void_p]),
    'solo_phase_timing': (C.c_int, [C.c_void_p, C.c_int32]),
    'solo_phase_times': (C.c_int64, [C.c_void_p, C.POINTER(C.c_double)]),
    'solo_block_size': (C.c_int, [C.c_void_p]),
    'solo_solver_stats': (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)]),
    'solo_bind_mesh': (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32]),
    'solo_render_scratch_bytes': (C.c_int64, [C.c_void_p, C.c_int32]),
    'solo_render': (C.c_int, [C.c_void_p, C.POINTER(_abi.SoloCamera), C.c_int32, C.c_int32, C.c_void_p,
                              C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    'solo_render_static': (C.c_int, [C.c_void_p, C.POINTER(_abi.SoloCamera), C.c_int32, C.c_int32, C.c_void_p,
                                     C.c_void_p, C.c_void_p, C.c_void_p]),
}


def load():
  """dlopen the library and type every export; raises if it is not built."""
  global _lib
  if _lib is None:
    if not os.path.exists(LIB_PATH):
      raise RuntimeError(
          'gym_solo_b200: %s is not built (run `python -c "import __graft_entry__ as g; g.build()"`); '
          'the Solo 8 step path has no CPU fallback' % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in EXPORTS.items():
      fn = getattr(lib, name)
      fn.restype = res
      fn.argtypes = args
    _lib = lib
  return _lib


class SoloError(RuntimeError):
  pass


def check(rc, handle=None):
  if rc != 0:
    msg = load().solo_last_error(handle)
    raise SoloError('solo8_b200 error %d: %s' % (rc, msg.decode() if msg else '?'))
