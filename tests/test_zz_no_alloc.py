"""solo_render_static (include/solo8_b200.h): render() for callers that cannot allocate device
memory -- the scratch of the two render kernels lives in a file-scope `static __device__`
array of SOLO_RENDER_STATIC_BYTES.  No counterpart in the reference.
Called through the C-ABI and compared with the existing entry point,
solo_render with caller scratch, on two calls in a row whose contents differ; the same two
kernels run on the same state, so the images have to be identical -- asserted exactly, which
is inside the render tests' tolerance (tests/test_gpu_render.py: at most 0.3 % of the pixels
off by more than one grey level).  The file sorts last on
purpose."""
import ctypes as C

import numpy as np
import pytest
import torch

from gym_solo_b200 import _abi, _lib, camera

pytestmark = pytest.mark.gpu

N_ENVS = 64
SOLO_EINVAL = -1          # include/solo8_b200.h


def _env():
  from gym_solo_b200.envs import Solo8VanillaEnv, Solo8VanillaConfig
  from gym_solo_b200.core import obs, rewards, termination
  cfg = Solo8VanillaConfig()
  cfg.dt = 1.0 / 240.0
  env = Solo8VanillaEnv(config=cfg, num_envs=N_ENVS, device='cuda:0')
  env.obs_factory.register_observation(obs.TorsoIMU(env.robot))
  env.reward_factory.register_reward(1, rewards.UprightReward(env.robot))
  env.termination_factory.register_termination(termination.PerpetualTermination())
  env.reset()
  return env


def _shake(env, steps, seed):
  g = torch.Generator(device='cuda').manual_seed(seed)
  for _ in range(steps):
    env.step((torch.rand((N_ENVS, env.client.n_api), generator=g, device='cuda') * 2 - 1) * 2 * np.pi)


def _camera(cfg, width, height):
  view = camera.compute_view_matrix_from_yaw_pitch_roll(cfg.render_pos, cfg.render_cam_distance, cfg.render_yaw,
                                                        cfg.render_pitch, cfg.render_roll, 2)
  proj = camera.compute_projection_matrix_fov(cfg.render_fov, width / height, 0.01, 100)
  cam, _ = _abi.make_camera(view, proj, width, height)
  return view, proj, cam


def _static(env, cam, begin, count, width, height, aux):
  """One call of the new entry point through ctypes: outputs are the caller's, nothing else."""
  c = env.client
  dev = c.device
  rgba = torch.zeros((count, height, width, 4), dtype=torch.uint8, device=dev)
  dep = torch.zeros((count, height, width), dtype=torch.float32, device=dev) if aux else None
  seg = torch.zeros((count, height, width), dtype=torch.int32, device=dev) if aux else None
  rc = c.lib.solo_render_static(c.handle, C.byref(cam), begin, count, C.c_void_p(rgba.data_ptr()),
                                C.c_void_p(dep.data_ptr() if aux else 0), C.c_void_p(seg.data_ptr() if aux else 0),
                                c._stream())
  return rc, rgba, dep, seg


def test_static_scratch_render_matches_caller_scratch_render_twice():
  env = _env()
  c = env.client
  if c._mesh is None:
    c.bind_mesh()
  assert hasattr(c.lib, 'solo_render_static')
  # ---- first call: 40 environments, the default camera, all three images ----
  _shake(env, 30, 11)
  w, h = env.config.render_width, env.config.render_height
  view, proj, cam = _camera(env.config, w, h)
  assert 0 < c.lib.solo_render_scratch_bytes(c.handle, 40) <= _abi.RENDER_STATIC_BYTES
  rc, rgba, dep, seg = _static(env, cam, 3, 40, w, h, True)
  _lib.check(rc, c.handle)
  _, _, want, want_dep, want_seg = c.getCameraImage(w, h, view, proj, env_begin=3, env_count=40)
  torch.cuda.synchronize()
  assert torch.equal(rgba, want) and torch.equal(dep, want_dep) and torch.equal(seg, want_seg)
  assert int((seg == 1).sum()) > 0 and not torch.equal(rgba[0], rgba[1])   # robots are in view, and differ
  first = rgba.clone()

  # ---- second call in a row: other states, fewer environments, another image size, colour only;
  # the array still holds the first call's records and boxes ----
  _shake(env, 25, 12)
  w2, h2 = 200, 120
  view2, proj2, cam2 = _camera(env.config, w2, h2)
  rc, rgba2, _, _ = _static(env, cam2, 0, 7, w2, h2, False)
  _lib.check(rc, c.handle)
  _, _, want2, _, _ = c.getCameraImage(w2, h2, view2, proj2, env_begin=0, env_count=7, depth=False,
                                       segmentation=False)
  torch.cuda.synchronize()
  assert torch.equal(rgba2, want2)

  # ---- and the first configuration again on the new states: differs from before, equals the reference ----
  rc, rgba3, dep3, seg3 = _static(env, cam, 3, 40, w, h, True)
  _lib.check(rc, c.handle)
  _, _, want3, want_dep3, want_seg3 = c.getCameraImage(w, h, view, proj, env_begin=3, env_count=40)
  torch.cuda.synchronize()
  assert torch.equal(rgba3, want3) and torch.equal(dep3, want_dep3) and torch.equal(seg3, want_seg3)
  assert not torch.equal(rgba3, first)
  env.close()


def test_static_scratch_render_rejects_what_does_not_fit():
  env = _env()
  c = env.client
  if c._mesh is None:
    c.bind_mesh()
  w, h = 64, 48
  _, _, cam = _camera(env.config, w, h)
  assert c.lib.solo_render_scratch_bytes(c.handle, N_ENVS) > _abi.RENDER_STATIC_BYTES
  rc, rgba, _, _ = _static(env, cam, 0, N_ENVS, w, h, False)
  assert rc == SOLO_EINVAL
  msg = c.lib.solo_last_error(c.handle).decode()
  assert 'SOLO_RENDER_STATIC_BYTES' in msg and str(_abi.RENDER_STATIC_BYTES) in msg
  torch.cuda.synchronize()
  assert int(rgba.sum()) == 0                                       # nothing ran
  # the argument checks of solo_render still apply
  rc, _, _, _ = _static(env, cam, N_ENVS - 2, 5, w, h, False)
  assert rc == SOLO_EINVAL
  env.close()
