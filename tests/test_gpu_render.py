"""render() on the GPU against the image oracle
(oracle/solo_render_oracle.py, float64): the CUDA rasteriser computes in float32,
so a pixel whose centre lies within rounding of a triangle edge, a cell boundary
of the checkerboard or a depth tie may pick the other side.  Stated tolerance:
at most 0.3 % of the pixels differ by more than 1 grey level in any channel."""
import numpy as np
import pytest
import torch

from gym_solo_b200 import _abi, camera
from oracle import solo_render_oracle as ro

pytestmark = pytest.mark.gpu

MAX_BAD_FRACTION = 3e-3


def _mismatch(a, b):
  return float((np.abs(a.astype(int) - b.astype(int)).max(-1) > 1).mean())


def _env(n, urdf=None):
  from gym_solo_b200.envs import Solo8VanillaEnv, Solo8VanillaConfig, Solo8MMRConfig
  from gym_solo_b200.core import obs, rewards, termination
  cfg = Solo8MMRConfig() if urdf == 'v4' else Solo8VanillaConfig()
  cfg.dt = 1.0 / 240.0
  env = Solo8VanillaEnv(config=cfg, num_envs=n, device='cuda:0')
  env.obs_factory.register_observation(obs.TorsoIMU(env.robot))
  env.reward_factory.register_reward(1, rewards.UprightReward(env.robot))
  env.termination_factory.register_termination(termination.PerpetualTermination())
  env.reset()
  return env


def _shake(env, steps, seed):
  import torch
  g = torch.Generator(device='cuda').manual_seed(seed)
  n_api = env.client.n_api
  for _ in range(steps):
    a = (torch.rand((env.num_envs, n_api), generator=g, device='cuda') * 2 - 1) * 2 * np.pi
    env.step(a)


@pytest.mark.parametrize('urdf', ['v2', 'v4'])
def test_render_matches_the_oracle_per_environment(urdf):
  env = _env(6, urdf)
  _shake(env, 40, 7)
  cfg = env.config
  imgs = env.render(envs=(0, 6))
  assert imgs.shape == (6, cfg.render_height, cfg.render_width, 4) and imgs.dtype.is_floating_point is False
  imgs = imgs.cpu().numpy()
  states = env.get_state().cpu().numpy().astype(np.float64)
  view = camera.compute_view_matrix_from_yaw_pitch_roll(cfg.render_pos, cfg.render_cam_distance, cfg.render_yaw,
                                                        cfg.render_pitch, cfg.render_roll, 2)
  proj = camera.compute_projection_matrix_fov(cfg.render_fov, cfg.render_aspect, 0.01, 100)
  vis = _abi.load_visual(env.client.blob)
  assert not np.array_equal(imgs[0], imgs[1])          # the environments really differ
  for e in range(6):
    want = ro.render(env.client.blob, vis, states[e], view, proj, cfg.render_width, cfg.render_height,
                     **_abi.RENDER_DEFAULTS)
    bad = _mismatch(imgs[e], want)
    assert bad <= MAX_BAD_FRACTION, (e, bad)
    robot = (want[..., 0] == want[..., 1]) & (want[..., 0] < 200)
    assert robot.sum() > 800                            # the robot is in the picture


def test_render_default_call_is_the_reference_shape():
  env = _env(3)
  img = env.render()
  assert isinstance(img, np.ndarray) and img.shape == (369, 369, 4) and img.dtype == np.uint8
  assert (img[..., 3] == 255).all()
  np.testing.assert_array_equal(img, env.render(envs=(0, 1))[0].cpu().numpy())
  with pytest.raises(ValueError):
    env.render(mode='human')


def test_render_other_cameras_sizes_and_shading():
  env = _env(4)
  _shake(env, 25, 3)
  c = env.client
  states = env.get_state().cpu().numpy().astype(np.float64)
  vis = _abi.load_visual(c.blob)
  cases = [
      dict(size=(200, 120), ypr=(35., -35., 0.), dist=0.8, fov=60, shade={}),
      dict(size=(97, 131), ypr=(-120., -5., 10.), dist=1.5, fov=100,
           shade=dict(light_direction=(1., 0., 0.3), ambient=0.3, diffuse=0.7, background=(0.1, 0.2, 0.3),
                      plane_rgb_a=(0.9, 0.2, 0.2), plane_rgb_b=(0.2, 0.9, 0.2), checker_size=0.25)),
      dict(size=(64, 64), ypr=(0., -89., 0.), dist=2.0, fov=45, shade={}),     # straight down
  ]
  for k in cases:
    w, h = k['size']
    view = c.computeViewMatrixFromYawPitchRoll([0, 0, .2], k['dist'], *k['ypr'], 2)
    proj = c.computeProjectionMatrixFOV(k['fov'], w / h, 0.01, 100)
    # pybullet's five return values: width, height, rgba, depth buffer, segmentation mask
    ww, hh, rgba, depth, seg = c.getCameraImage(w, h, view, proj, env_begin=1, env_count=3, **k['shade'])
    assert (ww, hh) == (w, h) and tuple(rgba.shape) == (3, h, w, 4)
    assert tuple(depth.shape) == (3, h, w) and depth.dtype == torch.float32
    assert tuple(seg.shape) == (3, h, w) and seg.dtype == torch.int32
    rgba, depth, seg = rgba.cpu().numpy(), depth.cpu().numpy(), seg.cpu().numpy()
    opts = dict(_abi.RENDER_DEFAULTS, **k['shade'])
    for i in range(3):
      want, want_depth, want_seg = ro.render(c.blob, vis, states[1 + i], view, proj, w, h, aux=True, **opts)
      bad = _mismatch(rgba[i], want)
      assert bad <= MAX_BAD_FRACTION, (k['size'], i, bad)
      same = seg[i] == want_seg
      assert 1.0 - same.mean() <= MAX_BAD_FRACTION, (k['size'], i, 1.0 - same.mean())
      assert set(np.unique(seg[i])) <= {-1, 0, 1} and (seg[i] == 1).any()
      # where both agree on what was hit, the depth buffers agree: 1e-6 on average (float32 ndc
      # depth has 6e-8 resolution near the far plane, the oracle is float64), 3e-4 at worst (sliver
      # triangles: their affine depth equation has coefficients of order 1 / area)
      err = np.abs(depth[i] - want_depth)[same]
      assert err.mean() <= 1e-6 and err.max() <= 3e-4, (err.mean(), err.max())
      assert (depth[i][seg[i] == -1] == 1.0).all() and depth[i].min() >= 0.0
    # the images that were not asked for are not produced
    _, _, _, d0, s0 = c.getCameraImage(w, h, view, proj, env_begin=0, env_count=1, depth=False, segmentation=False)
    assert d0 is None and s0 is None


def test_render_argument_errors():
  from gym_solo_b200._lib import SoloError
  env = _env(2)
  c = env.client
  view = c.computeViewMatrixFromYawPitchRoll([0, 0, .2], 1, 0, -20, 0, 2)
  proj = c.computeProjectionMatrixFOV(80, 1.0, 0.01, 100)
  with pytest.raises(SoloError):
    c.getCameraImage(32, 32, view, proj, env_begin=1, env_count=2)       # past the last environment
  with pytest.raises(SoloError):
    c.getCameraImage(32, 32, view, [1.0] * 16)                           # not a perspective projection
  with pytest.raises(SoloError):
    c.getCameraImage(0, 32, view, proj)
