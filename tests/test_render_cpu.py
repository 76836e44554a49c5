"""render() host side: camera matrices against hand-derived known
answers, the image oracle's size-independent properties, the compiled meshes.
The reference pins nothing here (no render test in /root/reference/tests)."""
import numpy as np
import pytest

from gym_solo_b200 import _abi, camera
from oracle import solo_render_oracle as ro


def _mat(m):
  return np.asarray(m, dtype=np.float64).reshape(4, 4).T


def test_projection_matrix_known_answers():
  p = _mat(camera.compute_projection_matrix_fov(90, 1.0, 1.0, 3.0))
  # tan(45 deg) = 1; (n + f) / (n - f) = -2; 2 n f / (n - f) = -3
  want = np.array([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, -2, -3], [0, 0, -1, 0]], dtype=np.float64)
  np.testing.assert_allclose(p, want, atol=1e-12)
  p = _mat(camera.compute_projection_matrix_fov(80, 2.0, 0.01, 100))     # solo8_base_env.py:129-131
  assert p[1, 1] == pytest.approx(1 / np.tan(np.radians(40)))
  assert p[0, 0] == pytest.approx(p[1, 1] / 2)
  # the near plane maps to ndc z = -1, the far plane to +1
  for z, ndc in ((-0.01, -1.0), (-100.0, 1.0)):
    c = p @ np.array([0, 0, z, 1.0])
    assert c[2] / c[3] == pytest.approx(ndc)


def test_view_matrix_known_answers():
  # Z up, no angles: the camera sits `distance` behind the target on -Y and looks along +Y
  v = _mat(camera.compute_view_matrix_from_yaw_pitch_roll([1, 2, 3], 4, 0, 0, 0, 2))
  eye = -v[:3, :3].T @ v[:3, 3]
  np.testing.assert_allclose(eye, [1, -2, 3], atol=1e-12)
  np.testing.assert_allclose(v[:3, :3], [[1, 0, 0], [0, 0, 1], [0, -1, 0]], atol=1e-12)
  # the target lands on the optical axis at depth `distance`
  np.testing.assert_allclose(v @ np.array([1, 2, 3, 1.0]), [0, 0, -4, 1], atol=1e-12)
  # yaw 90 degrees turns the camera about +Z: it now sits on +X looking along -X
  v = _mat(camera.compute_view_matrix_from_yaw_pitch_roll([0, 0, 0], 2, 90, 0, 0, 2))
  np.testing.assert_allclose(-v[:3, :3].T @ v[:3, 3], [2, 0, 0], atol=1e-12)
  # pitch -20 (the reference's default, configs.py:32) lifts the camera above the target
  v = _mat(camera.compute_view_matrix_from_yaw_pitch_roll([0, 0, .2], 1, 0, -20, 0, 2))
  np.testing.assert_allclose(-v[:3, :3].T @ v[:3, 3],
                             [0, -np.cos(np.radians(20)), .2 + np.sin(np.radians(20))], atol=1e-12)
  # Y up: the camera starts on -Z
  v = _mat(camera.compute_view_matrix_from_yaw_pitch_roll([0, 0, 0], 3, 0, 0, 0, 1))
  np.testing.assert_allclose(-v[:3, :3].T @ v[:3, 3], [0, 0, -3], atol=1e-12)
  with pytest.raises(ValueError):
    camera.compute_view_matrix_from_yaw_pitch_roll([0, 0, 0], 1, 0, 0, 0, 0)
  # rigid: rotation block orthonormal, right-handed
  R = v[:3, :3]
  np.testing.assert_allclose(R @ R.T, np.eye(3), atol=1e-12)
  assert np.linalg.det(R) == pytest.approx(1.0)


@pytest.mark.parametrize('name', ['solo8v2', 'solo8_mmr_v4', 'solo8_v3'])
def test_compiled_meshes_are_closed_outward_hulls(name):
  blob = _abi.load_model_blob({'solo8_v3': 'solo8_urdf_v3'}.get(name, name))
  vis = _abi.load_visual(blob)
  body = np.asarray(vis['tri_body'])
  verts = np.asarray(vis['tri_verts']).reshape(-1, 3, 3)
  assert vis['n_tri'] == len(body) == len(verts) == len(vis['tri_rgba']) > 100
  assert body.min() >= 0 and body.max() <= blob['ndof']
  # a closed surface with outward counter-clockwise faces has positive volume by the
  # divergence theorem, and its signed face-area vectors sum to zero
  vol = np.einsum('ni,ni->', verts[:, 0], np.cross(verts[:, 1], verts[:, 2])) / 6
  area_vec = 0.5 * np.cross(verts[:, 1] - verts[:, 0], verts[:, 2] - verts[:, 0]).sum(0)
  assert vol > 1e-5
  np.testing.assert_allclose(area_vec, 0, atol=1e-5)


def _standing_state(blob, z=0.30):
  nd = blob['ndof']
  st = np.zeros(13 + 2 * nd)
  st[2], st[6] = z, 1.0
  return st


def _default_camera(w=123, h=123):
  view = camera.compute_view_matrix_from_yaw_pitch_roll([0, 0, .2], 1, 0, -20, 0, 2)
  proj = camera.compute_projection_matrix_fov(80, w / h, 0.01, 100)
  return view, proj


def test_oracle_image_properties():
  blob = _abi.load_model_blob('solo8v2')
  vis = _abi.load_visual(blob)
  view, proj = _default_camera()
  img = ro.render(blob, vis, _standing_state(blob), view, proj, 123, 123, **_abi.RENDER_DEFAULTS)
  assert img.shape == (123, 123, 4) and img.dtype == np.uint8
  assert (img[..., 3] == 255).all()
  # sky above the horizon, checkerboard below it, grey robot in between
  assert (img[0, :, :3] == 255).all()
  grey = (img[..., 0] == img[..., 1]) & (img[..., 1] == img[..., 2]) & (img[..., 0] < 200)
  blue = img[..., 2] > img[..., 0].astype(int) + 20
  assert 300 < grey.sum() < 123 * 123 // 3
  assert blue.sum() > 1000
  rows = np.nonzero(grey.any(1))[0]
  # lifting the robot moves it up in the image; pulling the camera back shrinks it
  up = ro.render(blob, vis, _standing_state(blob, 0.6), view, proj, 123, 123, **_abi.RENDER_DEFAULTS)
  g2 = (up[..., 0] == up[..., 1]) & (up[..., 1] == up[..., 2]) & (up[..., 0] < 200)
  assert np.nonzero(g2.any(1))[0].mean() < rows.mean() - 10
  far_view = camera.compute_view_matrix_from_yaw_pitch_roll([0, 0, .2], 3, 0, -20, 0, 2)
  far = ro.render(blob, vis, _standing_state(blob), far_view, proj, 123, 123, **_abi.RENDER_DEFAULTS)
  g3 = (far[..., 0] == far[..., 1]) & (far[..., 1] == far[..., 2]) & (far[..., 0] < 200)
  assert g3.sum() < grey.sum() / 4
  # a robot behind the camera leaves only plane and sky
  st = _standing_state(blob)
  st[1] = -5.0
  gone = ro.render(blob, vis, st, view, proj, 123, 123, **_abi.RENDER_DEFAULTS)
  g4 = (gone[..., 0] == gone[..., 1]) & (gone[..., 1] == gone[..., 2]) & (gone[..., 0] < 200)
  assert g4.sum() == 0


def test_make_camera_validates():
  view, proj = _default_camera()
  cam, opts = _abi.make_camera(view, proj, 64, 48, ambient=0.5)
  assert cam.width == 64 and cam.height == 48 and cam.ambient == 0.5 and opts['diffuse'] == 0.35
  with pytest.raises(ValueError):
    _abi.make_camera(view[:15], proj, 64, 48)
  with pytest.raises(ValueError):
    _abi.make_camera(view, proj, 64, 48, shadows=True)
