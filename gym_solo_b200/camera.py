"""Camera matrices of render() -- the two pybullet helpers the reference calls
(gym_solo/envs/solo8_base_env.py:129-135), restated from Bullet's published
definitions (b3ComputeProjectionMatrixFOV / b3ComputeViewMatrixFromYawPitchRoll,
pybullet is absent from /root/reference).  Both return 16 floats, column-major,
like pybullet does.
"""
import math

import numpy as np


def compute_projection_matrix_fov(fov, aspect, near_val, far_val):
  """OpenGL perspective frustum: fov is the vertical field of view in degrees."""
  y_scale = 1.0 / math.tan(math.radians(fov) / 2.0)
  x_scale = y_scale / aspect
  n, f = float(near_val), float(far_val)
  return (x_scale, 0., 0., 0.,
          0., y_scale, 0., 0.,
          0., 0., (n + f) / (n - f), -1.,
          0., 0., 2. * n * f / (n - f), 0.)


def compute_view_matrix(eye, target, up):
  """Right-handed look-at (b3ComputeViewMatrixFromPositions)."""
  eye, target, up = (np.asarray(v, dtype=np.float64) for v in (eye, target, up))
  f = target - eye
  f = f / np.linalg.norm(f)
  u = up / np.linalg.norm(up)
  s = np.cross(f, u)
  s = s / np.linalg.norm(s)
  u = np.cross(s, f)
  return (s[0], u[0], -f[0], 0.,
          s[1], u[1], -f[1], 0.,
          s[2], u[2], -f[2], 0.,
          -float(s @ eye), -float(u @ eye), float(f @ eye), 1.)


def _rot(axis, a):
  c, s = math.cos(a), math.sin(a)
  if axis == 0:
    return np.array([[1, 0, 0], [0, c, -s], [0, s, c]], dtype=np.float64)
  if axis == 1:
    return np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]], dtype=np.float64)
  return np.array([[c, -s, 0], [s, c, 0], [0, 0, 1]], dtype=np.float64)


def compute_view_matrix_from_yaw_pitch_roll(target, distance, yaw, pitch, roll, up_axis_index):
  """Camera orbiting `target` at `distance`; angles in degrees; up axis 1 (Y) or
  2 (Z).  With Z up the camera starts at -distance along Y looking along +Y and
  is rotated by Rz(yaw) Ry(roll) Rx(pitch); with Y up it starts at -distance
  along Z and is rotated by Rz(roll) Ry(yaw) Rx(-pitch)."""
  yaw, pitch, roll = (math.radians(v) for v in (yaw, pitch, roll))
  if up_axis_index == 1:
    forward, up = 2, np.array([0., 1., 0.])
    rot = _rot(2, roll) @ _rot(1, yaw) @ _rot(0, -pitch)
  elif up_axis_index == 2:
    forward, up = 1, np.array([0., 0., 1.])
    rot = _rot(2, yaw) @ _rot(1, roll) @ _rot(0, pitch)
  else:
    raise ValueError('up_axis_index must be 1 or 2')
  eye = np.zeros(3)
  eye[forward] = -float(distance)
  eye = rot @ eye
  up = rot @ up
  target = np.asarray(target, dtype=np.float64)
  return compute_view_matrix(eye + target, target, up)
