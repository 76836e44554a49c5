"""CPU restatement of render() -- TEST INFRASTRUCTURE ONLY (tests/, smoke()).

What it restates: gym_solo/envs/solo8_base_env.py:116-142 asks pybullet for
getCameraImage(width, height, view, proj) and reshapes the RGBA bytes to
(h, w, 4).  pybullet (and its TinyRenderer) is absent from /root/reference and
from this image, and no graphics stack exists on either box, so there is nothing
to pin the pixels against: **parity unpinned**.  This file fixes the image
definition the CUDA rasteriser is tested against, in float64 numpy:

  * link triangles (body coordinates) -> world by forward kinematics of the state
    (base pose, then origin / fixed rotation / axis-angle per joint);
  * clip = proj * view * (x, y, z, 1); pixel x = (ndc_x / 2 + 1/2) W,
    y = (1/2 - ndc_y / 2) H; samples at pixel centres; a pixel is inside a
    triangle when its three edge functions are >= 0; depth = ndc z, nearest wins,
    the earlier triangle on a tie; back faces (clockwise seen from outside) culled;
  * colour = rgb * (ambient + diffuse * max(0, n . l)), flat per triangle,
    rounded to uint8 as floor(255 c + 0.5), alpha 255;
  * ground plane z = 0: ray through the pixel centre, checkerboard of
    checker_size cells, parity of floor(x / s) + floor(y / s);
  * background colour elsewhere.
"""
import numpy as np


def _quat_to_mat(q):
  x, y, z, w = q
  s = 2.0 / (x * x + y * y + z * z + w * w)
  return np.array([[1 - s * (y * y + z * z), s * (x * y - w * z), s * (x * z + w * y)],
                   [s * (x * y + w * z), 1 - s * (x * x + z * z), s * (y * z - w * x)],
                   [s * (x * z - w * y), s * (y * z + w * x), 1 - s * (x * x + y * y)]])


def _axis_angle(a, th):
  a = np.asarray(a, dtype=np.float64)
  K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
  return np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * (K @ K)


def body_poses(blob, state):
  """[(R, o)] per body: body 0 = base (state pos / quat), body d + 1 = link of dof d."""
  state = np.asarray(state, dtype=np.float64)
  ndof = blob['ndof']
  q = state[13:13 + ndof]
  poses = [(_quat_to_mat(state[3:7]), state[0:3].copy())]
  for d in range(ndof):
    Rp, op = poses[blob['joint_parent'][d]]
    o = op + Rp @ np.asarray(blob['joint_origin'][d])
    R = Rp @ np.asarray(blob['joint_rot'][d]).reshape(3, 3) @ _axis_angle(blob['joint_axis'][d], q[d])
    poses.append((R, o))
  return poses


def _u8(c):
  return np.floor(np.clip(c, 0.0, 1.0) * 255.0 + 0.5).astype(np.uint8)


def render(blob, visual, state, view, proj, width, height, light_direction, ambient, diffuse,
           background, plane_rgb_a, plane_rgb_b, checker_size, aux=False):
  """aux=True also returns getCameraImage's depth buffer (0.5 ndc_z + 0.5, 1 where nothing was
  hit) and segmentation mask (plane 0, robot 1, nothing -1)."""
  W, H = int(width), int(height)
  V = np.asarray(view, dtype=np.float64).reshape(4, 4).T     # column-major input
  P = np.asarray(proj, dtype=np.float64).reshape(4, 4).T
  VP = P @ V
  light = np.asarray(light_direction, dtype=np.float64)
  light = light / np.linalg.norm(light)
  img = np.empty((H, W, 4), dtype=np.uint8)
  img[..., :3] = _u8(np.asarray(background, dtype=np.float64))
  img[..., 3] = 255
  depth = np.full((H, W), 1.0)
  seg = np.full((H, W), -1, dtype=np.int32)

  poses = body_poses(blob, state)
  xs = np.arange(W) + 0.5
  ys = np.arange(H) + 0.5
  for b, verts, rgba in zip(visual['tri_body'], visual['tri_verts'], visual['tri_rgba']):
    R, o = poses[b]
    w = o + (R @ np.asarray(verts, dtype=np.float64).reshape(3, 3).T).T       # [3 vertices][xyz]
    clip = (VP @ np.concatenate([w, np.ones((3, 1))], 1).T).T
    if np.any(clip[:, 3] <= 1e-6):
      continue
    ndc = clip[:, :3] / clip[:, 3:4]
    sx = (ndc[:, 0] * 0.5 + 0.5) * W
    sy = (0.5 - ndc[:, 1] * 0.5) * H
    sz = ndc[:, 2]
    area = (sx[1] - sx[0]) * (sy[2] - sy[0]) - (sx[2] - sx[0]) * (sy[1] - sy[0])
    if not area < -1e-12:
      continue
    n = np.cross(w[1] - w[0], w[2] - w[0])
    n = n / np.linalg.norm(n)
    colour = _u8(np.asarray(rgba[:3]) * (ambient + diffuse * max(0.0, float(n @ light))))
    # swap v1 <-> v2: positive area with y pointing down
    X, Y, Z = sx[[0, 2, 1]], sy[[0, 2, 1]], sz[[0, 2, 1]]
    x0, x1 = int(max(0, np.floor(X.min() - 0.5))), int(min(W, np.ceil(X.max() + 0.5)))
    y0, y1 = int(max(0, np.floor(Y.min() - 0.5))), int(min(H, np.ceil(Y.max() + 0.5)))
    if x0 >= x1 or y0 >= y1:
      continue
    fx, fy = np.meshgrid(xs[x0:x1], ys[y0:y1])
    e = lambda a, c: (X[c] - X[a]) * (fy - Y[a]) - (Y[c] - Y[a]) * (fx - X[a])
    w0, w1, w2 = e(1, 2), e(2, 0), e(0, 1)
    z = (w0 * Z[0] + w1 * Z[1] + w2 * Z[2]) / -area
    sub = depth[y0:y1, x0:x1]
    hit = (w0 >= 0) & (w1 >= 0) & (w2 >= 0) & (z < sub) & (z >= -1.0)
    sub[hit] = z[hit]
    img[y0:y1, x0:x1, :3][hit] = colour
    seg[y0:y1, x0:x1][hit] = 1

  # ground plane
  Rv, tv = V[:3, :3], V[:3, 3]
  eye = -Rv.T @ tv
  nx = 2.0 * xs / W - 1.0
  ny = 1.0 - 2.0 * ys / H
  NX, NY = np.meshgrid(nx, ny)
  ray = (-Rv[2])[None, None, :] + NX[..., None] * (Rv[0] / P[0, 0]) + NY[..., None] * (Rv[1] / P[1, 1])
  A, Bz = P[2, 2], P[2, 3]
  near, far = Bz / (A - 1.0), Bz / (A + 1.0)
  with np.errstate(divide='ignore', invalid='ignore'):
    t = -eye[2] / ray[..., 2]
    z = -A + Bz / t
  hit = (ray[..., 2] != 0) & (t > near) & (t < far) & (z < depth)
  hx = eye[0] + t * ray[..., 0]
  hy = eye[1] + t * ray[..., 1]
  with np.errstate(invalid='ignore'):
    cell = (np.floor(hx / checker_size) + np.floor(hy / checker_size))
  odd = np.mod(np.where(hit, cell, 0.0), 2.0) == 1.0
  li = ambient + diffuse * max(0.0, float(light[2]))
  ca, cb = _u8(np.asarray(plane_rgb_a) * li), _u8(np.asarray(plane_rgb_b) * li)
  img[..., :3][hit & ~odd] = ca
  img[..., :3][hit & odd] = cb
  if not aux:
    return img
  depth = np.where(hit, z, depth)
  seg[hit] = 0
  return img, (0.5 * depth + 0.5).astype(np.float32), seg
