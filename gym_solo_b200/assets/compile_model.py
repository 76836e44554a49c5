"""URDF + STL -> compiled Solo model blob (JSON).

Offline tool.  Reads a Solo URDF and its collision
meshes and writes the constant model blob that the CUDA path and the CPU oracle
both consume through ``SoloModel`` (include/solo8_b200.h).  The blob is
committed; the URDF/STL sources are not (they stay in the reference tree:
gym_solo/assets/solo8v2/solo.urdf:8-583, gym_solo/assets/solo8_URDF_v4/*.urdf).

What it mirrors of pybullet's URDF import (Bullet recollection):
  * link/joint flattening order = depth-first from the base, children in file
    order -> pybullet joint index (golden: test_obs_observations.py:237-239);
  * ``URDF_USE_INERTIA_FROM_FILE``: mass / COM / inertia tensor from <inertial>;
    the *base* frame that the getters report is the base inertial frame (COM,
    principal axes), so the base body is re-expressed in that frame;
  * fixed joints carry no dof: the child is merged into its dynamic parent
    (rigidly equivalent); its API slot stays (12-wide API, reads 0.0);
  * collision mesh -> convex hull, 1 mm outward margin (btConvexHullShape).

Collision reduction (ours): the ground is the plane z = 0 and a convex hull touches
a plane at its support point in the -n direction, so every dynamic body (a link plus
the links fixed to it) is reduced to a support-point SET: the hull vertices that are
the support points along -z and +z of each of its links (the foot's lowest point gives
the 0.33698 standing height of the examples), then greedily the hull vertex that is furthest
outside the set's own hull, until the body's budget is used (an epsilon-kernel for
the support function; tools/support_error.py reports the error left).  Budgets:
torso 24 (the robot rests on it after every reset: 3.5 mm, was 18 mm with 7 mm on the
underside), 8 per leg body + 4 where a foot is attached (its 8 rim points) -> 104
points for Solo 8 (the kernels' limit is 128).  More points on the legs (14 each:
6-9 mm instead of 13) were measured and not kept: 40 % more speculative contact
rows, 2 % of the environments above 32 contacts, solver 9-11 % slower.

usage: python -m gym_solo_b200.assets.compile_model <urdf> <out.json> [--name N]
"""
import json
import math
import os
import struct
import sys
import xml.etree.ElementTree as ET

import numpy as np

MAX_LEGS = 4
MAX_CHAIN = 3
MAX_CAND = 128     # kMaxCand of the kernels


def rpy_to_mat(r, p, y):
  cr, sr, cp, sp, cy, sy = (math.cos(r), math.sin(r), math.cos(p), math.sin(p),
                            math.cos(y), math.sin(y))
  return np.array([[cy * cp, cy * sp * sr - sy * cr, cy * sp * cr + sy * sr],
                   [sy * cp, sy * sp * sr + cy * cr, sy * sp * cr - cy * sr],
                   [-sp, cp * sr, cp * cr]])


def read_stl(path):
  d = open(path, 'rb').read()
  n = struct.unpack('<I', d[80:84])[0]
  if 84 + 50 * n != len(d):
    raise ValueError('only binary STL supported: ' + path)
  rec = np.dtype([('n', '<f4', 3), ('v', '<f4', (3, 3)), ('a', '<u2')])
  a = np.frombuffer(d, dtype=rec, count=n, offset=84)
  return np.unique(a['v'].reshape(-1, 3).astype(np.float64), axis=0)


def _floats(s, n):
  v = [float(x) for x in s.split()]
  assert len(v) == n, s
  return v


def _origin(el):
  o = el.find('origin') if el is not None else None
  if o is None:
    return np.zeros(3), np.eye(3)
  xyz = np.array(_floats(o.get('xyz', '0 0 0'), 3))
  rpy = _floats(o.get('rpy', '0 0 0'), 3)
  return xyz, rpy_to_mat(*rpy)


def parse_urdf(path):
  root = ET.parse(path).getroot()
  links, joints = {}, []
  for l in root.findall('link'):
    inert = l.find('inertial')
    com, Rin = _origin(inert)
    mass = float(inert.find('mass').get('value'))
    i = inert.find('inertia')
    g = lambda k: float(i.get(k, '0'))
    I = np.array([[g('ixx'), g('ixy'), g('ixz')],
                  [g('ixy'), g('iyy'), g('iyz')],
                  [g('ixz'), g('iyz'), g('izz')]])
    I = Rin @ I @ Rin.T          # inertia in link-frame axes
    vis = l.find('visual')
    rgba = [0.8, 0.8, 0.8, 1.0]
    if vis is not None and vis.find('material') is not None and vis.find('material').find('color') is not None:
      rgba = _floats(vis.find('material').find('color').get('rgba'), 4)
    col = l.find('collision')
    mesh, cxyz, cR = None, np.zeros(3), np.eye(3)
    if col is not None:
      cxyz, cR = _origin(col)
      m = col.find('geometry').find('mesh')
      if m is not None:
        mesh = m.get('filename')
    c = l.find('contact')
    mu = 0.5  # Bullet's default lateral friction when no <contact> tag
    if c is not None and c.find('lateral_friction') is not None:
      mu = float(c.find('lateral_friction').get('value'))
    links[l.get('name')] = dict(name=l.get('name'), mass=mass, com=com, I=I,
                                mesh=mesh, cxyz=cxyz, cR=cR, mu=mu, rgba=rgba)
  for j in root.findall('joint'):
    xyz, R = _origin(j)
    ax = j.find('axis')
    axis = np.array(_floats(ax.get('xyz'), 3)) if ax is not None else np.array([1., 0, 0])
    lim = j.find('limit')
    lo = float(lim.get('lower', '0')) if lim is not None else 0.
    hi = float(lim.get('upper', '0')) if lim is not None else 0.
    dyn = j.find('dynamics')
    joints.append(dict(name=j.get('name'), type=j.get('type'),
                       parent=j.find('parent').get('link'),
                       child=j.find('child').get('link'),
                       xyz=xyz, R=R, axis=axis, lower=lo, upper=hi,
                       effort=float(lim.get('effort', '0')) if lim is not None else 0.,
                       velocity=float(lim.get('velocity', '0')) if lim is not None else 0.,
                       damping=float(dyn.get('damping', '0')) if dyn is not None else 0.,
                       friction=float(dyn.get('friction', '0')) if dyn is not None else 0.))
  children = {n: [] for n in links}
  has_parent = set()
  for j in joints:
    children[j['parent']].append(j)
    has_parent.add(j['child'])
  base = [n for n in links if n not in has_parent]
  assert len(base) == 1
  return links, joints, children, base[0]


def _greedy_dirs():
  rng = np.random.RandomState(20260)
  d = rng.standard_normal((20000, 3))
  return d / np.linalg.norm(d, axis=1, keepdims=True)


_DIRS = None


def support_set(full, seeds, budget):
  """Greedy support-point set of a convex point cloud `full` ([n, 3], body coordinates):
  starts from `seeds` (indices into full) and keeps adding the vertex that realises the largest
  support-function error h_full(d) - h_set(d) over 20 000 fixed directions."""
  global _DIRS
  if _DIRS is None:
    _DIRS = _greedy_dirs()
  hf = full @ _DIRS.T
  hfull = hf.max(0)
  sel = []
  for i in seeds:
    if i not in sel and len(sel) < budget:
      sel.append(int(i))
  hset = hf[sel].max(0) if sel else np.full(len(_DIRS), -1e30)
  while len(sel) < budget:
    d = int((hfull - hset).argmax())
    if hfull[d] - hset[d] <= 1e-9:
      break
    v = int(hf[:, d].argmax())
    sel.append(v)
    hset = np.maximum(hset, hf[v])
  return sel, float((hfull - hset).max())


def fibonacci_dirs(n):
  """n roughly uniform unit directions."""
  k = np.arange(n) + 0.5
  z = 1 - 2 * k / n
  r = np.sqrt(1 - z * z)
  th = math.pi * (1 + 5 ** 0.5) * k
  return np.stack([r * np.cos(th), r * np.sin(th), z], 1)


def visual_hull(hv, n_dirs):
  """Triangles (outward, counter-clockwise) of the hull of the support points of
  `hv` in n_dirs directions plus the six axis directions: a decimated convex
  stand-in for the link mesh that render() rasterises."""
  from scipy.spatial import ConvexHull
  dirs = np.concatenate([fibonacci_dirs(n_dirs), np.eye(3), -np.eye(3)])
  ext = np.maximum(hv.max(0) - hv.min(0), 1e-9)
  pts = np.unique(hv[((hv / ext) @ dirs.T).argmax(0)], axis=0)
  hull = ConvexHull(pts)
  tris = []
  for simplex, eq in zip(hull.simplices, hull.equations):
    a, b, c = pts[simplex]
    if np.cross(b - a, c - a) @ eq[:3] < 0:
      b, c = c, b
    tris.append(np.concatenate([a, b, c]))
  return np.array(tris)


def principal_frame(I):
  """Bullet diagonalises a non-diagonal inertia; returns (diag, R) with
  I = R diag R^T and R the rotation closest to identity (axes not sorted)."""
  off = abs(I[0, 1]) + abs(I[0, 2]) + abs(I[1, 2])
  if off == 0.0:
    return np.diag(I).copy(), np.eye(3)
  w, V = np.linalg.eigh(I)
  # assign eigenvectors to the axis they are most aligned with
  cols = []
  remaining = list(range(3))
  for axis in range(3):
    k = max(remaining, key=lambda c: abs(V[axis, c]))
    remaining.remove(k)
    v = V[:, k] * (1 if V[axis, k] >= 0 else -1)
    cols.append((w[k], v))
  R = np.stack([c[1] for c in cols], axis=1)
  if np.linalg.det(R) < 0:
    R[:, 2] *= -1
  return np.array([c[0] for c in cols]), R


def mat_to_quat(R):
  """Rotation matrix -> (x, y, z, w), w >= 0."""
  t = R[0, 0] + R[1, 1] + R[2, 2]
  if t > 0:
    s = math.sqrt(t + 1.0) * 2
    q = [(R[2, 1] - R[1, 2]) / s, (R[0, 2] - R[2, 0]) / s, (R[1, 0] - R[0, 1]) / s, 0.25 * s]
  else:
    i = int(np.argmax([R[0, 0], R[1, 1], R[2, 2]]))
    j, k = (i + 1) % 3, (i + 2) % 3
    s = math.sqrt(R[i, i] - R[j, j] - R[k, k] + 1.0) * 2
    q = [0.0, 0.0, 0.0, (R[k, j] - R[j, k]) / s]
    q[i] = 0.25 * s
    q[j] = (R[j, i] + R[i, j]) / s
    q[k] = (R[k, i] + R[i, k]) / s
  if q[3] < 0:
    q = [-v for v in q]
  return [float(v) + 0.0 for v in q]


def compile_model(urdf_path, name=None, fallback_boxes=None, visual=None, visual_dirs=40,
                  budget_base=24, budget_other=8, budget_foot_extra=4):
  """`visual`: optional dict that receives the render() mesh (tri_body, tri_verts
  in body coordinates, tri_rgba): the decimated convex hull of every link's
  collision mesh (the reference's URDFs use the same STL for <visual> and
  <collision>), or the fallback box where the STL is absent."""
  links, joints, children, base = parse_urdf(urdf_path)
  mesh_dir = os.path.dirname(urdf_path)
  fallback_boxes = fallback_boxes or {}

  # ---- flatten (DFS, file order) -> pybullet joint indices -------------------
  api = []          # list of joints in pybullet index order

  def dfs(link):
    for j in children[link]:
      api.append(j)
      dfs(j['child'])
  dfs(base)

  # ---- base body in its inertial (COM, principal) frame ----------------------
  bl = links[base]
  Ib_diag, Rb = principal_frame(bl['I'])
  # X_base maps URDF-base-link coords -> base-body coords: p_body = Rb^T (p - com)
  base_com = bl['com'].copy()

  bodies = []       # dynamic bodies: dict(mass, com, I, parts=[(link, p, R)])
  # each "part" = original link rigidly attached: offset p, rotation R of the
  # link frame expressed in the body frame
  bodies.append(dict(name=base, mass=0., com=np.zeros(3), I=np.zeros((3, 3)),
                     parts=[]))

  legs = []         # list of lists of dof dicts

  def attach(body, link, p, R):
    """Rigidly add `link` (frame at p,R in body coords) to `body`."""
    body['parts'].append((link, p, R))
    for j in children[link]:
      if j['type'] == 'fixed':
        attach(body, j['child'], p + R @ j['xyz'], R @ j['R'])

  def merge_inertia(body):
    m = sum(links[l]['mass'] for l, _, _ in body['parts'])
    com = sum(links[l]['mass'] * (p + R @ links[l]['com'])
              for l, p, R in body['parts']) / m
    I = np.zeros((3, 3))
    for l, p, R in body['parts']:
      L = links[l]
      d = p + R @ L['com'] - com
      I += R @ L['I'] @ R.T + L['mass'] * (d @ d * np.eye(3) - np.outer(d, d))
    body.update(mass=m, com=com, I=I)

  attach(bodies[0], base, -Rb.T @ base_com, Rb.T)
  merge_inertia(bodies[0])

  dof_of_joint = {}

  def walk_chain(parent_body_idx, link, p, R, chain):
    """Follow movable joints below `link` (frame p,R in parent body coords)."""
    for j in children[link]:
      jp, jR = p + R @ j['xyz'], R @ j['R']
      if j['type'] == 'fixed':
        walk_chain(parent_body_idx, j['child'], jp, jR, chain)
        continue
      if j['type'] not in ('revolute', 'continuous'):
        raise ValueError('unsupported joint type ' + j['type'])
      if parent_body_idx == 0:
        chain = []
        legs.append(chain)
      body = dict(name=j['child'], mass=0., com=np.zeros(3), I=np.zeros((3, 3)),
                  parts=[])
      attach(body, j['child'], np.zeros(3), np.eye(3))
      merge_inertia(body)
      bodies.append(body)
      idx = len(bodies) - 1
      ax = j['axis'] / np.linalg.norm(j['axis'])
      chain.append(dict(joint=j, body=idx, parent=parent_body_idx,
                        origin=jp, rot=jR, axis=ax))
      dof_of_joint[j['name']] = idx - 1
      walk_chain(idx, j['child'], np.zeros(3), np.eye(3), chain)

  walk_chain(0, base, -Rb.T @ base_com, Rb.T, None)
  assert len(legs) <= MAX_LEGS and all(len(c) <= MAX_CHAIN for c in legs)
  # dof order must be chain-contiguous
  d = 0
  for c in legs:
    for e in c:
      assert e['body'] == d + 1, 'dof order is not chain-contiguous'
      d += 1
  ndof = d

  # ---- collision support points ---------------------------------------------
  cand_body, cand_pos, cand_mu, cand_link = [], [], [], []
  hull_cache = {}
  tri_body, tri_verts, tri_rgba = [], [], []
  body_cloud = []      # per body: (points [n, 3] in body coordinates, link of every point, seeds)
  for bi, body in enumerate(bodies):
    pts_b, link_b, seeds = [], [], []
    for l, p, R in body['parts']:
      L = links[l]
      hv = None
      if L['mesh'] is not None:
        mp = os.path.join(mesh_dir, L['mesh'])
        if os.path.exists(mp):
          if mp not in hull_cache:
            from scipy.spatial import ConvexHull
            v = read_stl(mp)
            hull_cache[mp] = v[ConvexHull(v).vertices]
          hv = hull_cache[mp]
      if hv is None and l in fallback_boxes:
        lo, hi = fallback_boxes[l]
        hv = np.array([[x, y, z] for x in (lo[0], hi[0]) for y in (lo[1], hi[1])
                       for z in (lo[2], hi[2])])
      if hv is None:
        continue
      if visual is not None:
        for t in visual_hull(hv, visual_dirs):
          w = [p + R @ (L['cxyz'] + L['cR'] @ t[3 * k:3 * k + 3]) for k in range(3)]
          tri_body.append(bi)
          tri_verts.append(np.concatenate(w).round(6).tolist())
          tri_rgba.append(L['rgba'])
      if 'foot' in l.lower():
        # the foot is a wheel-like disc (axis y): 8 points of its rim at the mid plane stand in for
        # the rolling contact line (the lowest one gives the examples' 0.33698 standing height and a
        # contact on the foot's centre line: the straight-legged stand comes to rest to 1e-9)
        lo_, hi_ = hv.min(0), hv.max(0)
        mid = 0.5 * (lo_ + hi_)
        rim = []
        for k8 in range(8):
          th = 2 * math.pi * k8 / 8
          d = np.array([math.cos(th), 0., math.sin(th)])
          rim.append(mid + ((hv - mid) @ d).max() * d)
        hv = np.concatenate([np.array(rim), hv])
      in_body = p + (R @ (L['cxyz'][:, None] + L['cR'] @ hv.T)).T
      base_index = sum(len(x) for x in pts_b)
      if 'foot' in l.lower():
        seeds.extend(range(base_index, base_index + 8))
      # seeds: the link's support points along its own -z / +z (link frame = columns of R): the
      # exact lowest point of a link that stands or lies in its reference orientation
      if 'foot' not in l.lower():
        for sgn in (-1.0, 1.0):
          seeds.append(base_index + int((in_body @ (sgn * R[:, 2])).argmax()))
      pts_b.append(in_body)
      link_b.extend([l] * len(hv))
    body_cloud.append((np.concatenate(pts_b) if pts_b else np.zeros((0, 3)), link_b, seeds))
  meshed = [bi for bi, c in enumerate(body_cloud) if len(c[0])]
  budgets = {}
  left = MAX_CAND
  if 0 in meshed:
    budgets[0] = budget_base
    left -= budget_base
  others = [bi for bi in meshed if bi != 0]
  for bi in others:
    has_foot = any('foot' in l.lower() for l, _, _ in bodies[bi]['parts'])
    budgets[bi] = min(budget_other + (budget_foot_extra if has_foot else 0), left // max(1, len(others)))
  support_error = {}
  for bi in meshed:
    cloud, link_b, seeds = body_cloud[bi]
    sel, err = support_set(cloud, seeds, budgets[bi])
    support_error[bodies[bi]['name']] = err
    for v in sel:
      cand_body.append(bi)
      cand_pos.append([float(x) for x in cloud[v]])
      cand_mu.append(links[link_b[v]]['mu'])
      cand_link.append(link_b[v])

  model = dict(
      name=name or os.path.splitext(os.path.basename(urdf_path))[0],
      source=os.path.relpath(urdf_path, '/root/reference'),
      n_legs=len(legs), chain_len=[len(c) for c in legs], ndof=ndof,
      n_api=len(api), api_names=[j['name'] for j in api],
      api_dof=[dof_of_joint.get(j['name'], -1) for j in api],
      api_lower=[j['lower'] for j in api], api_upper=[j['upper'] for j in api],
      body_names=[b['name'] for b in bodies],
      body_mass=[b['mass'] for b in bodies],
      body_com=[b['com'].tolist() for b in bodies],
      body_inertia=[b['I'].tolist() for b in bodies],
      joint_parent=[e['parent'] for c in legs for e in c],
      joint_origin=[e['origin'].tolist() for c in legs for e in c],
      joint_rot=[e['rot'].tolist() for c in legs for e in c],
      joint_axis=[e['axis'].tolist() for c in legs for e in c],
      base_inertial_pos=base_com.tolist(), base_inertial_rot=Rb.tolist(),
      n_cand=len(cand_body), cand_body=cand_body, cand_pos=cand_pos,
      cand_mu=cand_mu, cand_link=cand_link, collision_margin=0.001,
      # max over 20 000 directions of h_hull(d) - h_points(d) per body, metres
      support_error=support_error,
      # pre-merge data, for getJointInfo (assets/joint_info.py): every URDF link with the body it
      # belongs to, its frame in that body's coordinates and its own inertial parameters
      links=[dict(name=l, body=bi, pos=(p + 0.0).tolist(), rot=(R + 0.0).tolist(), mass=links[l]['mass'],
                  com=links[l]['com'].tolist(), inertia=links[l]['I'].tolist())
             for bi, body in enumerate(bodies) for l, p, R in body['parts']],
      api_joints=[dict(name=j['name'], type=j['type'], parent=j['parent'], child=j['child'],
                       effort=j['effort'], velocity=j['velocity'], damping=j['damping'],
                       friction=j['friction']) for j in api],
  )
  if visual is not None:
    visual.update(name=model['name'], source=model['source'], n_tri=len(tri_body),
                  tri_body=tri_body, tri_verts=tri_verts, tri_rgba=tri_rgba)
  return model


def main(argv):
  urdf, out = argv[1], argv[2]
  name = argv[argv.index('--name') + 1] if '--name' in argv else None
  m = compile_model(urdf, name)
  with open(out, 'w') as f:
    json.dump(m, f, indent=1)
  print('wrote', out, 'ndof', m['ndof'], 'bodies', len(m['body_mass']),
        'cands', m['n_cand'], 'mass', sum(m['body_mass']))


if __name__ == '__main__':
  main(sys.argv)
