"""pybullet.getJointInfo tuples from a compiled model blob.

Golden: gym_solo/core/test_obs_observations.py:123-162 holds the 12 tuples the real engine
returns for solo8v2 -- the one place the reference pins Bullet's URDF import conventions
(tests/test_joint_info_golden.py asserts them to 1e-9).  Bullet puts every link frame at the
link's inertial frame: COM, axes of the diagonalised inertia (URDF_USE_INERTIA_FROM_FILE).
With Rci / Rpi the inertial rotations of the child / parent link, c_p the parent COM and
(xyz, Rj) the joint frame in the parent link frame:
  [13] joint axis in the child inertial frame        Rci^T axis
  [14] joint position in the parent inertial frame   Rpi^T (xyz - c_p)
  [15] parent-to-child inertial rotation at q = 0    quat(Rci^T Rj^T Rpi)
  [16] parent link index (-1: base)
Everything is derived from the constants the kernels simulate with (joint_origin, joint_rot,
joint_axis, in body coordinates) and the blob's per-link table (`links`: the frame of every
URDF link inside its body, pre-merge mass / COM / inertia), so the golden tuples pin the
simulated model, not a side table.
"""
import numpy as np

from gym_solo_b200.assets.compile_model import mat_to_quat, principal_frame


def merged_bodies(blob):
  """(mass, com, inertia) of every dynamic body recomputed from its links: what
  compile_model's merge of fixed-joint children must equal (body_mass, body_com,
  body_inertia of the blob)."""
  out = []
  for b in range(blob['ndof'] + 1):
    parts = [L for L in blob['links'] if L['body'] == b]
    m = sum(L['mass'] for L in parts)
    cs = [np.asarray(L['pos']) + np.asarray(L['rot']) @ np.asarray(L['com']) for L in parts]
    com = sum(L['mass'] * c for L, c in zip(parts, cs)) / m
    I = np.zeros((3, 3))
    for L, c in zip(parts, cs):
      R, d = np.asarray(L['rot']), c - com
      I += R @ np.asarray(L['inertia']) @ R.T + L['mass'] * (d @ d * np.eye(3) - np.outer(d, d))
    out.append((m, com, I))
  return out


def joint_info(blob):
  """List of 17-tuples, one per pybullet joint index, shaped like getJointInfo's."""
  links = {L['name']: L for L in blob['links']}
  frames = {n: principal_frame(np.asarray(L['inertia']))[1] for n, L in links.items()}
  index_of_link = {j['child']: i for i, j in enumerate(blob['api_joints'])}
  out = []
  for i, j in enumerate(blob['api_joints']):
    d = blob['api_dof'][i]
    movable = d >= 0
    par, chi = links[j['parent']], links[j['child']]
    Rp, pp = np.asarray(par['rot']), np.asarray(par['pos'])
    if movable:
      # joint frame in the parent BODY's coordinates = what the kernels use
      o_body, R_body = np.asarray(blob['joint_origin'][d]), np.asarray(blob['joint_rot'][d])
      axis_link = np.asarray(blob['joint_axis'][d])
    else:
      # fixed joint: the child link sits in the same body as its parent
      o_body, R_body = np.asarray(chi['pos']), np.asarray(chi['rot'])
      axis_link = np.zeros(3)
    xyz = Rp.T @ (o_body - pp)            # joint frame in the parent LINK frame
    Rj = Rp.T @ R_body
    Rci, Rpi = frames[j['child']], frames[j['parent']]
    axis = Rci.T @ axis_link
    pos = Rpi.T @ (xyz - np.asarray(par['com']))
    orn = mat_to_quat(Rci.T @ Rj.T @ Rpi)
    out.append((i, j['name'].encode('UTF-8'), 0 if movable else 4,
                7 + d if movable else -1, 6 + d if movable else -1, 1 if movable else 0,
                float(j['damping']), float(j['friction']),
                float(blob['api_lower'][i]), float(blob['api_upper'][i]),
                float(j['effort']), float(j['velocity']), j['child'].encode('UTF-8'),
                tuple(float(v) + 0.0 for v in axis), tuple(float(v) + 0.0 for v in pos),
                tuple(orn), index_of_link.get(j['parent'], -1)))
  return out
