"""A16: the compiled model against the only numbers of the real engine the reference holds
for it -- the 12 pybullet.getJointInfo tuples of gym_solo/core/test_obs_observations.py:123-162
(tests/golden/joint_info_golden.json, extracted by tools/gen_joint_info_golden.py).

They pin Bullet's URDF import conventions that the model compiler restates: joint index
order, q/u indices behind the floating base, fixed joints, limits, the link inertial frames
(COM + axes of the diagonalised inertia) and the joint frames relative to them.  The tuples
are derived from the constants the kernels simulate with (joint_origin / joint_rot /
joint_axis in body coordinates) plus the blob's per-link table, and the merge of the
fixed-joint children is checked against that table, so they pin the simulated model.
"""
import json
import os
import struct

import numpy as np
import pytest

from gym_solo_b200 import _abi
from gym_solo_b200.assets import blob_for_urdf
from gym_solo_b200.assets.joint_info import joint_info, merged_bodies

HERE = os.path.dirname(os.path.abspath(__file__))
TOL = 1e-9


@pytest.fixture(scope='module')
def golden():
  with open(os.path.join(HERE, 'golden', 'joint_info_golden.json')) as f:
    return json.load(f)['joint_info']


@pytest.fixture(scope='module')
def blob():
  return _abi.load_model_blob(blob_for_urdf('assets/solo8v2/solo.urdf'))


def test_all_twelve_tuples(golden, blob):
  ours = joint_info(blob)
  assert len(ours) == len(golden) == 12
  for got, want in zip(ours, golden):
    assert len(got) == len(want) == 17
    # index, name, type, qIndex, uIndex, flags: exact
    assert got[0] == want[0]
    assert got[1] == want[1].encode('UTF-8')
    assert got[2:6] == tuple(want[2:6])
    # damping, friction, lower, upper, max force, max velocity
    np.testing.assert_allclose(got[6:12], want[6:12], rtol=0, atol=TOL)
    assert got[12] == want[12].encode('UTF-8')
    # joint axis in the child inertial frame, parent frame position / orientation
    np.testing.assert_allclose(got[13], want[13], rtol=0, atol=TOL)
    np.testing.assert_allclose(got[14], want[14], rtol=0, atol=TOL)
    np.testing.assert_allclose(got[15], want[15], rtol=0, atol=TOL)
    assert got[16] == want[16]


def test_merged_bodies_are_the_sum_of_their_links(blob):
  """Fixed-joint children (the feet) are merged into their parent body: mass, COM and inertia
  of every simulated body must be the rigid sum of the links the tuples were checked on."""
  for b, (m, com, I) in enumerate(merged_bodies(blob)):
    np.testing.assert_allclose(m, blob['body_mass'][b], rtol=0, atol=1e-15)
    np.testing.assert_allclose(com, blob['body_com'][b], rtol=0, atol=1e-15)
    np.testing.assert_allclose(I, blob['body_inertia'][b], rtol=0, atol=1e-15)
  assert abs(sum(blob['body_mass']) - 2.17785) < 1e-5   # SURVEY App. A total mass


def test_struct_carries_the_blob(blob):
  """SoloModel (what solo_create receives) holds the same constants."""
  m = _abi.make_model(blob, link_friction=0.5)
  for d in range(blob['ndof']):
    np.testing.assert_array_equal(list(m.joint_origin[d]), blob['joint_origin'][d])
    np.testing.assert_array_equal(list(m.joint_axis[d]), blob['joint_axis'][d])
    np.testing.assert_array_equal(np.array(list(m.joint_rot[d])).reshape(3, 3), blob['joint_rot'][d])
  for b in range(blob['ndof'] + 1):
    assert m.body_mass[b] == blob['body_mass'][b]
    np.testing.assert_array_equal(list(m.body_com[b]), blob['body_com'][b])
  for j in range(12):
    assert (m.api_lower[j], m.api_upper[j]) == (-10.0, 10.0)   # solo.urdf:47


# ---- N3: any URDF path at construction (examples/solo8_mmr/interactive_pos_control.py:15-18) ----
def _write_box_stl(path, lo, hi):
  x0, y0, z0 = lo
  x1, y1, z1 = hi
  v = [(x0, y0, z0), (x1, y0, z0), (x1, y1, z0), (x0, y1, z0), (x0, y0, z1), (x1, y0, z1), (x1, y1, z1), (x0, y1, z1)]
  faces = [(0, 2, 1), (0, 3, 2), (4, 5, 6), (4, 6, 7), (0, 1, 5), (0, 5, 4), (2, 3, 7), (2, 7, 6),
           (1, 2, 6), (1, 6, 5), (0, 4, 7), (0, 7, 3)]
  with open(path, 'wb') as f:
    f.write(b'\0' * 80 + struct.pack('<I', len(faces)))
    for a, b, c in faces:
      f.write(struct.pack('<12fH', 0, 0, 0, *v[a], *v[b], *v[c], 0))


def _toy_urdf(d, hip_mass):
  _write_box_stl(os.path.join(d, 'torso.stl'), (-0.2, -0.1, -0.02), (0.2, 0.1, 0.02))
  _write_box_stl(os.path.join(d, 'leg.stl'), (-0.01, -0.01, -0.15), (0.01, 0.01, 0.0))
  def link(name, mass, mesh, iyz=0.0):
    return ('<link name="%s"><inertial><origin xyz="0 0 -0.05" rpy="0 0 0"/><mass value="%g"/>'
            '<inertia ixx="1e-3" ixy="0" ixz="0" iyy="1e-3" iyz="%g" izz="2e-4"/></inertial>'
            '<collision><geometry><mesh filename="%s"/></geometry></collision></link>' % (name, mass, iyz, mesh))
  parts = ['<robot name="toy">', link('torso', 1.0, 'torso.stl')]
  for i, (sx, sy) in enumerate([(1, 1), (1, -1), (-1, 1), (-1, -1)]):
    parts.append(link('hip%d' % i, hip_mass, 'leg.stl', 1e-5 * sy))
    parts.append(link('shin%d' % i, 0.05, 'leg.stl'))
    parts.append('<joint name="j%da" type="revolute"><parent link="torso"/><child link="hip%d"/>'
                 '<origin xyz="%g %g 0" rpy="0 0 0"/><axis xyz="0 1 0"/>'
                 '<limit lower="-3" upper="3" effort="10" velocity="20"/></joint>' % (i, i, 0.19 * sx, 0.1 * sy))
    parts.append('<joint name="j%db" type="revolute"><parent link="hip%d"/><child link="shin%d"/>'
                 '<origin xyz="0 0 -0.15" rpy="0 0 0"/><axis xyz="0 1 0"/>'
                 '<limit lower="-3" upper="3" effort="10" velocity="20"/></joint>' % (i, i, i))
  parts.append('</robot>')
  path = os.path.join(d, 'toy.urdf')
  with open(path, 'w') as f:
    f.write('\n'.join(parts))
  return path


def test_unknown_urdf_path_is_compiled_on_first_use(tmp_path, monkeypatch):
  monkeypatch.setenv('GYM_SOLO_B200_CACHE', str(tmp_path / 'cache'))
  from gym_solo_b200.core.configs import Solo8BaseConfig
  urdf = _toy_urdf(str(tmp_path), 0.2)
  cfg = Solo8BaseConfig()
  cfg.urdf_path = urdf
  path = cfg.urdf                       # what Solo8VanillaEnv._create_client loads
  blob = _abi.load_model_blob(path)
  assert blob['ndof'] == 8 and blob['n_api'] == 8 and blob['chain_len'] == [2, 2, 2, 2]
  assert blob['n_cand'] > 0 and abs(sum(blob['body_mass']) - (1.0 + 4 * 0.25)) < 1e-12
  m = _abi.make_model(blob, link_friction=0.5)
  assert m.ndof == 8 and (m.api_lower[0], m.api_upper[0]) == (-3.0, 3.0)
  info = joint_info(blob)
  assert info[1][16] == 0 and info[0][16] == -1 and info[0][3] == 7
  assert _abi.load_visual(blob)['n_tri'] > 0
  assert cfg.urdf == path               # second use: the cached blob
  # a changed file is a different model
  urdf2 = _toy_urdf(str(tmp_path), 0.3)
  assert cfg.urdf != path


def test_unsupported_topology_and_missing_file_raise(tmp_path, monkeypatch):
  monkeypatch.setenv('GYM_SOLO_B200_CACHE', str(tmp_path / 'cache'))
  with pytest.raises(ValueError):
    blob_for_urdf('assets/no_such_robot.urdf')
  urdf = _toy_urdf(str(tmp_path), 0.2)
  text = open(urdf).read().replace('type="revolute"', 'type="prismatic"', 1)
  bad = str(tmp_path / 'bad.urdf')
  open(bad, 'w').write(text)
  with pytest.raises(ValueError):
    blob_for_urdf(bad)


# ---- contact geometry: the support-point sets against the full STL hulls ---------------------
def test_support_point_sets_are_within_a_centimetre_of_the_hulls(blob):
  """Every body collides through a support-point set of its convex hull (compile_model.py:
  axis seeds + greedy epsilon-kernel; 8 rim points per foot).  The compiler records, per body,
  the largest distance by which the real hull reaches beyond the set over 20 000 directions
  (tools/support_error.py recomputes it from the meshes): round 1's box-corner sets were off by
  18 / 20 / 13 mm (torso / upper leg / lower leg + foot), 7 mm on the torso's underside."""
  err = blob['support_error']
  assert set(err) == set(blob['body_names'])
  assert max(err.values()) <= 1.4e-2 and err["base_link"] <= 4e-3, err
  assert blob['n_cand'] <= _abi.MAX_CAND
  # the torso rests on its true bottom face (z = -0.025 in the mesh, solo_body.stl), not 7 mm above it
  import numpy as np
  pts = np.array(blob['cand_pos'])[np.array(blob['cand_body']) == 0]
  assert abs(pts[:, 2].min() - (-0.025)) < 5e-4, pts[:, 2].min()
