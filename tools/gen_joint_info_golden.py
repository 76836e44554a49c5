"""Extract the 12 golden pybullet.getJointInfo tuples the reference's own test holds
(gym_solo/core/test_obs_observations.py:123-162, "obtained from getJointInfo (pybullet) for
all the joints") into tests/golden/joint_info_golden.json.

The tuples are read from the unmodified reference file with `ast` (nothing is imported or
executed).  Runs only in the authoring container (needs /root/reference).

usage: python tools/gen_joint_info_golden.py
"""
import ast
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = '/root/reference/gym_solo/core/test_obs_observations.py'


def main():
  tree = ast.parse(open(SRC).read())
  found = None
  for node in ast.walk(tree):
    if (isinstance(node, ast.Assign) and len(node.targets) == 1
        and isinstance(node.targets[0], ast.Attribute) and node.targets[0].attr == 'joint_info'):
      found = (node.lineno, node.end_lineno, ast.literal_eval(node.value))
  assert found, 'joint_info literal not found'
  lo, hi, tuples = found

  def plain(v):
    if isinstance(v, bytes):
      return v.decode('UTF-8')
    if isinstance(v, tuple):
      return [plain(x) for x in v]
    return v
  out = dict(source='gym_solo/core/test_obs_observations.py:%d-%d' % (lo, hi),
             note='bytes fields ([1] joint name, [12] link name) stored as str',
             joint_info=[plain(t) for t in tuples])
  path = os.path.join(ROOT, 'tests', 'golden', 'joint_info_golden.json')
  with open(path, 'w') as f:
    json.dump(out, f, indent=1)
  print('wrote', path, len(tuples), 'tuples from', out['source'])


if __name__ == '__main__':
  main()
