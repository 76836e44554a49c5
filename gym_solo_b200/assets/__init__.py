"""Compiled robot models (JSON blobs written by compile_model.py)."""
import os

_HERE = os.path.dirname(os.path.abspath(__file__))

# reference URDF path (Solo8VanillaConfig.urdf_path, examples/solo8_mmr/*.py:16-17)
# -> compiled blob
_BLOBS = {
    'assets/solo8v2/solo.urdf': 'solo8v2.json',
    'assets/solo8_URDF_v4/solo8_URDF_v4.urdf': 'solo8_mmr_v4.json',
    'assets/solo8_URDF_v3/solo8_URDF_v3.urdf': 'solo8_urdf_v3.json',
}


def blob_for_urdf(urdf_path):
  if urdf_path.endswith('.json'):
    return urdf_path
  name = _BLOBS.get(urdf_path)
  if name is None:
    raise ValueError('no compiled model for %r; run python -m gym_solo_b200.assets.compile_model' % urdf_path)
  path = os.path.join(_HERE, name)
  if not os.path.exists(path):
    raise ValueError('compiled model %s is missing' % path)
  return path
