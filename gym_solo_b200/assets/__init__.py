"""Compiled robot models (JSON blobs written by compile_model.py)."""
import hashlib
import json
import os

_HERE = os.path.dirname(os.path.abspath(__file__))

# reference URDF path (Solo8VanillaConfig.urdf_path, examples/solo8_mmr/*.py:16-17)
# -> compiled blob shipped with the package
_BLOBS = {
    'assets/solo8v2/solo.urdf': 'solo8v2.json',
    'assets/solo8_URDF_v4/solo8_URDF_v4.urdf': 'solo8_mmr_v4.json',
    'assets/solo8_URDF_v3/solo8_URDF_v3.urdf': 'solo8_urdf_v3.json',
}


def cache_dir():
  base = os.environ.get('GYM_SOLO_B200_CACHE') or os.path.join(
      os.environ.get('XDG_CACHE_HOME') or os.path.join(os.path.expanduser('~'), '.cache'), 'gym_solo_b200')
  os.makedirs(base, exist_ok=True)
  return base


def _compile_cached(path):
  """Any other URDF file (callers set config.urdf_path freely:
  examples/solo8_mmr/interactive_pos_control.py:15-18): compiled on first use into
  the cache directory, keyed on the contents of the URDF and of every file next to it
  (its meshes).  Topology limits are the kernels': a floating base with up to four
  serial chains of up to three revolute joints (compile_model raises otherwise)."""
  from gym_solo_b200.assets import compile_model as cm
  h = hashlib.sha1()
  d = os.path.dirname(os.path.abspath(path))
  h.update(open(path, 'rb').read())
  for f in sorted(os.listdir(d)):
    if f.lower().endswith('.stl'):
      st = os.stat(os.path.join(d, f))
      h.update(('%s:%d:%d' % (f, st.st_size, int(st.st_mtime))).encode())
  name = 'urdf_' + h.hexdigest()[:16]
  out = os.path.join(cache_dir(), name + '.json')
  if not os.path.exists(out):
    visual = {}
    try:
      model = cm.compile_model(os.path.abspath(path), name, visual=visual)
    except AssertionError as e:
      raise ValueError('URDF %r is outside the supported topology (floating base, <= 4 serial '
                       'chains of <= 3 revolute joints): %s' % (path, e))
    tmp = out + '.%d.tmp' % os.getpid()
    with open(os.path.join(cache_dir(), name + '_visual.json.%d.tmp' % os.getpid()), 'w') as f:
      json.dump(visual, f)
    os.replace(f.name, os.path.join(cache_dir(), name + '_visual.json'))
    with open(tmp, 'w') as f:
      json.dump(model, f)
    os.replace(tmp, out)
  return out


def blob_for_urdf(urdf_path):
  if urdf_path.endswith('.json'):
    return urdf_path
  name = _BLOBS.get(urdf_path)
  if name is not None:
    path = os.path.join(_HERE, name)
    if not os.path.exists(path):
      raise ValueError('compiled model %s is missing' % path)
    return path
  if os.path.isfile(urdf_path):
    return _compile_cached(urdf_path)
  raise ValueError('no compiled model for %r and no such URDF file' % urdf_path)
