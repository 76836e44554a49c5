"""Contact geometry: how far is the reduced support-point set of every link from the link's full
STL convex hull?  For a plane the contact point of a convex hull is its support point in the
direction of the plane normal, so the error of the reduction in a direction d is
  h_full(d) - h_reduced(d) >= 0   (metres; the reduced set is a subset of the hull),
i.e. how far the real hull would reach below the lowest retained point.  Reported per BODY (a link plus
the links fixed to it) as the maximum, the 99th percentile and the mean over 10^4 uniformly random
directions, and the maximum over the directions within 30 degrees of the body's -z (the underside
the folded reset pose rests on).

Needs the reference's meshes (/root/reference): authoring container only.
  python tools/support_error.py [blob.json]
"""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gym_solo_b200.assets import compile_model as cm


def main():
  blob_path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, 'gym_solo_b200', 'assets', 'solo8v2.json')
  blob = json.load(open(blob_path))
  urdf = os.path.join('/root/reference', blob['source'])
  links, joints, children, base = cm.parse_urdf(urdf)
  mesh_dir = os.path.dirname(urdf)
  rng = np.random.RandomState(0)
  dirs = rng.standard_normal((10000, 3)); dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
  from scipy.spatial import ConvexHull
  cand_pos = np.array(blob['cand_pos']); cand_link = np.array(blob['cand_link'])
  out = {}
  for b, body in enumerate(blob['body_names']):
    clouds = []
    for P in blob['links']:
      if P['body'] != b:
        continue
      L = links[P['name']]
      if L['mesh'] is None or not os.path.exists(os.path.join(mesh_dir, L['mesh'])):
        continue
      v = cm.read_stl(os.path.join(mesh_dir, L['mesh']))
      hv = v[ConvexHull(v).vertices]
      p, R = np.array(P['pos']), np.array(P['rot'])
      clouds.append(p + (R @ (L['cxyz'][:, None] + L['cR'] @ hv.T)).T)      # BODY coordinates, like the candidates
    if not clouds:
      continue
    full = np.concatenate(clouds)
    red = cand_pos[np.array(blob['cand_body']) == b]
    e = (full @ dirs.T).max(0) - (red @ dirs.T).max(0)
    under = dirs[:, 2] < -np.cos(np.radians(30))          # within 30 degrees of the body's -z
    out[body] = dict(hull_points=int(len(full)), points=int(len(red)), max_mm=float(1e3 * e.max()),
                     p99_mm=float(1e3 * np.percentile(e, 99)), mean_mm=float(1e3 * e.mean()),
                     underside_max_mm=float(1e3 * e[under].max()),
                     recorded_by_compiler_mm=1e3 * blob.get('support_error', {}).get(body, float('nan')))
  print(json.dumps(out, indent=1))


if __name__ == '__main__':
  main()
