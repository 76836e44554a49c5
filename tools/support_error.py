"""Contact geometry: how far is the reduced support-point set of every link from the link's full
STL convex hull?  For a plane the contact point of a convex hull is its support point in the
direction of the plane normal, so the error of the reduction in a direction d is
  h_full(d) - h_reduced(d) >= 0   (metres; the reduced set is a subset of the hull),
i.e. how far the real hull would reach below the lowest retained point.  Reported per link as the
maximum and the 99th percentile over 10^4 uniformly random directions, and over the directions
within 30 degrees of the link's -z (the underside the folded reset pose rests on).

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
  parts = {L['name']: L for L in blob['links']}
  for name, L in links.items():
    if L['mesh'] is None or not os.path.exists(os.path.join(mesh_dir, L['mesh'])):
      continue
    v = cm.read_stl(os.path.join(mesh_dir, L['mesh']))
    hv = v[ConvexHull(v).vertices]
    # full hull in BODY coordinates, like the candidates
    P = parts[name]
    p, R = np.array(P['pos']), np.array(P['rot'])
    full = (p + (R @ (L['cxyz'][:, None] + L['cR'] @ hv.T)).T)
    red = cand_pos[cand_link == name]
    e = (full @ dirs.T).max(0) - (red @ dirs.T).max(0)
    down = -R[:, 2]                      # the link's own -z in body coordinates
    under = dirs @ down > np.cos(np.radians(30))
    out[name] = dict(hull_vertices=int(len(hv)), points=int(len(red)), max_mm=float(1e3 * e.max()),
                     p99_mm=float(1e3 * np.percentile(e, 99)), mean_mm=float(1e3 * e.mean()),
                     underside_max_mm=float(1e3 * e[under].max()))
  print(json.dumps(out, indent=1))


if __name__ == '__main__':
  main()
