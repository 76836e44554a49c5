"""A/B of solver builds / execution options on one box: for every variant given a fresh process
runs a bench workload at the sizes given and prints per-kernel times, the slowest environment's
solver cycles and a checksum of the state.  A variant is a library (build/variants/lib_*.so built
with -D switches; `product` = the in-tree library), optionally followed by @opt=value pairs for
solo_set_option (graph, zero_copy, occ).

  python tools/solver_ab.py 4096,65536 [cfg=2] product product@occ=1 build/variants/lib_x.so
"""
import json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = r'''
import json, os, sys
import numpy as np, torch
sys.path.insert(0, %(root)r)
from gym_solo_b200 import _lib
if %(lib)r:
  _lib.LIB_PATH = %(lib)r
import bench
from gym_solo_b200 import _abi
n = %(n)d
cfg = %(cfg)d
dev = torch.device('cuda', 0)
blob_override = [v for k, v in %(sopts)r if k == 'blob']
if blob_override:
  # another compiled model for the same workload (e.g. round 1's support-point sets)
  import gym_solo_b200.assets as _A
  _orig = _A.blob_for_urdf
  _A.blob_for_urdf = lambda p: os.path.abspath(blob_override[0]) if 'solo8v2' in p else _orig(p)
env = bench.make_env(cfg, n, dev)
for name, val in %(opts)r:
  env.client.set_option({'graph': _abi.OPT_GRAPH, 'zero_copy': _abi.OPT_ZERO_COPY, 'occ': _abi.OPT_SOLVER_OCCUPANCY, 'warm': _abi.OPT_FINISH_WARMUP}[name], val)
K = bench.CONFIGS[cfg]
gen = torch.Generator(device=dev).manual_seed(1234)
acts = [(torch.rand((n, K['n_api']), generator=gen, device=dev) * 2 - 1) * float(K['bound']) for _ in range(64)]
for k in range(40): env.step(acts[k %% 64])
torch.cuda.synchronize()
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
out = {}
reps = 200 if n <= 8192 else 60
# graph replay, per-step events, flushed (what bench.py's value measures)
s = [torch.cuda.Event(enable_timing=True) for _ in range(reps)]; e = [torch.cuda.Event(enable_timing=True) for _ in range(reps)]
for k in range(reps):
  flush.fill_(k & 0xff); s[k].record(); env.step(acts[k %% 64]); e[k].record()
torch.cuda.synchronize()
ms = np.array([a.elapsed_time(b) for a, b in zip(s, e)])
out['step_ms'] = float(ms.mean()); out['step_ms_p50'] = float(np.median(ms))
env.client.phase_timing(True)
mx = []
for k in range(reps):
  flush.fill_(k & 0xff); env.step(acts[k %% 64])
  if k %% 10 == 0:
    st = env.client.solver_stats(); mx.append(float(st[3].max()))
cnt, pm = env.client.phase_times()
env.client.phase_timing(False)
out['phase_ms'] = [round(m / cnt, 5) for m in pm]
out['max_cycles_mean'] = float(np.mean(mx))
out['checksum'] = float(env.get_state().double().abs().sum())
# end to end: synchronous step_host from a pool of pinned action buffers
import time
host = [torch.from_numpy(np.random.RandomState(k).uniform(-1, 1, (n, K['n_api'])).astype(np.float32) * float(K['bound'])).pin_memory() for k in range(16)]
for k in range(20): env.step_host(host[k %% 16])
t0 = time.perf_counter()
for k in range(reps): o, r, d, _ = env.step_host(host[k %% 16]); x = float(r[0])
out['e2e_us'] = 1e6 * (time.perf_counter() - t0) / reps
for k in range(reps): env.step(acts[k %% 64])
torch.cuda.synchronize()
t0 = time.perf_counter()
for k in range(reps): env.step(acts[k %% 64])
torch.cuda.synchronize()
out['b2b_us'] = 1e6 * (time.perf_counter() - t0) / reps
print(json.dumps(out))
'''

def main():
  sizes = [int(x) for x in sys.argv[1].split(',')]
  cfg = 1
  libs = sys.argv[2:]
  if libs and libs[0].startswith('cfg='):
    cfg = int(libs[0][4:]); libs = libs[1:]
  for n in sizes:
    for spec in libs:
      lib, *rest = spec.split('@')
      opts = [(x.split('=')[0], int(x.split('=')[1])) for x in rest if not x.startswith('blob=')]
      sopts = [(x.split('=')[0], x.split('=', 1)[1]) for x in rest if x.startswith('blob=')]
      code = CHILD % dict(root=ROOT, lib='' if lib == 'product' else os.path.abspath(lib), n=n, cfg=cfg, opts=opts,
                          sopts=sopts)
      r = subprocess.run([sys.executable, '-c', code], capture_output=True, text=True)
      line = r.stdout.strip().splitlines()[-1] if r.stdout.strip() else 'FAILED: ' + r.stderr[-400:]
      print(n, spec, line, flush=True)

if __name__ == '__main__':
  main()
