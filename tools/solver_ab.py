"""A/B of solver builds on one box: for every library given (build/variants/lib_*.so, built with
-DSOLO_V_* switches) a fresh process runs the bench workload (configs[1]) at the sizes given and
prints per-kernel times, the slowest environment's solver cycles and a checksum of the state.

  python tools/solver_ab.py 4096,65536 build/variants/lib_base.so build/variants/lib_fused.so ...
"""
import json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = r'''
import json, os, sys
import numpy as np, torch
sys.path.insert(0, %(root)r)
from gym_solo_b200 import _lib
_lib.LIB_PATH = %(lib)r
import bench
n = %(n)d
cfg = %(cfg)d
dev = torch.device('cuda', 0)
env = bench.make_env(cfg, n, dev)
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
print(json.dumps(out))
'''

def main():
  sizes = [int(x) for x in sys.argv[1].split(',')]
  cfg = 1
  libs = sys.argv[2:]
  if libs and libs[0].startswith('cfg='):
    cfg = int(libs[0][4:]); libs = libs[1:]
  for n in sizes:
    for lib in libs:
      code = CHILD % dict(root=ROOT, lib=os.path.abspath(lib), n=n, cfg=cfg)
      r = subprocess.run([sys.executable, '-c', code], capture_output=True, text=True)
      line = r.stdout.strip().splitlines()[-1] if r.stdout.strip() else 'FAILED: ' + r.stderr[-400:]
      print(n, os.path.basename(lib), line, flush=True)

if __name__ == '__main__':
  main()
