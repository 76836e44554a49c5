"""Where does a step's time go?  Per-kernel CUDA-event times (solo_phase_timing),
back-to-back step rate without L2 flush, and the Python-side cost of env.step."""
import os, sys, time, json
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
dev = torch.device('cuda', 0)
env = bench.make_env(n, dev)
gen = torch.Generator(device=dev).manual_seed(1234)
acts = [(torch.rand((n, 12), generator=gen, device=dev) * 2 - 1) * 2 * np.pi for _ in range(64)]
for k in range(20):
  env.step(acts[k % 64])
torch.cuda.synchronize()
out = {}
# 1. back to back, no flush
t0 = time.perf_counter()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
K = 300
for k in range(K):
  env.step(acts[k % 64])
e1.record()
t1 = time.perf_counter()
torch.cuda.synchronize()
out['b2b_gpu_ms_per_step'] = e0.elapsed_time(e1) / K
out['b2b_cpu_issue_ms_per_step'] = 1e3 * (t1 - t0) / K
# 2. phase timing
env.client.phase_timing(True)
for k in range(K):
  env.step(acts[k % 64])
cnt, ms = env.client.phase_times()
env.client.phase_timing(False)
out['phase_ms'] = [m / cnt for m in ms]
out['phase_steps'] = cnt
# 3. with flush
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
env.client.phase_timing(True)
for k in range(100):
  flush.fill_(k & 0xff)
  env.step(acts[k % 64])
cnt, ms = env.client.phase_times()
out['phase_ms_flushed'] = [m / cnt for m in ms]
st = env.client.solver_stats().cpu().numpy()
order = np.argsort(-st[3])[:8]
out['worst_envs_iters_nc_active_cycles'] = [[int(st[0, i]), int(st[1, i]), int(st[2, i]), int(st[3, i])] for i in order]
out['mean_iters_nc_active_cycles'] = [float(st[k].mean()) for k in range(4)]
out['max_nc'] = int(st[1].max())
print(json.dumps(out))
