"""bench.py -- Solo8 env-steps/sec on N B200s (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--config C] [--envs E] [--impl reference]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

--config selects the BASELINE.json configuration (default 1 = configs[1], the one the metric
is quoted on; 2, 3, 4 = configs[2], [3], [4] with the same JSON schema, see CONFIGS below).

A "step" is one env.step() of every environment of the batch: three kernels
(setup -> solve -> finish) per physics substep, replayed as one captured CUDA
graph per rank.  Workload = BASELINE.json configs[1]: Solo8VanillaEnv, 4096
envs per GPU, POSITION_CONTROL, flat ground, dt = 1/240, TorsoIMU +
MotorEncoder observations (21), 1 x UprightReward, PerpetualTermination, fresh
U(-2pi, 2pi) actions every step (generated on the device outside the timed
spans).  Weak scaling: every rank owns its own 4096 envs, no data-path
collective; NCCL only gathers episode statistics after the timed region.

value  = env-steps all ranks processed / max-over-ranks device time (CUDA events
         around every step on the launching stream; L2 flushed between steps)
e2e    = the same through Solo8VanillaEnv.step_host (numpy in / numpy out via
         the C-ABI's solo_step_host: H2D actions + step + D2H obs/reward/done)
roofline.achieved = 377 algorithmic bytes per env-step x envs per
         launch / mean launch duration, against the measured HBM copy peak
cpu_baseline / --impl reference = the double-precision CPU oracle (a port: the
         reference's own engine, pybullet, is not installable here) on the
         box's host threads.
"""
import argparse
import json
import os
import sys
import threading
import time
import warnings

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
  sys.path.insert(0, ROOT)

METRIC = 'solo8_env_steps_per_sec'
UNIT = 'env-steps/s'
DT = 1.0 / 240.0

# BASELINE.json configs[1..4].  envs: per GPU (weak scaling) or for the whole job (strong).
# bytes: algorithmic bytes per env-step (state read + written, action, obs, reward,
# done, counter, float32; substeps stay on chip).
CONFIGS = {
    1: dict(workload='Solo8VanillaEnv batched 4096 envs on 1xB200, position control, flat-ground contact, '
                     'dt=1/240 (BASELINE.json configs[1])',
            envs=4096, scaling='weak', model='solo8v2', control='position', substeps=1, n_api=12, obs_dim=21,
            reward='upright', term=None, bytes=377, bound=2 * np.pi,
            program='TorsoIMU+MotorEncoder, 1xUprightReward, PerpetualTermination'),
    2: dict(workload='Solo8VanillaEnv 65536 envs sharded over 8xB200 (8192 per GPU) with HomePositionReward+'
                     'UprightReward and termination resets (BASELINE.json configs[2])',
            envs=8192, scaling='weak', model='solo8v2', control='position', substeps=1, n_api=12, obs_dim=21,
            reward='upright+home', term=1000, bytes=377, bound=2 * np.pi,
            program='TorsoIMU+MotorEncoder, 1xUprightReward + 1xHomePositionReward, '
                    'TimeBasedTermination(1000) with in-kernel snapshot auto reset'),
    3: dict(workload='examples/solo8_mmr environment variant (MMR v4 URDF) batched 16384 envs on 1xB200 '
                     '(BASELINE.json configs[3])',
            envs=16384, scaling='weak', model='solo8_mmr_v4', control='position', substeps=1, n_api=10, obs_dim=19,
            reward='upright', term=None, bytes=393, bound=2 * np.pi,
            program='TorsoIMU+MotorEncoder (10 joints), 1xUprightReward, PerpetualTermination'),
    4: dict(workload='Solo8 torque-control mode, 4 substeps per action, 262144 envs across the GPUs of the job '
                     '(BASELINE.json configs[4], strong scaling)',
            envs=262144, scaling='strong', model='solo8v2', control='torque', substeps=4, n_api=12, obs_dim=21,
            reward='upright', term=None, bytes=377, bound=2.0,
            program='TorsoIMU+MotorEncoder, 1xUprightReward, PerpetualTermination; torque control, '
                    '4 physics substeps per env-step'),
}


_REAL_STDOUT = None


def emit(line):
  out = _REAL_STDOUT or sys.stdout
  out.write(json.dumps(line) + '\n')
  out.flush()


def parse_args():
  ap = argparse.ArgumentParser()
  ap.add_argument('--gpus', type=int, default=1)
  ap.add_argument('--steps', type=int, default=1000)
  ap.add_argument('--warmup', type=int, default=20)
  ap.add_argument('--config', type=int, default=1, choices=sorted(CONFIGS), help='BASELINE.json configs[C]')
  ap.add_argument('--envs', type=int, default=None,
                  help='environments per GPU (default: the configuration\'s; whole-job count / N when it scales strongly)')
  ap.add_argument('--stream', default='side', choices=['side', 'default'],
                  help='run on a torch side stream (captured directly) or on the legacy default stream')
  ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
  ap.add_argument('--no-cpu-baseline', action='store_true')
  ap.add_argument('--no-l2-flush', action='store_true')
  return ap.parse_args()


def hbm_peak():
  p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
  if os.path.exists(p):
    try:
      return float(json.load(open(p))['hbm_gbs']), 'measured'
    except Exception:
      pass
  return 6650.0, 'fallback'


def _capture(n_envs):
  """profiles/traffic.json: numbers of the committed `ncu --set full` captures, per batch size."""
  p = os.path.join(ROOT, 'profiles', 'traffic.json')
  try:
    return json.load(open(p)).get('by_envs', {}).get(str(int(n_envs)))
  except Exception:
    return None


def ncu_instructions(kernel, n_envs):
  """smsp__inst_executed.sum of one launch of `kernel` from the capture at this batch size, or None."""
  c = _capture(n_envs)
  return c['warp_instructions_per_launch'].get(kernel) if c else None


def ncu_traffic(kernel, n_envs):
  """dram__bytes_read.sum + dram__bytes_write.sum of one launch of `kernel` from the committed
  `ncu --set full` capture at this batch size (ncu's cold caches); None without one."""
  c = _capture(n_envs)
  return c['dram_bytes_per_launch'].get(kernel) if c else None


def issue_util(kernel, n_envs, launch_ms, clocks):
  """Secondary roofline for this latency/issue-bound path: warp instructions of one launch
  (ncu capture) / issue slots the launch had (148 SMs x 4 schedulers x cycles at the sampled
  SM clock).  None without a capture for this batch size."""
  inst = ncu_instructions(kernel, n_envs)
  mhz = clocks.get('sm_mhz') or clocks.get('sm_max_mhz')
  if not inst or not mhz or launch_ms <= 0:
    return None
  return inst / (148 * 4 * launch_ms * 1e-3 * mhz * 1e6)


# ---- clocks under load -------------------------------------------------------------
class ClockSampler(threading.Thread):
  REASONS = [(0x8, 'hw_slowdown'), (0x40, 'hw_thermal_slowdown'), (0x20, 'sw_thermal_slowdown'),
             (0x4, 'sw_power_cap'), (0x80, 'hw_power_brake_slowdown')]

  def __init__(self, index, period=0.02):
    super().__init__(daemon=True)
    self.index, self.period = index, period
    self.samples, self.reasons, self.max_mhz = [], set(), None
    self._stop = threading.Event()
    self.recording = False
    try:
      import pynvml
      pynvml.nvmlInit()
      self.nv = pynvml
      self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
      self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
    except Exception:
      self.nv = None

  def run(self):
    if self.nv is None:
      return
    while not self._stop.is_set():
      if self.recording:
        try:
          self.samples.append(float(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)))
          mask = int(self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
          for bit, name in self.REASONS:
            if mask & bit:
              self.reasons.add(name)
        except Exception:
          pass
      time.sleep(self.period)

  def stop(self):
    self._stop.set()

  def summary(self):
    if not self.samples:
      return {'sm_mhz': None, 'sm_max_mhz': self.max_mhz, 'reasons': sorted(self.reasons), 'samples': 0}
    return {'sm_mhz': float(np.median(self.samples)), 'sm_max_mhz': self.max_mhz,
            'reasons': sorted(self.reasons), 'samples': len(self.samples)}


# ---- the workload --------------------------------------------------------------------
def make_env(cfg_id, n_envs, device):
  from gym_solo_b200.envs import Solo8VanillaEnv, Solo8VanillaConfig
  from gym_solo_b200.envs.solo8v2vanilla import Solo8MMRConfig
  from gym_solo_b200.core import obs, rewards, termination
  K = CONFIGS[cfg_id]
  cfg = Solo8MMRConfig() if K['model'] == 'solo8_mmr_v4' else Solo8VanillaConfig()
  cfg.dt = DT
  cfg.control_mode = K['control']
  cfg.num_substeps = K['substeps']
  env = Solo8VanillaEnv(config=cfg, num_envs=n_envs, device=device, auto_reset=K['term'] is not None)
  env.obs_factory.register_observation(obs.TorsoIMU(env.robot))
  env.obs_factory.register_observation(obs.MotorEncoder(env.robot))
  with warnings.catch_warnings():
    warnings.simplefilter('ignore', DeprecationWarning)
    env.reward_factory.register_reward(1, rewards.UprightReward(env.robot))
    if K['reward'] == 'upright+home':
      env.reward_factory.register_reward(1, rewards.HomePositionReward(env.robot))
  env.termination_factory.register_termination(
      termination.TimeBasedTermination(K['term']) if K['term'] is not None else termination.PerpetualTermination())
  env.reset()
  return env


def cpu_oracle_setup(cfg_id):
  from gym_solo_b200 import _abi
  from oracle import solo_oracle as so
  K = CONFIGS[cfg_id]
  blob = _abi.load_model_blob(K['model'])
  model = _abi.make_model(blob, link_friction=0.5)
  config = _abi.make_config(dt=DT, num_substeps=K['substeps'],
                            control_mode=_abi.CONTROL_TORQUE if K['control'] == 'torque' else _abi.CONTROL_POSITION)
  orc = so.PhysicsOracle(model, config)
  if K['model'] == 'solo8v2':
    targets = np.array([np.pi / 2, np.pi, 0, np.pi / 2, np.pi, 0, -np.pi / 2, -np.pi, 0, -np.pi / 2, -np.pi, 0])
  else:
    # solo8v2vanilla.py:124-125: starting_joint_pos looked up by joint name, 0 for names it does not have
    start = {'FL_HFE': np.pi / 2, 'FL_KFE': np.pi, 'FR_HFE': np.pi / 2, 'FR_KFE': np.pi,
             'HL_HFE': -np.pi / 2, 'HL_KFE': -np.pi, 'HR_HFE': -np.pi / 2, 'HR_KFE': -np.pi}
    targets = np.array([start.get(j, 0.0) for j in blob['api_names']])
  snap = orc.settle(targets)
  return so, orc, model, snap


def cpu_step(K, so, orc, model, snap, states, counters, actions, threads):
  """One env.step of the batch on the host: physics (C, pthreads) + the numpy
  restatement of get_obs / get_reward / is_terminated (+ the snapshot auto reset)."""
  states = orc.step_batch(states, actions, n_steps=1, n_threads=threads)
  view = so.StateView(states, model)
  obs = np.concatenate([so.torso_imu(view)[0], so.motor_encoder(view)[0]], axis=1)
  rew = so.upright_reward(view)
  if K['reward'] == 'upright+home':
    rew = rew + so.home_position_reward(view)
  done = np.zeros(states.shape[0], dtype=bool)
  if K['term'] is not None:
    counters += 1
    done = counters > K['term']
    if done.any():
      states[done] = snap
      counters[done] = 0
  return states, obs, rew, done


def run_cpu(cfg_id, n_envs, steps, warmup, threads, seed=1234):
  K = CONFIGS[cfg_id]
  so, orc, model, snap = cpu_oracle_setup(cfg_id)
  rng = np.random.RandomState(seed)
  states = np.tile(snap, (n_envs, 1))
  counters = np.zeros(n_envs, dtype=np.int64)
  na, bound = K['n_api'], K['bound']
  for _ in range(warmup):
    states, *_ = cpu_step(K, so, orc, model, snap, states, counters, rng.uniform(-bound, bound, (n_envs, na)), threads)
  acts = [rng.uniform(-bound, bound, (n_envs, na)) for _ in range(min(steps, 64))]
  t0 = time.perf_counter()
  for k in range(steps):
    states, *_ = cpu_step(K, so, orc, model, snap, states, counters, acts[k % len(acts)], threads)
  dt = time.perf_counter() - t0
  return n_envs * steps / dt, dt


def cpu_model_name():
  try:
    for line in open('/proc/cpuinfo'):
      if line.startswith('model name'):
        return line.split(':', 1)[1].strip()
  except Exception:
    pass
  return 'unknown'


def resolve_envs(args, world):
  K = CONFIGS[args.config]
  if args.envs is not None:
    return args.envs
  return K['envs'] if K['scaling'] == 'weak' else K['envs'] // world


def reference_arm(args, rank, world):
  """--impl reference: the CPU path on the host cores.  The reference's engine
  (pybullet) cannot be installed (no wheel, no network), so this times the
  oracle port; one step = one env.step of one GPU's batch of the configuration."""
  if rank != 0:
    return
  K = CONFIGS[args.config]
  n = resolve_envs(args, world)
  if K['scaling'] == 'strong':
    n = min(n, 8192)          # bounded sample of the 262144-env job: the CPU rate does not depend on the batch
  threads = os.cpu_count() or 1
  value, total = run_cpu(args.config, n, args.steps, min(args.warmup, 3), threads)
  line = {
      'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': args.gpus,
      'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': 1e3 * total / args.steps,
      'higher_is_better': True, 'scaling': K['scaling'], 'vs_baseline': None, 'dtype': 'f64',
      'data': 'synthetic',
      'config': {'workload': K['workload'], 'envs_per_step': n, 'engine': 'oracle port (C, double)',
                 'program': K['program'], 'cpu_model': cpu_model_name()},
      'cpu_baseline': {'value': value, 'unit': UNIT, 'cores': threads, 'kind': 'port', 'cpu_model': cpu_model_name(),
                       'sample': '%d envs x %d steps, physics on %d pthreads + numpy epilogue; pybullet itself is '
                                 'not installed on this box' % (n, args.steps, threads)},
      'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
      'gpu_launches': 0,
  }
  emit(line)


def main():
  # stdout carries exactly ONE JSON line: native libraries (NCCL's version banner ...) write to
  # file descriptor 1 behind Python's back, so fd 1 is pointed at stderr for the whole run and
  # the line goes to a private duplicate of the original stdout
  global _REAL_STDOUT
  sys.stdout.flush()
  _REAL_STDOUT = os.fdopen(os.dup(1), 'w')
  os.dup2(2, 1)
  args = parse_args()
  rank = int(os.environ.get('RANK', '0'))
  world = int(os.environ.get('WORLD_SIZE', '1'))
  local = int(os.environ.get('LOCAL_RANK', '0'))
  if args.impl == 'reference':
    reference_arm(args, rank, world)
    return
  K = CONFIGS[args.config]

  import torch
  import torch.distributed as dist
  if not torch.cuda.is_available():
    raise SystemExit('bench.py: no CUDA device; the Solo 8 step path has no CPU fallback')
  torch.cuda.set_device(local)
  device = torch.device('cuda', local)
  if world > 1:
    os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
    # NCCL's own log lines (version banner, NCCL_DEBUG output) go to stderr: stdout carries
    # exactly one JSON line
    os.environ.setdefault('NCCL_DEBUG_FILE', '/dev/stderr')
    dist.init_process_group('nccl', device_id=device)

  n = resolve_envs(args, world)
  na, bound, od = K['n_api'], float(K['bound']), K['obs_dim']
  # everything below runs on ONE stream: a torch side stream (the step's graph is captured on
  # it directly) or the legacy default stream (--stream default)
  stream = torch.cuda.Stream(device=device) if args.stream == 'side' else torch.cuda.current_stream(device)
  with torch.cuda.stream(stream):
    env = make_env(args.config, n, device)
    gen = torch.Generator(device=device).manual_seed(1234 + rank)

    def fresh_actions():
      return (torch.rand((n, na), generator=gen, device=device) * 2 - 1) * bound

    flush = None if args.no_l2_flush else torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=device)
    sampler = ClockSampler(local)
    sampler.start()

    def barrier():
      if world > 1:
        dist.barrier()
      torch.cuda.synchronize(device)

    # ---- device-resident arm ------------------------------------------------------------
    for _ in range(args.warmup):
      env.step(fresh_actions())
    barrier()
    sampler.recording = True
    launches0 = env.client.launch_count()
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    ends = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    for k in range(args.steps):
      a = fresh_actions()
      if flush is not None:
        flush.fill_(k & 0xff)
      starts[k].record()
      env.step(a)
      ends[k].record()
    barrier()
    launches = env.client.launch_count() - launches0
    per_step_ms = np.array([s.elapsed_time(e) for s, e in zip(starts, ends)])
    total_ms = float(per_step_ms.sum())

    # ---- per-kernel durations, live (CUDA events on the launching stream inside the C-ABI; plain
    # launches with events between the kernels instead of the graph replay) --
    env.client.phase_timing(True)
    for k in range(min(args.steps, 200)):
      a = fresh_actions()
      if flush is not None:
        flush.fill_(k & 0xff)
      env.step(a)
    phase_steps, phase_ms = env.client.phase_times()
    env.client.phase_timing(False)
    phase_ms = [m / max(1, phase_steps) for m in phase_ms]      # per PHYSICS step (x substeps per env-step)
    # the solver's own clock: SM cycles every solver warp was resident for in one launch
    env.step(fresh_actions())
    torch.cuda.synchronize(device)
    solver_warp_cycles = float(env.client.solver_stats()[3].double().sum())

    # ---- end-to-end arm: host buffers through solo_step_host -----------------------------
    rng = np.random.RandomState(4321 + rank)
    # the step's inputs live in page-locked host memory; the kernels read them from there and write
    # the results to page-locked host memory, inside the step
    host_actions = [torch.from_numpy(rng.uniform(-bound, bound, (n, na)).astype(np.float32)).pin_memory()
                    for _ in range(16)]
    for k in range(min(args.warmup, 5)):
      env.step_host(host_actions[k % 16])
    barrier()
    e2e_steps = args.steps
    t0 = time.perf_counter()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    checksum = 0.0
    for k in range(e2e_steps):
      o, r, d, _ = env.step_host(host_actions[k % 16])
      checksum += float(r[0]) + float(o[n - 1, od - 1]) + float(d[n // 2])
    ev1.record()
    barrier()
    e2e_ms = max(ev0.elapsed_time(ev1), 0.0)
    e2e_wall_ms = 1e3 * (time.perf_counter() - t0)
    sampler.recording = False
    sampler.stop()

    t = torch.tensor([total_ms, e2e_ms, e2e_wall_ms], dtype=torch.float64, device=device)
    if world > 1:
      dist.all_reduce(t, op=dist.ReduceOp.MAX)
      stats = env.episode_statistics(all_gather=True)   # the framework's only collective
    else:
      stats = env.episode_statistics()[None]
    total_ms, e2e_ms, e2e_wall_ms = (float(x) for x in t.cpu())
    diag = env.client.diag.cpu().numpy()

  if rank == 0:
    nbytes = K['bytes']
    value = world * n * args.steps / (total_ms * 1e-3)
    e2e_value = world * n * e2e_steps / (max(e2e_ms, e2e_wall_ms) * 1e-3)
    peak, peak_kind = hbm_peak()
    step_ms = float(per_step_ms.mean())
    names = ['setup_legpar_kernel', 'solve_warp_kernel', 'finish_kernel']
    dom = int(np.argmax(phase_ms))
    launch_ms = phase_ms[dom]
    sub = K['substeps']
    # one launch of the dominant kernel advances n environments by 1 / substeps of an env-step
    achieved = nbytes * n / sub / (launch_ms * 1e-3) / 1e9
    clocks = sampler.summary()
    mhz = clocks.get('sm_mhz') or clocks.get('sm_max_mhz') or 1965.0
    line = {
        'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps,
        'warmup': args.warmup, 'ms_per_step': total_ms / args.steps, 'higher_is_better': True,
        'scaling': K['scaling'], 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
        'config': {'workload': K['workload'], 'baseline_config': args.config, 'envs_per_gpu': n, 'dt': DT,
                   'obs_dim': od, 'program': K['program'], 'physics_steps_per_env_step': sub,
                   'actions': 'U(-%.4g, %.4g)^%d fresh each step, generated on device outside the timed spans'
                              % (bound, bound, na),
                   'l2': 'flushed between timed steps (256 MiB fill)' if flush is not None else 'not flushed',
                   'block_threads': env.client.block_size(), 'solver': 'one warp per environment',
                   'stream': args.stream,
                   'execution': 'one captured CUDA graph per env.step (setup -> solve -> finish per physics '
                                'substep, solver and finish as programmatic dependent launches)',
                   'timing': 'CUDA events around every step on the launching stream, summed; max over ranks'},
        'roofline': {'bound': 'hbm', 'achieved': achieved, 'peak': peak, 'unit': 'GB/s',
                     'frac': achieved / peak, 'traffic': ncu_traffic(names[dom], n), 'peak_source': peak_kind,
                     'bytes_per_env_step': nbytes, 'kernel': names[dom],
                     'launch_ms_mean': launch_ms,
                     'kernels_ms_mean': dict(zip(names, phase_ms)), 'kernel_share_of_step': launch_ms / sum(phase_ms),
                     'kernels_timed': 'plain launches with events between the kernels (no overlap of dependent launches)',
                     'step_ms_mean': step_ms, 'step_ms_p50': float(np.median(per_step_ms)),
                     'step_ms_max': float(per_step_ms.max()),
                     'whole_step_frac': nbytes * n / (step_ms * 1e-3) / 1e9 / peak,
                     'issue_slot_util': issue_util(names[dom], n, launch_ms, clocks),
                     'issue_slot_util_source': 'warp instructions of profiles/traffic.json (ncu capture, same batch size) '
                                               '/ live issue slots; null without a capture for this batch size',
                     'solver_warp_residency': solver_warp_cycles / (148 * 4 * phase_ms[1] * 1e-3 * mhz * 1e6),
                     'solver_warp_residency_source': 'live: sum over environments of the solver warp\'s clock64 span in one '
                                                     'launch / (592 schedulers x launch cycles)',
                     'note': 'path is FP32-issue/latency bound, not HBM bound (SURVEY.md 8d)'},
        'e2e': {'value': e2e_value, 'unit': UNIT, 'h2d_bytes_per_step': world * n * na * 4,
                'd2h_bytes_per_step': world * n * (od * 4 + 4 + 1), 'steps': e2e_steps,
                'ratio_to_value': e2e_value / value,
                'api': 'Solo8VanillaEnv.step_host(pinned float32 CPU tensor) -> solo_step_host, synchronous: the '
                       'kernels read the actions from and write obs|reward|done to page-locked host memory '
                       '(zero-copy over PCIe, inside the step), stream sync; results read on the host',
                'checksum': checksum},
        'gpu_launches': int(launches),
        'clocks': clocks,
        'solver': {'mean_contacts_per_env_step': float(diag[0]) / max(1.0, float(diag[3])),
                   'mean_pgs_iterations': float(diag[1]) / max(1.0, float(diag[3])),
                   'nan_restores': int(diag[2])},
        'episode_stats_ranks': int(stats.shape[0]),
    }
    if world == 1 and not args.no_cpu_baseline:
      threads = os.cpu_count() or 1
      # ~10 s of host work on all threads (3.4e5 / 2.1e5 / 1.5e5 / 0.7e5 env-steps/s on the 16-thread box)
      c_envs, c_steps = {1: (2048, 1600), 2: (2048, 1000), 3: (2048, 750), 4: (2048, 350)}[args.config]
      cv, ct = run_cpu(args.config, c_envs, c_steps, 2, threads)
      line['cpu_baseline'] = {'value': cv, 'unit': UNIT, 'cores': threads, 'kind': 'port',
                              'cpu_model': cpu_model_name(),
                              'sample': '%d envs x %d steps of the same workload (%.1f s), oracle port: '
                                        'C double physics on %d pthreads + numpy epilogue; pybullet itself is not '
                                        'installed on this box' % (c_envs, c_steps, ct, threads)}
    emit(line)
  env.close()
  if world > 1:
    dist.barrier()
    dist.destroy_process_group()


if __name__ == '__main__':
  main()
