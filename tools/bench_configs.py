"""env-steps/s of BASELINE.json configs[2..4] on ONE GPU (their per-GPU shard sizes), same
timing rules as bench.py: CUDA events around every step on the launching stream, 256 MiB L2
flush between steps, fresh actions generated outside the timed spans.

  python tools/bench_configs.py [steps]
"""
import json, os, sys, warnings
import numpy as np
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gym_solo_b200.envs import Solo8VanillaEnv, Solo8VanillaConfig
from gym_solo_b200.envs.solo8v2vanilla import Solo8MMRConfig
from gym_solo_b200.core import obs, rewards, termination

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 100
dev = torch.device('cuda', 0)
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)


def run(name, env, n, width, bound, physics_per_step=1):
  gen = torch.Generator(device=dev).manual_seed(1234)
  act = lambda: (torch.rand((n, width), generator=gen, device=dev) * 2 - 1) * bound
  for _ in range(30):
    env.step(act())
  torch.cuda.synchronize()
  s = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
  e = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
  for k in range(steps):
    a = act()
    flush.fill_(k & 0xff)
    s[k].record(); env.step(a); e[k].record()
  torch.cuda.synchronize()
  ms = float(np.sum([x.elapsed_time(y) for x, y in zip(s, e)]))
  out = dict(config=name, envs_per_gpu=n, steps=steps, ms_per_step=ms / steps,
             env_steps_per_s=n * steps / (ms * 1e-3), physics_steps_per_s=physics_per_step * n * steps / (ms * 1e-3),
             episodes=float(env.episode_statistics()[2]), nan_restores=int(env.client.diag[2]))
  print(json.dumps(out), flush=True)
  env.close()


def base(cfg, n, auto_reset=False):
  cfg.dt = 1.0 / 240.0
  env = Solo8VanillaEnv(config=cfg, num_envs=n, device=dev, auto_reset=auto_reset)
  env.obs_factory.register_observation(obs.TorsoIMU(env.robot))
  env.obs_factory.register_observation(obs.MotorEncoder(env.robot))
  return env


with warnings.catch_warnings():
  warnings.simplefilter('ignore', DeprecationWarning)
  # configs[2]: 65536 envs over 8 GPUs = 8192 per GPU, Upright + HomePosition, termination resets
  env = base(Solo8VanillaConfig(), 8192, auto_reset=True)
  env.reward_factory.register_reward(1, rewards.UprightReward(env.robot))
  env.reward_factory.register_reward(1, rewards.HomePositionReward(env.robot))
  env.termination_factory.register_termination(termination.TimeBasedTermination(100))
  env.reset()
  run('configs[2] shard: 8192 envs, Upright+HomePosition, TimeBasedTermination(100) auto reset', env, 8192, 12, 2 * np.pi)
  # configs[3]: MMR v4 model, 16384 envs on one GPU
  env = base(Solo8MMRConfig(), 16384)
  env.reward_factory.register_reward(1, rewards.UprightReward(env.robot))
  env.termination_factory.register_termination(termination.PerpetualTermination())
  env.reset()
  run('configs[3]: MMR v4 model, 16384 envs', env, 16384, 10, 2 * np.pi)
  # configs[4]: torque control, 4 substeps, 262144 envs over 8 GPUs = 32768 per GPU
  cfg = Solo8VanillaConfig()
  cfg.control_mode = 'torque'
  cfg.num_substeps = 4
  env = base(cfg, 32768)
  env.reward_factory.register_reward(1, rewards.UprightReward(env.robot))
  env.termination_factory.register_termination(termination.PerpetualTermination())
  env.reset()
  run('configs[4] shard (8 GPUs): torque control, 4 substeps, 32768 envs', env, 32768, 12, 2.0, physics_per_step=4)
