"""Frames/s of render(): `count` environments per call at the
reference's default camera and 369 x 369 image, CUDA events around `reps` calls.

  python tools/bench_render.py [n_envs] [count] [reps]
"""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
  n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
  count = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
  reps = int(sys.argv[3]) if len(sys.argv) > 3 else 20
  from gym_solo_b200.envs import Solo8VanillaEnv, Solo8VanillaConfig
  from gym_solo_b200.core import obs, rewards, termination
  cfg = Solo8VanillaConfig()
  cfg.dt = 1.0 / 240.0
  env = Solo8VanillaEnv(config=cfg, num_envs=n, device='cuda:0')
  env.obs_factory.register_observation(obs.TorsoIMU(env.robot))
  env.reward_factory.register_reward(1, rewards.UprightReward(env.robot))
  env.termination_factory.register_termination(termination.PerpetualTermination())
  env.reset()
  g = torch.Generator(device='cuda').manual_seed(0)
  for _ in range(30):
    env.step((torch.rand((n, 12), generator=g, device='cuda') * 2 - 1) * 2 * np.pi)
  c = env.client
  view = c.computeViewMatrixFromYawPitchRoll(cfg.render_pos, cfg.render_cam_distance, cfg.render_yaw,
                                             cfg.render_pitch, cfg.render_roll, 2)
  proj = c.computeProjectionMatrixFOV(cfg.render_fov, cfg.render_aspect, 0.01, 100)
  w, h = cfg.render_width, cfg.render_height
  out = torch.empty((count, h, w, 4), dtype=torch.uint8, device='cuda')
  for _ in range(3):
    c.getCameraImage(w, h, view, proj, 0, count, out=out, depth=False, segmentation=False)
  torch.cuda.synchronize()
  e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
  e0.record()
  for _ in range(reps):
    c.getCameraImage(w, h, view, proj, 0, count, out=out, depth=False, segmentation=False)
  e1.record()
  torch.cuda.synchronize()
  ms = e0.elapsed_time(e1) / reps
  bytes_out = count * h * w * 4
  print(json.dumps(dict(n_envs=n, envs_per_call=count, width=w, height=h, ms_per_call=ms,
                        frames_per_s=count / ms * 1e3, output_GBps=bytes_out / ms / 1e6,
                        n_tri=int(c._mesh[0].numel()))))


if __name__ == '__main__':
  main()
