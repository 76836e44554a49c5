"""Multi-GPU plumbing of the env-step path: independent shards, one process per
GPU.  Environments never interact
(gym_solo/envs/test_solo8v2vanilla.py:177-194), so there is NO data-path
collective; the only exchange is an all-gather of four floats of episode
statistics per rank (NCCL over NVLink on GPUs, gloo in the CPU tests)."""
import os

import torch


def rank_world():
  """(rank, world, local_rank) from torch.distributed if initialised, else from
  the torchrun environment, else a single process."""
  try:
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
      return dist.get_rank(), dist.get_world_size(), int(os.environ.get('LOCAL_RANK', dist.get_rank()))
  except Exception:
    pass
  return (int(os.environ.get('RANK', '0')), int(os.environ.get('WORLD_SIZE', '1')),
          int(os.environ.get('LOCAL_RANK', '0')))


def shard_range(total_envs: int, rank: int, world: int):
  """Contiguous block [lo, hi) of a `total_envs` batch owned by `rank`: the
  first `total % world` ranks take one extra environment."""
  if world < 1 or not 0 <= rank < world:
    raise ValueError('bad rank / world size')
  if total_envs < 0:
    raise ValueError('total_envs must be >= 0')
  base, extra = divmod(int(total_envs), int(world))
  lo = rank * base + min(rank, extra)
  return lo, lo + base + (1 if rank < extra else 0)


def rank_seed(base_seed: int, rank: int) -> int:
  """Per-rank RNG seed for synthetic actions (bench.py: 1234 + rank)."""
  return int(base_seed) + int(rank)


def gather_statistics(mine: torch.Tensor) -> torch.Tensor:
  """all_gather of one small per-rank vector -> [world, k].  Without an
  initialised process group this is the single-rank answer [1, k]."""
  import torch.distributed as dist
  if not (dist.is_available() and dist.is_initialized()):
    return mine[None].clone()
  out = [torch.empty_like(mine) for _ in range(dist.get_world_size())]
  dist.all_gather(out, mine)
  return torch.stack(out)


def summarize(gathered: torch.Tensor):
  """[world, 4] rows (return sum, length sum, episodes, nan restores) -> job-wide
  dict with episode-weighted means."""
  g = gathered.to(torch.float64).cpu()
  episodes = float(g[:, 2].sum())
  return dict(episodes=episodes,
              mean_return=float(g[:, 0].sum()) / episodes if episodes else 0.0,
              mean_length=float(g[:, 1].sum()) / episodes if episodes else 0.0,
              nan_restores=float(g[:, 3].sum()), ranks=int(g.shape[0]))
