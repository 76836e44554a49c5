"""Multi-rank host logic on CPU: world_size-2 gloo process group, shard ranges,
per-rank seeds and the framework's only collective (all-gather of episode
statistics).  SURVEY.md section 8e."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gym_solo_b200 import sharding


def _free_port():
  s = socket.socket()
  s.bind(('127.0.0.1', 0))
  p = s.getsockname()[1]
  s.close()
  return p


def _worker(rank, world, port, total, out_dir):
  os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), RANK=str(rank),
                    WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
  dist.init_process_group('gloo', rank=rank, world_size=world)
  try:
    r, w, _ = sharding.rank_world()
    assert (r, w) == (rank, world)
    lo, hi = sharding.shard_range(total, r, w)
    # this rank's "episode statistics": return sum, length sum, episodes, nan restores
    mine = torch.tensor([10.0 * (rank + 1), 100.0 * (hi - lo), float(hi - lo), float(rank)])
    g = sharding.gather_statistics(mine)
    assert g.shape == (world, 4)
    for k in range(world):
      klo, khi = sharding.shard_range(total, k, world)
      assert g[k].tolist() == [10.0 * (k + 1), 100.0 * (khi - klo), float(khi - klo), float(k)]
    s = sharding.summarize(g)
    assert s['episodes'] == total and s['ranks'] == world
    assert abs(s['mean_length'] - 100.0) < 1e-12
    torch.save(dict(lo=lo, hi=hi, seed=sharding.rank_seed(1234, r)), os.path.join(out_dir, 'r%d.pt' % rank))
  finally:
    dist.destroy_process_group()


@pytest.mark.parametrize('total', [65536, 4097])
def test_two_rank_gloo_shards_and_statistics(tmp_path, total):
  world = 2
  mp.spawn(_worker, args=(world, _free_port(), total, str(tmp_path)), nprocs=world, join=True)
  parts = [torch.load(os.path.join(str(tmp_path), 'r%d.pt' % r)) for r in range(world)]
  assert parts[0]['lo'] == 0 and parts[-1]['hi'] == total
  assert parts[0]['hi'] == parts[1]['lo']                       # contiguous, disjoint, complete
  assert abs((parts[0]['hi'] - parts[0]['lo']) - (parts[1]['hi'] - parts[1]['lo'])) <= 1
  assert parts[0]['seed'] != parts[1]['seed']


def test_shard_range_properties():
  for total in (0, 1, 7, 4096, 262144):
    for world in (1, 2, 4, 8):
      spans = [sharding.shard_range(total, r, world) for r in range(world)]
      assert spans[0][0] == 0 and spans[-1][1] == total
      assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
      sizes = [hi - lo for lo, hi in spans]
      assert max(sizes) - min(sizes) <= 1
  with pytest.raises(ValueError):
    sharding.shard_range(8, 2, 2)


def test_single_process_gather_is_identity():
  mine = torch.tensor([1.0, 2.0, 3.0, 0.0])
  g = sharding.gather_statistics(mine)
  assert g.shape == (1, 4) and g[0].tolist() == mine.tolist()
