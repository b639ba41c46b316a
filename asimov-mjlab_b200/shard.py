"""Multi-GPU plumbing: worlds shard across ranks with NO collective on the step path.

One process per GPU (torchrun); each rank owns `envs_per_rank` worlds, seeds its generators with seed + rank
(the reference seeds everything from one int, src/mjlab/utils/random.py:9-26) and the only reductions are of
timing / episode statistics.
"""
from __future__ import annotations

import os


def rank_world():
  return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


def shard_range(total: int, rank: int, world: int) -> range:
  """Contiguous shard of `total` worlds owned by `rank` (sizes differ by at most one)."""
  base, rem = divmod(total, world)
  lo = rank * base + min(rank, rem)
  return range(lo, lo + base + (1 if rank < rem else 0))


def rank_seed(seed: int, rank: int) -> int:
  return int(seed) + int(rank)


def reduce_max(value: float, device=None) -> float:
  """Max over ranks (timing is reported as the slowest rank)."""
  import torch
  import torch.distributed as dist
  if not (dist.is_available() and dist.is_initialized()):
    return float(value)
  t = torch.tensor([value], dtype=torch.float64, device=device)
  dist.all_reduce(t, op=dist.ReduceOp.MAX)
  return float(t[0])


def reduce_mean_stats(stats: dict, count: float, device=None) -> dict:
  """Count-weighted mean of scalar episode statistics over ranks (one flat all-reduce)."""
  import torch
  import torch.distributed as dist
  keys = sorted(stats)
  t = torch.tensor([stats[k] * count for k in keys] + [count], dtype=torch.float64, device=device)
  if dist.is_available() and dist.is_initialized():
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
  n = float(t[-1]) if float(t[-1]) > 0 else 1.0
  return {k: float(t[i]) / n for i, k in enumerate(keys)}
