"""N > 1 host logic on CPU: world_size-2 gloo run of the sharding / reduction helpers the bench uses."""
import os

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from asimov_mjlab_b200 import shard


def _worker(rank, world, port, q):
  os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
  dist.init_process_group("gloo", rank=rank, world_size=world)
  r = shard.shard_range(4097, rank, world)
  rng = np.random.default_rng(shard.rank_seed(7, rank))
  ms = shard.reduce_max(10.0 + 5.0 * rank)
  st = shard.reduce_mean_stats({"reward": float(rank + 1), "len": 10.0 * (rank + 1)}, count=float(len(r)))
  dist.barrier()
  q.put((rank, r.start, r.stop, ms, st, float(rng.random())))
  dist.destroy_process_group()


def test_two_rank_sharding_and_reductions():
  world = 2
  ctx = mp.get_context("spawn")
  q = ctx.Queue()
  port = 29500 + os.getpid() % 2000
  ps = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
  for p in ps:
    p.start()
  out = sorted(q.get(timeout=120) for _ in ps)
  for p in ps:
    p.join(timeout=60)
    assert p.exitcode == 0
  (r0, a0, b0, ms0, st0, u0), (r1, a1, b1, ms1, st1, u1) = out
  assert (a0, b0, a1, b1) == (0, 2049, 2049, 4097)          # disjoint, contiguous, covers everything
  assert ms0 == ms1 == 15.0                                   # max over ranks
  w0, w1 = 2049, 2048
  assert st0 == st1 and abs(st0["reward"] - (1 * w0 + 2 * w1) / 4097) < 1e-12
  assert u0 != u1                                             # per-rank seeds differ


def test_shard_range_properties():
  for total in (1, 7, 4096, 65536):
    for world in (1, 2, 4, 8):
      parts = [shard.shard_range(total, r, world) for r in range(world)]
      assert sum(len(p) for p in parts) == total
      assert all(parts[i].stop == parts[i + 1].start for i in range(world - 1))
      assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1


def _policy_worker(rank, q):
  """What rl/train.py does on rank `rank`: the env seeds the process with seed + rank, then PPO builds the policy."""
  from asimov_mjlab_b200.rl.ppo import init_policy
  torch.manual_seed(100 + rank)                      # the env's per-rank seed
  before = torch.rand(4)                             # env draws ...
  ac = init_policy(0, 12, 15, 5, "cpu")
  after = torch.rand(4)                              # ... and keeps drawing from ITS stream afterwards
  torch.manual_seed(100 + rank)
  ref = torch.rand(8)
  w = torch.cat([p.detach().reshape(-1) for p in ac.parameters()])
  q.put((rank, w.numpy(), torch.cat([before, after]).numpy(), ref.numpy()))


def test_policy_init_is_identical_across_ranks_and_leaves_the_per_rank_streams_alone():
  """PPO used to call torch.manual_seed(seed) AFTER the env had seeded itself with seed + rank, which put
  every rank on the same random stream.  Two spawned 'ranks': same weights, different streams, stream not disturbed."""
  ctx = mp.get_context("spawn")
  q = ctx.Queue()
  ps = [ctx.Process(target=_policy_worker, args=(r, q)) for r in range(2)]
  for p in ps:
    p.start()
  out = sorted((q.get(timeout=120) for _ in ps), key=lambda t: t[0])
  for p in ps:
    p.join(timeout=60)
    assert p.exitcode == 0
  (_, w0, s0, ref0), (_, w1, s1, ref1) = out
  assert np.array_equal(w0, w1)                      # identical initial weights
  assert np.array_equal(s0, ref0) and np.array_equal(s1, ref1)      # the policy init consumed nothing from the rank's stream
  assert not np.array_equal(s0, s1)                  # and the ranks' streams differ
