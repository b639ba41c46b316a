"""PPO rollout + update on the engine, one process per GPU (BASELINE.json configs[4]; reference scripts/train.py:36-124).

  python -m asimov_mjlab_b200.rl.train --envs 4096 --iters 10
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \
      -m asimov_mjlab_b200.rl.train --envs 4096 --iters 10

Rank 0 prints one JSON line: env-steps/s over all ranks (rollout + update, timed on the device, max over ranks), the
split between rollout and update, and the training statistics of the last iteration.  Logging is offline (no W&B).
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch
import torch.distributed as dist


def main():
  ap = argparse.ArgumentParser()
  ap.add_argument("--envs", type=int, default=4096, help="envs per GPU")
  ap.add_argument("--iters", type=int, default=10)
  ap.add_argument("--warmup", type=int, default=2)
  ap.add_argument("--steps-per-env", type=int, default=24)
  ap.add_argument("--seed", type=int, default=0)
  ap.add_argument("--task", default="velocity", choices=["velocity", "tracking"], help="G1 velocity-flat or G1 tracking-flat (synthetic motion)")
  ap.add_argument("--origin-sites", action="store_true", help="velocity task: the scene with mjlab's one world-body site per env (SURVEY.md H3)")
  ap.add_argument("--log", action="store_true", help="rank 0 prints the training statistics of every iteration to stderr")
  a = ap.parse_args()
  rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
  dev = torch.device(f"cuda:{local}")
  torch.cuda.set_device(dev)
  if world > 1:
    dist.init_process_group("nccl", device_id=dev)
  from .ppo import PPO
  from .velocity_env import G1VelocityFlatEnv
  if a.task == "tracking":
    from .tracking_env import G1TrackingFlatEnv
    env = G1TrackingFlatEnv(a.envs, str(dev), seed=a.seed + rank)
  else:
    env = G1VelocityFlatEnv(a.envs, str(dev), seed=a.seed + rank, origin_sites=a.origin_sites)         # per-rank seed = seed + rank (utils/random.py:9-26)
  algo = PPO(env, dev, steps_per_env=a.steps_per_env, seed=a.seed)
  ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
  t_roll = t_upd = 0.0
  info = {}
  for it in range(a.warmup + a.iters):
    if it == a.warmup:
      t_roll = t_upd = 0.0
      if world > 1:
        dist.barrier()
      torch.cuda.synchronize(dev)
    ev[0].record(); algo.rollout(); ev[1].record(); info = algo.update(); ev[2].record()
    torch.cuda.synchronize(dev)
    t_roll += ev[0].elapsed_time(ev[1]); t_upd += ev[1].elapsed_time(ev[2])
    if a.log:
      st = algo.episode_stats()
      if rank == 0:
        print(json.dumps({"iter": it, **info, **st, "mean_reward_per_step": float(algo.b_rew.mean()), "max_abs_reward": float(algo.b_rew.abs().max()),
                          "max_abs_action": float(algo.b_act.abs().max())}), file=sys.stderr, flush=True)
  t = torch.tensor([t_roll, t_upd], device=dev)
  if world > 1:
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
  stats = algo.episode_stats()
  if rank == 0:
    steps = a.iters * a.steps_per_env * a.envs * world
    total_ms = float(t[0] + t[1])
    print(json.dumps({
      "metric": "ppo_env_steps_per_s", "value": steps / (total_ms * 1e-3), "unit": "env-steps/s (4 physics sub-steps each), rollout + update",
      "n_gpus": world, "envs_per_gpu": a.envs, "iters": a.iters, "steps_per_env": a.steps_per_env,
      "rollout_ms_per_iter": float(t[0]) / a.iters, "update_ms_per_iter": float(t[1]) / a.iters,
      "rollout_env_steps_per_s": steps / (float(t[0]) * 1e-3), "train": {**info, **stats},
      "config": {"workload": ("Mjlab-Tracking-Flat-Unitree-G1 (synthetic motion)" if a.task == "tracking" else "Mjlab-Velocity-Flat-Unitree-G1") + ", PPO (rsl_rl hyper-parameters), whole env step in one CUDA graph",
                 "parallelism": f"env-shard x{world}; NCCL: gradient all-reduce per mini-batch, normaliser / KL / statistics reductions"}}), flush=True)
  if world > 1:
    dist.destroy_process_group()


if __name__ == "__main__":
  main()
