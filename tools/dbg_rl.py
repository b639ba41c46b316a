"""Per-term reward magnitudes of the G1 velocity env under random actions (debug aid)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import importlib, torch
pkg = importlib.import_module("asimov_mjlab_b200")
from asimov_mjlab_b200.rl.velocity_env import G1VelocityFlatEnv

env = G1VelocityFlatEnv(1024, "cuda:0", seed=0)
torch.manual_seed(1)
worst = {}
for it in range(200):
  a = torch.randn(env.num_envs, env.num_actions, device=env.device)
  obs, rew, dones, ex = env.step(a)
  if it % 20 == 0 or not torch.isfinite(rew).all() or rew.abs().max() > 100:
    print(it, "rew min/max", float(rew.min()), float(rew.max()), "dones", int(dones.sum()), "qvel max", float(env.D.qvel.abs().max()),
          "obs max", float(obs.abs().max()), "finite", bool(torch.isfinite(env.D.qpos).all()))
for k, v in env.term_sums.items():
  print(f"{k:28s} mean {float(v.mean()):12.4g}  min {float(v.min()):12.4g}  max {float(v.max()):12.4g}")
