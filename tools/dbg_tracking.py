import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from asimov_mjlab_b200.rl import tracking_env as te
env = te.G1TrackingFlatEnv(256, "cuda:0", seed=2)
for pol in ("motion", "zero"):
  for v in env.term_sums.values(): v.zero_()
  nd = 0
  for i in range(40):
    ts = (env.time_steps + 1).clamp(max=env.T - 1)
    act = (env.mo_jp[ts] - env.default_joint_pos) / env.action_scale if pol == "motion" else torch.zeros(256, env.num_actions, device="cuda:0")
    obs, rew, dones, ex = env.step(act); nd += int(dones.sum())
  print(pol, "dones", nd, {k: round(float(v.mean()) / 40, 5) for k, v in env.term_sums.items()})
  print({k: round(float(v.median()), 4) for k, v in env.metrics.items()})
  print("terminated by: z", float(((env.mo_bpos[env.time_steps, env.anchor_k, 2] - env.D.xpos[:, env.body_ids[env.anchor_k], 2]).abs() > 0.25).float().mean()))
