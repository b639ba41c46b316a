"""Profile one env step of the G1 velocity env (for `ncu --profile-from-start off`), and time env.step / policy act.
  ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file out.csv python tools/prof_env.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

from asimov_mjlab_b200.rl.velocity_env import G1VelocityFlatEnv

W = int(os.environ.get("W", "4096"))
if os.environ.get("ENV", "velocity") == "tracking":
  from asimov_mjlab_b200.rl.tracking_env import G1TrackingFlatEnv
  env = G1TrackingFlatEnv(W, "cuda:0", seed=0)
else:
  env = G1VelocityFlatEnv(W, "cuda:0", seed=0, use_fused=bool(int(os.environ.get("FUSED", "1"))))
act = torch.randn(64, W, env.num_actions, device="cuda:0")
for i in range(40):
  env.step(act[i % 64])
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(100):
  env.step(act[i % 64])
e1.record(); torch.cuda.synchronize()
print("env.step ms", e0.elapsed_time(e1) / 100, "env-steps/s", W * 100 / (e0.elapsed_time(e1) * 1e-3))
torch.cuda.profiler.start()
env.step(act[0])
torch.cuda.synchronize()
torch.cuda.profiler.stop()
