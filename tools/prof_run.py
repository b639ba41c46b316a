"""Short profiling workload: G1-flat, W worlds, warm into contact, then a few eager steps (for ncu)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import torch

from asimov_mjlab_b200 import engine, scenes
from asimov_mjlab_b200.hostmodel import HostData

W = int(os.environ.get("W", "4096"))
warm = int(os.environ.get("WARM", "60"))
steps = int(os.environ.get("STEPS", "4"))
asset = os.environ.get("ASSET", "g1_flat")
m = scenes.load_asset(asset)
m.opt.ls_parallel = bool(int(os.environ.get("LSP", "1")))
M = engine.put_model(m)
key = m.keys["robot/init_state"]
hd = HostData(m)
hd.qpos[:] = key["qpos"]; hd.ctrl[:] = key["ctrl"]
D = engine.put_data(m, hd, W, int(os.environ.get("NCONMAX", "50")), int(os.environ.get("NJMAX", "300")), model=M)
g = torch.Generator().manual_seed(0)
if "terrain_origins" in m.meta:       # spawn on the generated terrain like bench.py does
  import numpy as np
  from asimov_mjlab_b200 import terrains
  eo, _, _ = terrains.env_origins_curriculum(W, np.asarray(m.meta["terrain_origins"]), seed=0)
  org = torch.as_tensor(eo, dtype=torch.float32)
  org[:, 0:2] += (torch.rand(W, 2, generator=g) - 0.5) * 2.0
  xy = org[:, 0:2].numpy()
  surf = np.max([terrains.surface_height(m, xy + np.array(o)) for o in ((0, 0), (0.3, 0.2), (0.3, -0.2), (-0.3, 0.2), (-0.3, -0.2))], axis=0)
  org[:, 2] = torch.as_tensor(surf, dtype=torch.float32) + 0.05
  D.qpos[:, 0:3] += org.cuda()
D.qpos[:, 7:] += (0.05 * torch.randn(W, m.nq - 7, generator=g)).cuda()
D.ctrl[:] += (0.1 * torch.randn(W, m.nu, generator=g)).cuda()
if not int(os.environ.get("STAND", "0")):
  engine.step(M, D, warm)
if int(os.environ.get("STAND", "0")):     # the bench's first 40 steps / the RL regime: robots still upright, both feet loaded (nefc ~ 50, 5 Newton iterations)
  import bench
  key = m.keys["robot/init_state"]
  D.qpos[:] = torch.tensor(key["qpos"], dtype=torch.float32, device="cuda")
  D.qpos[:, 7:] += (0.02 * torch.randn(W, m.nq - 7, generator=g)).cuda()
  D.qvel.zero_(); D.qacc_warmstart.zero_(); D.ctrl[:] = torch.tensor(key["ctrl"], dtype=torch.float32, device="cuda")
  engine.forward(M, D)
  pool = torch.as_tensor(bench.action_targets(m, 16, W, seed=1000)).cuda()
  for k in range(10):
    D.ctrl.copy_(pool[k]); engine.step(M, D, 4)
elif int(os.environ.get("RAND", "0")):      # the bench's state distribution: random action targets, resampled every 4 sub-steps
  import bench
  pool = torch.as_tensor(bench.action_targets(m, 40, W, seed=1000)).cuda()
  for k in range(40):
    D.ctrl.copy_(pool[k]); engine.step(M, D, 4)
torch.cuda.synchronize()
print("ncon mean %.1f nefc mean %.1f niter mean %.2f" % (D.ncon.float().mean().item(), D.nefc.float().mean().item(), D.solver_niter.float().mean().item()))
print("nefc hist", torch.bincount(D.nefc.cpu() // 8).tolist())
print("niter hist", torch.bincount(D.solver_niter.cpu()).tolist())
print("ncon hist", torch.bincount(D.ncon.cpu()).tolist())
engine.step(M, D, steps)
torch.cuda.synchronize()
print("stage ms", engine.profile_step(M, D, 10))
