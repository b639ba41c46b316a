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
m = scenes.load_asset("g1_flat")
m.opt.ls_parallel = bool(int(os.environ.get("LSP", "1")))
M = engine.put_model(m)
key = m.keys["robot/init_state"]
hd = HostData(m)
hd.qpos[:] = key["qpos"]; hd.ctrl[:] = key["ctrl"]
D = engine.put_data(m, hd, W, 50, 300, model=M)
g = torch.Generator().manual_seed(0)
D.qpos[:, 7:] += (0.05 * torch.randn(W, m.nq - 7, generator=g)).cuda()
D.ctrl[:] += (0.1 * torch.randn(W, m.nu, generator=g)).cuda()
engine.step(M, D, warm)
torch.cuda.synchronize()
print("ncon mean %.1f nefc mean %.1f niter mean %.2f" % (D.ncon.float().mean().item(), D.nefc.float().mean().item(), D.solver_niter.float().mean().item()))
print("nefc hist", torch.bincount(D.nefc.cpu() // 8).tolist())
print("niter hist", torch.bincount(D.solver_niter.cpu()).tolist())
print("ncon hist", torch.bincount(D.ncon.cpu()).tolist())
engine.step(M, D, steps)
torch.cuda.synchronize()
print("stage ms", engine.profile_step(M, D, 10))
