"""Engine-vs-oracle comparison on a GPU box (development aid; the pytest versions live in tests/)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import numpy as np
import torch

from asimov_mjlab_b200 import engine, scenes
from asimov_mjlab_b200.hostmodel import HostData
from oracle import Oracle
from util import g1_states

m = scenes.load_asset("g1_flat")
m.opt.ls_parallel = bool(int(os.environ.get("LSP", "1")))
W = int(os.environ.get("W", "16"))
nstep = int(os.environ.get("NSTEP", "20"))
qpos, qvel, ctrl = g1_states(m, W, seed=1)
M = engine.put_model(m)
D = engine.put_data(m, HostData(m), W, 50, 300, model=M)
D.qpos[:] = torch.as_tensor(qpos, dtype=torch.float32, device="cuda")
D.qvel[:] = torch.as_tensor(qvel, dtype=torch.float32, device="cuda")
D.ctrl[:] = torch.as_tensor(ctrl, dtype=torch.float32, device="cuda")
orc = Oracle(m, "f32", 50, 300, ls_parallel=int(m.opt.ls_parallel))
worlds = [orc.new_world(qpos[w], qvel[w], ctrl[w]) for w in range(W)]
FIELDS = ["xpos", "xquat", "xipos", "geom_xpos", "site_xpos", "subtree_com", "cinert", "cdof", "cvel", "cdof_dot", "qfrc_bias",
          "qfrc_passive", "actuator_force", "qfrc_smooth", "qacc_smooth", "qacc", "qfrc_constraint", "cacc", "sensordata", "qpos", "qvel"]


def compare(tag):
  torch.cuda.synchronize()
  worst = {}
  for f in FIELDS:
    g = getattr(D, f).cpu().numpy().reshape(W, -1)
    o = np.stack([w[f] for w in worlds]).reshape(W, -1)
    err = np.abs(g - o).max(axis=1) / (1e-6 + np.abs(o).max())
    worst[f] = (err.max(), int(err.argmax()))
  ncon_g = D.ncon.cpu().numpy(); nefc_g = D.nefc.cpu().numpy()
  ncon_o = np.array([w["ncon"][0] for w in worlds]); nefc_o = np.array([w["nefc"][0] for w in worlds])
  nit_g = D.solver_niter.cpu().numpy(); nit_o = np.array([w["solver_niter"][0] for w in worlds])
  print(f"[{tag}] ncon match {np.array_equal(ncon_g, ncon_o)} nefc match {np.array_equal(nefc_g, nefc_o)} niter g {nit_g.tolist()} o {nit_o.tolist()}")
  if not np.array_equal(ncon_g, ncon_o):
    print("   ncon g", ncon_g.tolist(), "o", ncon_o.tolist())
  print("   " + "  ".join(f"{k}:{v[0]:.1e}@{v[1]}" for k, v in worst.items()))


engine.forward(M, D)
for w in worlds:
  orc.forward(w)
compare("forward")
for s in range(nstep):
  engine.step(M, D)
  for w in worlds:
    orc.step(w)
  if s in (0, 1, 4, 9, 19, 49, 99) or s == nstep - 1:
    compare(f"step {s + 1}")
print("stage ms:", engine.profile_step(M, D, 5))
