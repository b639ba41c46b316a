import os, sys, copy
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")]
import numpy as np
from asimov_mjlab_b200 import scenes
from util import g1_states
from test_gpu_parity import _setup
g1 = scenes.load_asset("g1_flat")
W = 16
qpos, qvel, ctrl = g1_states(g1, W, seed=5)
for lsp in (True, False):
  m = copy.copy(g1); m.opt = copy.copy(g1.opt); m.opt.solver, m.opt.iterations = 1, 30
  engine, M, D, orc, worlds = _setup(m, W, qpos, qvel, ctrl, lsp)
  engine.forward(M, D)
  for w in worlds: orc.forward(w)
  qa_g, qa_o = D.qacc.cpu().numpy(), np.stack([w["qacc"] for w in worlds])
  print("lsp", lsp)
  print("nit g", D.solver_niter.cpu().numpy().tolist()); print("nit o", [int(w["solver_niter"][0]) for w in worlds])
  print("cost g", np.round(D.solver_cost.cpu().numpy(), 1).tolist()); print("cost o", [round(float(w["solver_cost"][0]), 1) for w in worlds])
  print("qacc err", np.round(np.abs(qa_g - qa_o).max(1), 3).tolist(), "scale", np.abs(qa_o).max())
