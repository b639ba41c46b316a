import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import numpy as np, torch
np.set_printoptions(precision=4, suppress=True, linewidth=220)
from asimov_mjlab_b200 import scenes
from util import asimov_states
import test_gpu_parity as T
from oracle import Oracle
m = scenes.load_asset("asimov_toe_flat")
W = 32
qpos, qvel, ctrl = asimov_states(m, W, seed=9, spread=0.2, vel=0.3, zlo=0.70, zhi=0.76)
for lsp in (1, 0):
  for k in (1, 2, 3, 4, 6, 10, 20, 50):
    m.opt.iterations = k
    engine, M, D, orc, worlds = T._setup(m, W, qpos, qvel, ctrl, bool(lsp))
    o64 = Oracle(m, "f64", D.nconmax, D.njmax, ls_parallel=lsp)
    engine.forward(M, D)
    ws = [21, 5]
    for w in ws:
      orc.forward(worlds[w])
      x = o64.new_world(qpos[w], qvel[w], ctrl[w]); o64.forward(x)
      g = D.qacc.cpu().numpy()[w]
      print("lsp", lsp, "iters", k, "world", w, "niter g/o32/o64", D.solver_niter.cpu().numpy()[w], worlds[w]["solver_niter"][0], x["solver_niter"][0],
            "cost", D.solver_cost.cpu().numpy()[w], worlds[w]["solver_cost"][0], x["solver_cost"][0],
            "dq g-o64 %.4g o32-o64 %.4g" % (np.abs(g - x["qacc"]).max(), np.abs(worlds[w]["qacc"] - x["qacc"]).max()))
