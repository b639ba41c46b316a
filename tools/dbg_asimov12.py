import sys; sys.path[:0]=['/root/repo','/root/repo/oracle','/root/repo/tests']
import numpy as np, torch
from asimov_mjlab_b200 import scenes
from test_gpu_parity import _setup
from util import legged_states
m=scenes.load_asset("asimov_flat")
W=32
qpos,qvel,ctrl=legged_states(m,W,seed=3,spread=0.1,vel=0.2,zlo=0.725,zhi=0.74); qpos[:,3:7]=[1,0,0,0]
for it in (10,50):
  import copy
  mm=copy.copy(m); mm.opt=copy.copy(m.opt); mm.opt.iterations=it
  engine,M,D,orc,worlds=_setup(mm,W,qpos,qvel,ctrl,True)
  engine.forward(M,D)
  for w in worlds: orc.forward(w)
  g=D.qacc.cpu().numpy(); o=np.stack([w["qacc"] for w in worlds])
  err=np.abs(g-o).max(axis=1)/(1e-6+np.abs(o).max())
  print("iterations",it,"niter gpu",D.solver_niter.cpu().numpy().tolist()); print("niter orc",[int(w["solver_niter"][0]) for w in worlds]); print("err",np.round(err,5).tolist()); print("ncon",D.ncon.cpu().numpy().tolist())
  cg=D.solver_cost.cpu().numpy(); co=np.array([float(w["solver_cost"][0]) for w in worlds]); print("cost rel diff", np.round(np.abs(cg-co)/(1+np.abs(co)),6).tolist())
