"""Small run that touches every kernel variant, for compute-sanitizer (one --tool per gpurun call):
  compute-sanitizer --tool memcheck  python tools/sanitize_run.py
  compute-sanitizer --tool racecheck python tools/sanitize_run.py
G1 (k_solve<5,0>, k_collision<2,0>), G1 with friction loss / elliptic cones / exact line search + CG (k_solve<5,1|3|7>),
Go1 on boxes (grid broadphase, k_collision<8,0>) and on height fields, Asimov-toe full collision (k_collision<8,1>: GJK / EPA),
the fixed-tendon model, overflowing nconmax / njmax, the allocation-free re-solve and the fused manager kernels."""
import copy
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import numpy as np
import torch

from asimov_mjlab_b200 import engine, scenes
from asimov_mjlab_b200.hostmodel import HostData
from util import asimov_states, g1_states, legged_states, rough_states

W = int(os.environ.get("W", "8"))
NSTEP = int(os.environ.get("NSTEP", "3"))


def run(tag, m, states, nconmax=50, njmax=300, opt=None, expand=()):
  qpos, qvel, ctrl = states
  m = copy.copy(m); m.opt = copy.copy(m.opt)
  for k, v in (opt or {}).items():
    setattr(m.opt, k, v)
  M = engine.put_model(m, "cuda:0")
  D = engine.put_data(m, HostData(m), len(qpos), nconmax, njmax, model=M)
  for f in expand:
    engine.expand_model_fields(M, len(qpos), [f])
  if "dof_frictionloss" in expand:
    M.dof_frictionloss[:, 6:] = 0.2
  D.qpos[:] = torch.as_tensor(qpos, dtype=torch.float32, device="cuda:0")
  D.qvel[:] = torch.as_tensor(qvel, dtype=torch.float32, device="cuda:0")
  D.ctrl[:] = torch.as_tensor(ctrl, dtype=torch.float32, device="cuda:0")
  engine.forward(M, D)
  engine.step(M, D, NSTEP)
  engine.solve_noalloc(M, D)
  _ = D.efc_J
  torch.cuda.synchronize()
  print(f"[{tag}] ok: ncon {int(D.ncon.sum())}, nefc {int(D.nefc.sum())}, niter max {int(D.solver_niter.max())}, finite {bool(torch.isfinite(D.qpos).all())}", flush=True)


g1 = scenes.load_asset("g1_flat")
run("g1", g1, g1_states(g1, W, seed=1), opt=dict(ls_parallel=True))
run("g1 overflow", g1, g1_states(g1, W, seed=41, spread=0.5, zlo=0.15, zhi=0.5), nconmax=5, njmax=14, opt=dict(ls_parallel=True))
run("g1 frictionloss", g1, g1_states(g1, W, seed=2), opt=dict(ls_parallel=True), expand=("dof_frictionloss", "geom_friction"))
run("g1 elliptic", g1, g1_states(g1, W, seed=3), opt=dict(ls_parallel=True, cone=1))
run("g1 exact ls", g1, g1_states(g1, W, seed=4), opt=dict(ls_parallel=False))
run("g1 cg euler", g1, g1_states(g1, W, seed=5), opt=dict(ls_parallel=False, solver=1, integrator=0))
go1r = scenes.load_asset("go1_rough")
run("go1 boxes", go1r, rough_states(go1r, W, seed=7, tumble=1.2), opt=dict(ls_parallel=True))
go1h = scenes.load_asset("go1_rough_hf")
qpos, qvel, ctrl = legged_states(go1h, W, seed=41)
org = np.asarray(go1h.meta["terrain_origins"])
for w in range(W):
  qpos[w, 0:2] = org[w % org.shape[0], (3 * w) % org.shape[1], :2]; qpos[w, 2] = max(org[w % org.shape[0], (3 * w) % org.shape[1], 2], 0.0) + 0.35
run("go1 hfield", go1h, (qpos, qvel, ctrl), opt=dict(ls_parallel=True))
asf = scenes.load_asset("asimov_toe_full_flat")
qpos, qvel, ctrl = asimov_states(asf, W, seed=19, spread=0.3, zlo=0.45, zhi=0.7)
rng = np.random.default_rng(19)
for w in range(W):
  ax = rng.normal(size=3); ax /= np.linalg.norm(ax); ang = rng.normal(0, 1.0)
  qpos[w, 3:7] = [np.cos(ang / 2), *(ax * np.sin(ang / 2))]
run("asimov full (GJK/EPA)", asf, (qpos, qvel, ctrl), opt=dict(ls_parallel=True))
if int(os.environ.get("ENV", "1")):
  from asimov_mjlab_b200.rl.velocity_env import G1VelocityFlatEnv
  env = G1VelocityFlatEnv(16, "cuda:0", seed=1, use_graph=False)
  for _ in range(3):
    env.step(torch.randn(16, env.num_actions, device="cuda:0"))
  torch.cuda.synchronize()
  print("[velocity env, fused manager kernels] ok", flush=True)
print("sanitize_run done", flush=True)
