"""Per-world solve time (cycles; needs a -DSOLVE_TIMING build via B2S_LIB) in the standing and the fallen regime."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import numpy as np, torch
import bench
from asimov_mjlab_b200 import engine
from asimov_mjlab_b200.sim import MujocoCfg, Simulation, SimulationCfg
E = int(os.environ.get("W", "4096"))
m = bench.load_model()
cfg = SimulationCfg(nconmax=50, njmax=300, mujoco=MujocoCfg(timestep=0.005, iterations=10, ls_iterations=20))
sim = Simulation(E, cfg, m, "cuda:0"); M, D = sim.model, sim.data
key = m.keys["robot/init_state"]
q0 = torch.tensor(key["qpos"], dtype=torch.float32, device="cuda"); c0 = torch.tensor(key["ctrl"], dtype=torch.float32, device="cuda")
pool = torch.as_tensor(bench.action_targets(m, 64, E, seed=1000), device="cuda")
D.qpos[:] = q0; D.qpos[:, 7:] += 0.02 * torch.randn(E, m.nq - 7, device="cuda"); D.ctrl[:] = c0; sim.forward()
for label, nsteps in (("standing", 10), ("fallen", 140)):
  for i in range(nsteps):
    D.ctrl.copy_(pool[i % 64]); engine.step(M, D, 4)
  torch.cuda.synchronize()
  st = engine.profile_step(M, D, 1)
  cyc = D.solver_cost.cpu().numpy().astype(np.float64); nit = D.solver_niter.cpu().numpy(); nefc = D.nefc.cpu().numpy()
  us = cyc / 1965.0
  print(f"[{label}] solve kernel {st['solve']*1e3:.0f} us; per-world us: mean {us.mean():.1f} p50 {np.median(us):.1f} p90 {np.quantile(us,0.9):.1f} p99 {np.quantile(us,0.99):.1f} max {us.max():.1f}; sum/148 SMs {us.sum()/148:.0f} us")
  for k in range(0, 11):
    sel = nit == k
    if sel.sum():
      print(f"   niter {k:2d}: {sel.sum():5d} worlds, mean nefc {nefc[sel].mean():5.1f}, mean {us[sel].mean():6.1f} us, max {us[sel].max():6.1f}")
