"""Per-step device time of the bench loop right after start-up (why do 20 steps run slower than 300?)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import numpy as np, torch
import bench
from asimov_mjlab_b200 import engine
from asimov_mjlab_b200.sim import MujocoCfg, Simulation, SimulationCfg
E = 4096
m = bench.load_model()
cfg = SimulationCfg(nconmax=50, njmax=300, mujoco=MujocoCfg(timestep=0.005, iterations=10, ls_iterations=20))
sim = Simulation(E, cfg, m, "cuda:0"); M, D = sim.model, sim.data
sim.expand_model_fields(("geom_friction",))
sim.create_graph(decimation=4)
key = m.keys["robot/init_state"]
q0 = torch.tensor(key["qpos"], dtype=torch.float32, device="cuda"); c0 = torch.tensor(key["ctrl"], dtype=torch.float32, device="cuda")
pool = torch.as_tensor(bench.action_targets(m, 64, E, seed=1000), device="cuda")
noise = 0.02 * torch.randn(E, m.nq - 7, device="cuda")
def reset():
  D.qpos[:] = q0; D.qpos[:, 7:] += noise; D.qvel.zero_(); D.qacc_warmstart.zero_(); D.time.zero_(); D.ctrl[:] = c0; sim.forward()
for trial in range(2):
  reset()
  N = 300
  ev = [torch.cuda.Event(enable_timing=True) for _ in range(N + 1)]
  ev[0].record()
  stats = []
  for i in range(N):
    D.ctrl.copy_(pool[i % 64]); sim.step_decimation(); ev[i + 1].record()
    if i in (0, 5, 10, 20, 40, 80, 150, 299):
      stats.append((i, float(D.nefc.float().mean()), float(D.solver_niter.float().mean()), float((D.qpos[:, 2] < 0.4).float().mean())))
  torch.cuda.synchronize()
  t = [ev[i].elapsed_time(ev[i + 1]) for i in range(N)]
  print("trial", trial, "ms/step:", " ".join(f"{i}:{np.mean(t[i:j]):.3f}" for i, j in ((0, 1), (1, 5), (5, 10), (10, 20), (20, 40), (40, 80), (80, 150), (150, 300))))
  print("   (step, mean nefc, mean niter, fallen):", stats)
reset()
for i in range(10):
  D.ctrl.copy_(pool[i % 64]); sim.step_decimation()
torch.cuda.synchronize()
print("standing regime: stage ms", engine.profile_step(M, D, 8))
print("nefc hist/8", torch.bincount(D.nefc.cpu() // 8).tolist()); print("niter hist", torch.bincount(D.solver_niter.cpu()).tolist()); print("ncon hist", torch.bincount(D.ncon.cpu()).tolist()); print("nl mean", float(D.nl.float().mean()))
