"""Solver row capacity sweep (Model.opt.solve_rowcap) on one workload: rate, rows per world, share of worlds in the second launch."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import torch
import bench
from asimov_mjlab_b200 import engine
for wl in os.environ.get("WL", "asimov_toe_flat,g1_flat").split(","):
  for cap in [int(c) for c in os.environ.get("CAPS", "128,100000,64,192").split(",")]:
    pb = bench.PhysBench(wl, 4096, "cuda:0", 0)
    pb.M.opt.solve_rowcap = cap
    pb.sim.create_graph(decimation=bench.DECIMATION)
    ms, _ = pb.timed(40, 10)
    nefc = pb.D.nefc.float()
    print(f"[{wl}] rowcap {cap}: {pb.rate(ms, 40) / 1e6:.3f} M sub-steps/s; nefc mean {nefc.mean().item():.1f} p99 {nefc.quantile(0.99).item():.0f} max {nefc.max().item():.0f}; worlds over the cap {(nefc > cap).float().mean().item() * 100:.1f} %; niter mean {pb.D.solver_niter.float().mean().item():.2f}",
          "stage ms", {k: round(v * 1e3, 1) for k, v in engine.profile_step(pb.M, pb.D, 5).items() if k == "solve"})
    del pb
    torch.cuda.empty_cache()
