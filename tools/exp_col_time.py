"""Where k_collision spends its time (needs a -DCOL_TIMING build via B2S_LIB): per-world phase cycles and pair counts."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import numpy as np, torch
import bench
from asimov_mjlab_b200 import engine
W = int(os.environ.get("W", "4096"))
for wl in os.environ.get("WL", "asimov_toe_full_flat,go1_rough_hf").split(","):
  pb = bench.PhysBench(wl, W, "cuda:0", 0)
  pb.reset_all()
  for i in range(1, int(os.environ.get('NSTEP', '12'))):
    pb.one_step(i)
  engine.forward(pb.M, pb.D)
  torch.cuda.synchronize()
  t = pb.D.efc_force[:, -8:].cpu().numpy().astype(np.float64)
  us = t[:, :3] / 1965.0
  print(f"[{wl}] per world: broadphase {us[:,0].mean():.1f} us, narrowphase {us[:,1].mean():.1f} us (max {us[:,1].max():.0f}), of which convex-pair loop {us[:,2].mean():.1f} us; "
        f"survivors {t[:,3].mean():.1f} (max {t[:,3].max():.0f}), convex pairs {t[:,4].mean():.1f} (max {t[:,4].max():.0f}), with a contact {t[:,5].mean():.2f}; ncon {pb.D.ncon.float().mean():.2f}")
  tot = us[:, 0] + us[:, 1]
  print("   per-world total us percentiles 50/90/99/99.9/max:", [round(float(np.percentile(tot, q)), 1) for q in (50, 90, 99, 99.9, 100)], "sum over worlds %.1f ms" % (tot.sum() / 1e3),
        " max-of-4 (CTA) mean %.1f us; worlds > 60 us: %d" % (tot.reshape(-1, 4).max(axis=1).mean(), (tot > 60).sum()))
  hv = np.argsort(-tot)[:5]
  print("   heaviest worlds:", [(int(w), round(float(tot[w]), 1), int(t[w, 3]), int(pb.D.ncon[w].item())) for w in hv], "(world, us, survivors, ncon)")
  if "cvx_task" in pb.D.field_names and os.environ.get("TASKS", "1") == "1":
    nl, nh = int(pb.D.cvx_ntask.flatten()[0].item()), int(pb.D.cvx_ntask.flatten()[2].item()); nt = nl + nh
    if nt:
      allrec = pb.D.cvx_task.reshape(-1, 8)[:, :2]; rec = torch.cat([allrec[:nl], allrec[allrec.shape[0] - nh:]]).cpu().numpy(); print(f'   expensive-first tasks: {nh}'); cyc = (rec[:, 1] & 0xffff).astype(np.float64) * 64 / 1965.0
      gjk = rec[:, 0] & 255; epa = (rec[:, 0] >> 8) & 255; nvt = (rec[:, 0] >> 16) & 4095; epa_us = ((rec[:, 1] >> 16) & 0xffff).astype(np.float64) * 64 / 1965.0
      print(f"   convex tasks: {nt}; per task us 50/90/99/99.9/max:", [round(float(np.percentile(cyc, q)), 1) for q in (50, 90, 99, 99.9, 100)], "sum %.1f ms" % (cyc.sum() / 1e3))
      top = np.argsort(-cyc)[:8]
      print("   slowest tasks (us, gjk iterations, epa rounds, epa us, hull vertices):", [(round(float(cyc[t]), 1), int(gjk[t]), int(epa[t]), round(float(epa_us[t]), 1), int(nvt[t])) for t in top])
      print("   tasks with EPA: %d (%.1f ms of which EPA %.1f ms); GJK iterations mean %.1f, > 20: %d" % ((epa > 0).sum(), cyc[epa > 0].sum() / 1e3, epa_us.sum() / 1e3, gjk.mean(), (gjk > 20).sum()))
  print("   stage ms", engine.profile_step(pb.M, pb.D, 5))
