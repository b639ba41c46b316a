import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import numpy as np, torch
np.set_printoptions(precision=6, suppress=True, linewidth=220)
import fixtures as fx
import test_gpu_parity as T
from oracle import Oracle
ground = sys.argv[1] if len(sys.argv) > 1 else "plane"
rng = np.random.default_rng(31)
nrow = ncol = 17
elev = (0.5 + 0.5 * np.sin(np.linspace(0, 7, nrow))[:, None] * np.cos(np.linspace(0, 5, ncol))[None, :]).reshape(-1).tolist()
movers = ("mesh", "box", "sphere", "capsule", "mesh")
m = fx.compile_xml(fx.convex_scene(ground, movers, hfield_elev=elev, hfield_size=(1.0, 1.0, 0.15, 0.1), nrow=nrow, ncol=ncol, margin=0.004))
W = 16
qpos = np.tile(m.qpos0, (W, 1)); qvel = np.zeros((W, m.nv)); ctrl = np.zeros((W, m.nu))
for w in range(W):
  for k in range(len(movers)):
    ax = rng.normal(size=3); ax /= np.linalg.norm(ax); ang = rng.normal(0, 0.8)
    qpos[w, 7 * k: 7 * k + 3] = [rng.uniform(-0.6, 0.6), rng.uniform(-0.6, 0.6), rng.uniform(0.05, 0.3) + (0.1 if ground == "hfield" else 0.0)]
    qpos[w, 7 * k + 3: 7 * k + 7] = [np.cos(ang / 2), *(ax * np.sin(ang / 2))]
  qvel[w] = rng.normal(0, 0.5, m.nv)
engine, M, D, orc, worlds = T._setup(m, W, qpos, qvel, ctrl, True, nconmax=48, njmax=200)
o64 = Oracle(m, "f64", 48, 200, ls_parallel=1)
nbad = 0
for chunk in range(10):
  if chunk:
    for w in worlds: orc.step(w, 10)
    for f in ("qpos", "qvel", "qacc_warmstart"):
      getattr(D, f)[:] = torch.as_tensor(np.stack([w[f] for w in worlds]), dtype=torch.float32, device="cuda:0")
  engine.forward(M, D)
  for w in worlds: orc.forward(w)
  n_all = D.ncon.cpu().numpy()
  G = D.contact_geom.cpu().numpy(); P = D.contact_pos.cpu().numpy(); Fr = D.contact_frame.cpu().numpy().reshape(W, -1, 9); Di = D.contact_dist.cpu().numpy()
  for i, w in enumerate(worlds):
    n = int(w["ncon"][0])
    if int(n_all[i]) != n or G[i, :n].reshape(-1).tolist() != w["con_geom"][: 2 * n].tolist():
      print("chunk", chunk, "world", i, "LIST MISMATCH gpu", n_all[i], G[i, :n_all[i]].reshape(-1).tolist(), "orc", n, w["con_geom"][:2*n].tolist())
      continue
    x = o64.new_world(w["qpos"].astype(np.float64), w["qvel"].astype(np.float64)); o64.forward(x)
    for c in range(n):
      dp = np.abs(P[i, c] - w["con_pos"][3*c:3*c+3]).max(); dd = abs(Di[i, c] - w["con_dist"][c]); dn = np.abs(Fr[i, c, :3] - w["con_frame"][9*c:9*c+3]).max()
      if dp > 2e-4 or dd > 2e-5 or dn > 2e-3:
        nbad += 1
        g1, g2 = w["con_geom"][2*c:2*c+2]
        print("chunk", chunk, "world", i, "c", c, "types", m.geom_type[g1], m.geom_type[g2], "dd %.2e dp %.2e dn %.2e" % (dd, dp, dn))
        print("   gpu dist %.6f pos %s n %s" % (Di[i, c], P[i, c], Fr[i, c, :3]))
        print("   o32 dist %.6f pos %s n %s" % (w["con_dist"][c], w["con_pos"][3*c:3*c+3], w["con_frame"][9*c:9*c+3]))
        if int(x["ncon"][0]) == n: print("   o64 dist %.6f pos %s n %s" % (x["con_dist"][c], x["con_pos"][3*c:3*c+3], x["con_frame"][9*c:9*c+3]))
print("bad", nbad)
