"""Generate tests/golden/*.npz: seeded rollouts recorded from the float64 CPU oracle.

These are REGRESSION pins of this repository's own restatement (the reference's step arithmetic cannot be run
here: SURVEY.md §8c), committed so the GPU box -- which has no /root/reference -- and later rounds compare
against fixed numbers.  Usage: python tools/make_golden.py
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import numpy as np

import fixtures as fx
from asimov_mjlab_b200 import scenes
from oracle import Oracle
from util import asimov_states, g1_states, legged_states

OUT = os.path.join(ROOT, "tests", "golden")
os.makedirs(OUT, exist_ok=True)


def record(m, qpos, qvel, ctrl, nstep, every, nconmax, njmax, ls_parallel):
  orc = Oracle(m, "f64", nconmax, njmax, ls_parallel=ls_parallel)
  W = len(qpos)
  worlds = [orc.new_world(qpos[w], qvel[w], ctrl[w]) for w in range(W)]
  rec = {k: [] for k in ("qpos", "qvel", "qacc", "sensordata", "ncon", "nefc", "con_geom")}
  for s in range(nstep):
    for w in worlds:
      orc.step(w)
    if (s + 1) % every == 0 or s == 0:
      for k in ("qpos", "qvel", "qacc", "sensordata"):
        rec[k].append(np.stack([w[k] for w in worlds]))
      rec["ncon"].append(np.array([w["ncon"][0] for w in worlds]))
      rec["nefc"].append(np.array([w["nefc"][0] for w in worlds]))
      rec["con_geom"].append(np.stack([w["con_geom"].copy() for w in worlds]))
  return {k: np.stack(v) for k, v in rec.items()}


g1 = scenes.load_asset("g1_flat")
for lsp in (0, 1):
  qpos, qvel, ctrl = g1_states(g1, 4, seed=21, spread=0.1, vel=0.1, zlo=0.74, zhi=0.77)
  r = record(g1, qpos, qvel, ctrl, 40, 10, 50, 300, lsp)
  np.savez_compressed(os.path.join(OUT, f"g1_rollout_ls{lsp}.npz"), qpos0=qpos, qvel0=qvel, ctrl0=ctrl, **r)
go1 = scenes.load_asset("go1_flat")
qpos, qvel, ctrl = legged_states(go1, 4, seed=22, spread=0.1, vel=0.1, zlo=0.26, zhi=0.29)
r = record(go1, qpos, qvel, ctrl, 40, 10, 50, 300, 1)
np.savez_compressed(os.path.join(OUT, "go1_flat_rollout_ls1.npz"), qpos0=qpos, qvel0=qvel, ctrl0=ctrl, **r)
asv = scenes.load_asset("asimov_toe_flat")
# short horizon: saturating toe actuators (5 Nm on ~1.5e-4 kg m^2) make this model chaotic within ~10 steps
qpos, qvel, ctrl = asimov_states(asv, 4, seed=23, spread=0.05, vel=0.1, zlo=0.725, zhi=0.75)
c0 = np.array(asv.keys["robot/init_state"]["ctrl"])
ctrl = c0 + 0.2 * (ctrl - c0)
r = record(asv, qpos, qvel, ctrl, 8, 4, 50, 300, 1)
np.savez_compressed(os.path.join(OUT, "asimov_toe_flat_rollout_ls1.npz"), qpos0=qpos, qvel0=qvel, ctrl0=ctrl, **r)
m = fx.compile_xml(fx.LEGGED)
rng = np.random.default_rng(3)
qpos = np.tile(m.qpos0, (3, 1)); qpos[:, 2] = [0.3, 0.33, 0.36]
qvel = rng.normal(0, 0.2, (3, m.nv)); ctrl = rng.normal(0, 0.3, (3, m.nu))
r = record(m, qpos, qvel, ctrl, 100, 25, 16, 64, 0)
np.savez_compressed(os.path.join(OUT, "legged_rollout.npz"), qpos0=qpos, qvel0=qvel, ctrl0=ctrl, **r)
print("golden written:", sorted(os.listdir(OUT)))
