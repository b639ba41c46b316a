"""Committed golden rollouts (tests/golden/, made by tools/make_golden.py from the float64 oracle).

CPU: the float64 oracle reproduces them exactly and the float32 oracle within fp32 tolerance.
GPU (marked): the CUDA path reproduces them through the C ABI.
"""
import os

import numpy as np
import pytest

import fixtures as fx
from oracle import Oracle

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = [("g1_rollout_ls0.npz", "g1", 0, 50, 300, 40, 10), ("g1_rollout_ls1.npz", "g1", 1, 50, 300, 40, 10),
         ("legged_rollout.npz", "legged", 0, 16, 64, 100, 25), ("go1_flat_rollout_ls1.npz", "go1", 1, 50, 300, 40, 10),
         ("asimov_toe_flat_rollout_ls1.npz", "asimov_toe", 1, 50, 300, 8, 4)]


def _model(kind, g1):
  if kind == "g1":
    return g1
  if kind in ("go1", "asimov_toe"):
    from asimov_mjlab_b200 import scenes
    return scenes.load_asset(kind + "_flat")
  return fx.compile_xml(fx.LEGGED)


def _rollout_oracle(m, z, dtype, lsp, nconmax, njmax, nstep, every):
  orc = Oracle(m, dtype, nconmax, njmax, ls_parallel=lsp)
  worlds = [orc.new_world(z["qpos0"][w], z["qvel0"][w], z["ctrl0"][w]) for w in range(len(z["qpos0"]))]
  out = []
  for s in range(nstep):
    for w in worlds:
      orc.step(w)
    if (s + 1) % every == 0 or s == 0:
      out.append({k: np.stack([w[k] for w in worlds]) for k in ("qpos", "qvel", "qacc", "sensordata")}
                 | {"ncon": np.array([w["ncon"][0] for w in worlds]), "nefc": np.array([w["nefc"][0] for w in worlds]),
                    "con_geom": np.stack([w["con_geom"].copy() for w in worlds])})
  return out


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_oracle_f64_reproduces_golden_exactly(case, g1):
  fn, kind, lsp, ncm, njm, nstep, every = case
  z = np.load(os.path.join(GOLD, fn))
  out = _rollout_oracle(_model(kind, g1), z, "f64", lsp, ncm, njm, nstep, every)
  for i, o in enumerate(out):
    assert np.array_equal(o["ncon"], z["ncon"][i]) and np.array_equal(o["nefc"], z["nefc"][i])
    assert np.array_equal(o["con_geom"], z["con_geom"][i])
    assert np.allclose(o["qpos"], z["qpos"][i], rtol=0, atol=1e-12) and np.allclose(o["qvel"], z["qvel"][i], rtol=0, atol=1e-10)


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_oracle_f32_tracks_golden(case, g1):
  fn, kind, lsp, ncm, njm, nstep, every = case
  z = np.load(os.path.join(GOLD, fn))
  out = _rollout_oracle(_model(kind, g1), z, "f32", lsp, ncm, njm, nstep, every)
  # first record is after ONE step: fp32 vs fp64 of the same algorithm
  assert np.array_equal(out[0]["ncon"], z["ncon"][0]) and np.array_equal(out[0]["nefc"], z["nefc"][0])
  # the grid line search picks among discrete step sizes, so fp32-vs-fp64 differences are ~100x larger with it
  tq, tv = (2e-4, 3e-2) if lsp else (2e-6, 2e-3)
  assert np.abs(out[0]["qpos"] - z["qpos"][0]).max() < tq
  assert np.abs(out[0]["qvel"] - z["qvel"][0]).max() < tv
  assert np.abs(out[-1]["qpos"] - z["qpos"][-1]).max() < 5e-3


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_cuda_path_tracks_golden(case, g1):
  import torch
  from asimov_mjlab_b200 import engine
  from asimov_mjlab_b200.hostmodel import HostData
  fn, kind, lsp, ncm, njm, nstep, every = case
  z = np.load(os.path.join(GOLD, fn))
  m = _model(kind, g1)
  m.opt.ls_parallel = bool(lsp)
  W = len(z["qpos0"])
  M = engine.put_model(m)
  D = engine.put_data(m, HostData(m), W, ncm, njm, model=M)
  for k in ("qpos", "qvel", "ctrl"):
    getattr(D, k)[:] = torch.as_tensor(z[k + "0"], dtype=torch.float32, device="cuda:0")
  engine.step(M, D)
  assert np.array_equal(D.ncon.cpu().numpy(), z["ncon"][0]) and np.array_equal(D.nefc.cpu().numpy(), z["nefc"][0])
  n = z["ncon"][0]
  g = D.contact_geom.cpu().numpy()
  for w in range(W):
    assert g[w, : n[w]].reshape(-1).tolist() == z["con_geom"][0][w][: 2 * n[w]].tolist()
  tq, tv = (2e-4, 3e-2) if lsp else (2e-6, 2e-3)
  assert np.abs(D.qpos.cpu().numpy() - z["qpos"][0]).max() < tq
  assert np.abs(D.qvel.cpu().numpy() - z["qvel"][0]).max() < tv
  assert np.abs(D.qacc.cpu().numpy() - z["qacc"][0]).max() < (2e-2 if lsp else 2e-3) * (1 + np.abs(z["qacc"][0]).max())
  engine.step(M, D, nstep - 1)
  if kind == "asimov_toe":
    # chaotic within a few steps (Newton runs truncated at the iteration cap + saturating toe actuators): the
    # first step above is the pin; at the end only the slow coordinates (base position) are compared
    assert np.abs(D.qpos.cpu().numpy()[:, :3] - z["qpos"][-1][:, :3]).max() < 2e-2
    return
  assert np.abs(D.qpos.cpu().numpy() - z["qpos"][-1]).max() < 5e-3
