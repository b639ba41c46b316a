"""b2s_solve_noalloc (include/b200sim.h): the allocation-free variant of the constraint solve, scheduling scratch in a fixed
device array.  Collected last on purpose.  Through the C ABI, against the existing entry point
(b2s_forward), on two calls in a row with different states and world counts, plus the capacity check."""
import numpy as np
import pytest

import fixtures as fx
from test_gpu_parity import TOL_FORWARD, _setup
from util import g1_states

pytestmark = pytest.mark.gpu

FIELDS = ("qacc", "qfrc_constraint", "efc_force", "solver_cost")


def _resolve_and_compare(engine, M, D):
  """forward() through the existing entry point, keep its solve outputs, wipe them, re-solve without allocating, compare."""
  import torch
  engine.forward(M, D)
  torch.cuda.synchronize()
  ref = {f: getattr(D, f).clone() for f in FIELDS}
  nit = D.solver_niter.clone()
  for f in FIELDS:
    getattr(D, f).fill_(float("nan"))
  D.solver_niter.fill_(-1)
  before = torch.cuda.memory_allocated(), engine.launch_count()
  engine.solve_noalloc(M, D)
  torch.cuda.synchronize()
  assert engine.launch_count() - before[1] == (3 if D.njmax > M.opt.solve_rowcap else 2)     # queue kernel + k_solve (+ its launch for worlds with many rows): the CUDA path ran
  assert torch.cuda.memory_allocated() == before[0]
  assert torch.equal(D.solver_niter, nit)
  active = torch.arange(D.njmax, device="cuda:0")[None, :] < D.nefc[:, None]      # rows beyond nefc are not part of the result
  for f in FIELDS:
    g, o = getattr(D, f).double(), ref[f].double()
    if f == "efc_force":
      g, o = torch.where(active, g, 0.0), torch.where(active, o, 0.0)
    assert torch.isfinite(g).all(), f
    tol = TOL_FORWARD.get(f, 5e-4)
    assert float((g - o).abs().max()) <= tol * (1e-6 + float(o.abs().max())), f
  return int(D.nefc.sum())


def test_solve_noalloc_matches_forward_on_two_calls_in_a_row(g1):
  import torch
  W = 96
  qpos, qvel, ctrl = g1_states(g1, W, seed=12, spread=0.3)
  engine, M, D, _, _ = _setup(g1, W, qpos, qvel, ctrl, True)
  assert _resolve_and_compare(engine, M, D) > 500         # contact-rich: the solve did real work
  # second call: other states (the queue classes change), the array still holds the first call's queue
  engine.step(M, D, 7)
  D.ctrl[:] += 0.3 * torch.randn(W, g1.nu, generator=torch.Generator().manual_seed(0)).to("cuda:0")
  assert _resolve_and_compare(engine, M, D) > 500
  # third: FEWER worlds of another model on the same device -- stale queue entries beyond its world count must not be read
  m2 = fx.compile_xml(fx.LEGGED)
  q2 = np.tile(m2.qpos0, (5, 1)); q2[:, 2] = np.linspace(0.2, 0.3, 5)
  e2, M2, D2, _, _ = _setup(m2, 5, q2, None, None, True, nconmax=16, njmax=64)
  assert _resolve_and_compare(e2, M2, D2) > 0
  assert _resolve_and_compare(engine, M, D) > 500         # and the first data again


def test_solve_noalloc_rejects_sizes_beyond_the_static_capacity():
  from asimov_mjlab_b200 import engine
  m = fx.compile_xml(fx.SPHERE_DROP)
  cap = engine.noalloc_max_worlds()
  assert cap == 65536
  _, M, D, _, _ = _setup(m, 4, None, None, None, True, nconmax=4, njmax=8)
  engine.forward(M, D)
  engine.solve_noalloc(M, D)                               # within capacity
  M2 = engine.put_model(m, "cuda:0")
  from asimov_mjlab_b200.hostmodel import HostData
  D2 = engine.put_data(m, HostData(m), cap + 1, 4, 8, model=M2)
  engine.forward(M2, D2)
  n0 = engine.launch_count()
  with pytest.raises(engine.BackendError, match="exceeds the static scratch capacity"):
    engine.solve_noalloc(M2, D2)
  assert engine.launch_count() == n0                       # nothing launched
