"""`Simulation`: the drop-in seam (reference src/mjlab/sim/sim.py:41-199), same
names, argument meaning and error behaviour, driving the B200 backend.

Differences from the reference, all behind the same public surface:
  * graphs are `torch.cuda.CUDAGraph`s captured on torch's current stream (the
    reference captures on Warp's stream, sim_data.py:36-47); writes, launches
    and reads are therefore ordered without an external-stream shim;
  * `create_graph(decimation=n)` can capture n sub-steps as one graph
    (north_star: "a single CUDA graph per decimation");
  * `expand_model_fields` re-captures the graphs itself.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Literal

import numpy as np

from . import engine
from .hostmodel import HostData

_INTEGRATOR = {"euler": 0, "implicitfast": 3}
_CONE = {"pyramidal": 0, "elliptic": 1}
_SOLVER = {"pgs": 0, "cg": 1, "newton": 2}
_JACOBIAN = {"dense": 0, "sparse": 1, "auto": 2}


@dataclass
class MujocoCfg:
  """MuJoCo options (field-for-field the reference's MujocoCfg, sim.py:41-76)."""
  timestep: float = 0.002
  integrator: Literal["euler", "implicitfast"] = "implicitfast"
  impratio: float = 1.0
  cone: Literal["pyramidal", "elliptic"] = "pyramidal"
  jacobian: Literal["auto", "dense", "sparse"] = "auto"
  solver: Literal["newton", "cg", "pgs"] = "newton"
  iterations: int = 100
  tolerance: float = 1e-8
  ls_iterations: int = 50
  ls_tolerance: float = 0.01
  gravity: tuple = (0, 0, -9.81)

  def apply(self, model) -> None:
    o = model.opt
    o.jacobian = _JACOBIAN[self.jacobian]
    o.cone = _CONE[self.cone]
    o.integrator = _INTEGRATOR[self.integrator]
    o.solver = _SOLVER[self.solver]
    o.timestep = self.timestep
    o.impratio = self.impratio
    o.gravity = np.array(self.gravity, dtype=float)
    o.iterations = self.iterations
    o.tolerance = self.tolerance
    o.ls_iterations = self.ls_iterations
    o.ls_tolerance = self.ls_tolerance


@dataclass
class NanGuardCfg:
  """Field-for-field the reference's NanGuardCfg (utils/nan_guard.py:16-23)."""
  enabled: bool = False
  buffer_size: int = 100
  output_dir: str = "/tmp/mjlab/nan_dumps"
  max_envs_to_dump: int = 5


class NanGuard:
  """Rolling buffer of (qpos, qvel) snapshots, dumped to an .npz when a step leaves NaN / Inf in qpos, qvel, qacc or
  qacc_warmstart (reference utils/nan_guard.py:26-160: same fields, same trigger; the snapshot is the physics state as plain
  arrays instead of mj_getState's packed vector).  Disabled = every method is a no-op."""

  def __init__(self, cfg: NanGuardCfg, num_envs: int):
    self.enabled, self.num_envs = cfg.enabled, num_envs
    if not self.enabled:
      return
    from collections import deque
    self.cfg = cfg
    self.buffer = deque(maxlen=cfg.buffer_size)
    self.step_counter = 0
    self._dumped = False

  @staticmethod
  def detect_nans(data):
    import torch
    mask = torch.zeros(data.qpos.shape[0], dtype=torch.bool, device=data.qpos.device)
    for t in (data.qpos, data.qvel, data.qacc, data.qacc_warmstart):
      mask |= ~torch.isfinite(t).all(dim=-1)
    return mask

  def capture(self, data) -> None:
    if not self.enabled:
      return
    self.buffer.append({"step": self.step_counter, "qpos": data.qpos.cpu().numpy().copy(), "qvel": data.qvel.cpu().numpy().copy()})
    self.step_counter += 1

  def check_and_dump(self, data) -> bool:
    if not self.enabled or self._dumped:
      return False
    mask = self.detect_nans(data)
    if not bool(mask.any()):
      return False
    import os
    from datetime import datetime
    ids = mask.nonzero().reshape(-1).cpu().numpy().tolist()[: self.cfg.max_envs_to_dump]
    os.makedirs(self.cfg.output_dir, exist_ok=True)
    out = {}
    for item in self.buffer:
      out[f"qpos_step_{item['step']:06d}"] = item["qpos"][ids]
      out[f"qvel_step_{item['step']:06d}"] = item["qvel"][ids]
    out["_metadata"] = np.array({"num_envs_total": self.num_envs, "nan_env_ids": ids, "detection_step": self.step_counter}, dtype=object)
    self.dump_path = os.path.join(self.cfg.output_dir, f"nan_dump_{datetime.now().strftime('%Y%m%d_%H%M%S')}.npz")
    np.savez_compressed(self.dump_path, **out)
    self._dumped = True
    return True

  def watch(self, data):
    from contextlib import contextmanager

    @contextmanager
    def _cm():
      self.capture(data)
      yield
      self.check_and_dump(data)
    return _cm()


@dataclass(kw_only=True)
class SimulationCfg:
  """Field-for-field the reference's SimulationCfg (sim.py:78-94)."""
  nconmax: int | None = None
  njmax: int | None = None
  ls_parallel: bool = True
  contact_sensor_maxmatch: int = 64
  mujoco: MujocoCfg = field(default_factory=MujocoCfg)
  nan_guard: NanGuardCfg = field(default_factory=NanGuardCfg)


class Simulation:
  """Batched simulation on one GPU: `num_envs` worlds of `model`."""

  def __init__(self, num_envs: int, cfg: SimulationCfg, model, device: str, mj_data=None):
    self.cfg = cfg
    self.device = device
    self.num_envs = num_envs
    self._mj_model = model
    cfg.mujoco.apply(self._mj_model)
    self._mj_data = mj_data if mj_data is not None else HostData(model)
    self._model = engine.put_model(self._mj_model, device)
    self._model.opt.ls_parallel = cfg.ls_parallel
    self._model.opt.contact_sensor_maxmatch = cfg.contact_sensor_maxmatch
    self._data = engine.put_data(self._mj_model, self._mj_data, nworld=num_envs, nconmax=cfg.nconmax, njmax=cfg.njmax,
                                 model=self._model)
    self.use_cuda_graph = True
    self._decimation = 1
    self.create_graph()
    self.nan_guard = NanGuard(cfg.nan_guard, num_envs)

  def create_graph(self, decimation: int | None = None) -> None:
    import torch
    if decimation is not None:
      self._decimation = int(decimation)
    self.step_graph = self.forward_graph = self.decimation_graph = None
    if not self.use_cuda_graph:
      return
    dev = torch.device(self._model._device)
    side = torch.cuda.Stream(device=dev)
    side.wait_stream(torch.cuda.current_stream(dev))
    with torch.cuda.stream(side):   # warm up outside capture (lazy module loading)
      # create_graph() follows expand_model_fields + startup randomisation (manager_based_rl_env.py:107-109): refresh the
      # once-only frames of static bodies so terrain pose randomisation is honoured
      engine.refresh_static(self._model, self._data)
      engine.forward(self._model, self._data)
    torch.cuda.current_stream(dev).wait_stream(side)
    torch.cuda.synchronize(dev)
    self.step_graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(self.step_graph):
      engine.step(self._model, self._data)
    self.forward_graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(self.forward_graph):
      engine.forward(self._model, self._data)
    if self._decimation > 1:
      self.decimation_graph = torch.cuda.CUDAGraph()
      with torch.cuda.graph(self.decimation_graph):
        engine.step(self._model, self._data, self._decimation)

  @property
  def mj_model(self):
    return self._mj_model

  @property
  def mj_data(self):
    return self._mj_data

  @property
  def model(self):
    return self._model

  @property
  def data(self):
    return self._data

  wp_model = model
  wp_data = data

  def expand_model_fields(self, fields) -> None:
    invalid = [f for f in fields if not hasattr(self._mj_model, f)]
    if invalid:
      raise ValueError(f"Fields not found in model: {invalid}")
    engine.expand_model_fields(self._model, self.num_envs, list(fields))
    self.create_graph()

  def forward(self) -> None:
    if self.forward_graph is not None:
      self.forward_graph.replay()
    else:
      engine.forward(self._model, self._data)

  def step(self) -> None:
    if self.nan_guard.enabled:        # reference sim.py:194-199: the guard brackets every step
      with self.nan_guard.watch(self._data):
        self._step()
    else:
      self._step()

  def _step(self) -> None:
    if self.step_graph is not None:
      self.step_graph.replay()
    else:
      engine.step(self._model, self._data)

  def step_decimation(self) -> None:
    """All `decimation` sub-steps with ctrl held, as one graph launch."""
    if self.decimation_graph is not None:
      self.decimation_graph.replay()
    else:
      engine.step(self._model, self._data, self._decimation)
