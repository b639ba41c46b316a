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


@dataclass(kw_only=True)
class SimulationCfg:
  nconmax: int | None = None
  njmax: int | None = None
  ls_parallel: bool = True
  contact_sensor_maxmatch: int = 64
  mujoco: MujocoCfg = field(default_factory=MujocoCfg)


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
    self._data = engine.put_data(self._mj_model, self._mj_data, nworld=num_envs, nconmax=cfg.nconmax,
                                 njmax=cfg.njmax, model=self._model)
    self.use_cuda_graph = True
    self._decimation = 1
    self.create_graph()

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
