"""Host-side pieces of the drop-in boundary that need no GPU: the mujoco-enum translation for real MjModels, the config
schema of `Simulation` (reference src/mjlab/sim/sim.py:41-94) and the NaN guard (utils/nan_guard.py)."""
import enum
import sys
import types

import numpy as np
import pytest


def test_mujoco_sensor_and_object_enums_are_translated_by_name(monkeypatch):
  """A real mujoco.MjModel carries mjtSensor / mjtObj values of the INSTALLED mujoco (they moved between releases); the ABI
  has its own numbering, matched by name through the installed module's enum.  A fake `mujoco` with shifted values stands in."""
  from asimov_mjlab_b200 import abi
  fake = types.ModuleType("mujoco")
  fake.mjtSensor = enum.IntEnum("mjtSensor", {"mjSENS_TOUCH": 0, "mjSENS_ACCELEROMETER": 7, "mjSENS_GYRO": 9, "mjSENS_CONTACT": 50,
                                              "mjSENS_SUBTREEANGMOM": 33, "mjSENS_RANGEFINDER": 4})
  fake.mjtObj = enum.IntEnum("mjtObj", {"mjOBJ_UNKNOWN": 0, "mjOBJ_BODY": 1, "mjOBJ_XBODY": 2, "mjOBJ_GEOM": 5, "mjOBJ_SITE": 6, "mjOBJ_CAMERA": 7})
  monkeypatch.setitem(sys.modules, "mujoco", fake)
  tab = abi.mujoco_enum_map("mjtSensor", "mjSENS_", "B2S_SENS_")
  assert tab[7] == abi.ENUMS["B2S_SENS_ACCELEROMETER"] and tab[9] == abi.ENUMS["B2S_SENS_GYRO"] and tab[50] == abi.ENUMS["B2S_SENS_CONTACT"]
  assert 0 not in tab          # touch sensors have no counterpart
  MjModel = type("MjModel", (), {"__module__": "mujoco._structs"})
  m = MjModel()
  m.sensor_type = np.array([9, 7, 50]); m.sensor_objtype = np.array([6, 6, 2]); m.sensor_reftype = np.array([0, 0, 1])
  assert abi._model_array(m, "sensor_type").tolist() == [abi.ENUMS["B2S_SENS_GYRO"], abi.ENUMS["B2S_SENS_ACCELEROMETER"], abi.ENUMS["B2S_SENS_CONTACT"]]
  assert abi._model_array(m, "sensor_objtype").tolist() == [6, 6, 2]
  m.sensor_type = np.array([0])
  with pytest.raises(ValueError, match="mjtSensor"):
    abi._model_array(m, "sensor_type")
  m.sensor_objtype = np.array([7])
  with pytest.raises(ValueError, match="mjtObj"):
    abi._model_array(m, "sensor_objtype")


def test_simulation_cfg_is_constructed_like_the_reference_tasks():
  """tasks/velocity/velocity_env_cfg.py:53-61 and tasks/tracking/tracking_env_cfg.py:56-64 build SimulationCfg like this."""
  from asimov_mjlab_b200.sim import MujocoCfg, NanGuardCfg, SimulationCfg
  cfg = SimulationCfg(nconmax=50, njmax=300, mujoco=MujocoCfg(timestep=0.005, iterations=10, ls_iterations=20))
  assert cfg.ls_parallel is True and cfg.contact_sensor_maxmatch == 64
  assert cfg.nan_guard == NanGuardCfg() and cfg.nan_guard.enabled is False and cfg.nan_guard.buffer_size == 100
  assert cfg.mujoco.integrator == "implicitfast" and cfg.mujoco.solver == "newton" and cfg.mujoco.cone == "pyramidal"
  assert cfg.mujoco.jacobian == "auto" and cfg.mujoco.tolerance == 1e-8 and cfg.mujoco.gravity == (0, 0, -9.81)
  with pytest.raises(TypeError):
    SimulationCfg(50, 300)          # kw_only, as in the reference


def test_nan_guard_buffers_and_dumps(tmp_path):
  import torch
  from asimov_mjlab_b200.sim import NanGuard, NanGuardCfg
  data = types.SimpleNamespace(qpos=torch.zeros(4, 3), qvel=torch.zeros(4, 2), qacc=torch.zeros(4, 2), qacc_warmstart=torch.zeros(4, 2))
  off = NanGuard(NanGuardCfg(), 4)
  with off.watch(data):
    data.qvel[1, 0] = float("nan")
  assert not off.enabled
  g = NanGuard(NanGuardCfg(enabled=True, buffer_size=3, output_dir=str(tmp_path), max_envs_to_dump=2), 4)
  data.qvel.zero_()
  for k in range(5):
    with g.watch(data):
      data.qpos += 1.0
  assert len(g.buffer) == 3 and not g._dumped
  with g.watch(data):
    data.qacc[2, 1] = float("inf"); data.qpos[3, 0] = float("nan")
  assert g._dumped and NanGuard.detect_nans(data).tolist() == [False, False, True, True]
  z = np.load(g.dump_path, allow_pickle=True)
  meta = z["_metadata"].item()
  assert meta["nan_env_ids"] == [2, 3] and meta["num_envs_total"] == 4
  assert z["qpos_step_000005"].shape == (2, 3) and float(z["qpos_step_000005"][0, 0]) == 5.0
