"""ctypes wrapper around the CPU oracle (TEST INFRASTRUCTURE ONLY; see oracle.h).

Usage:
    orc = Oracle(host_model, dtype="f32", nconmax=50, njmax=300)
    w = orc.new_world()            # dict-like, numpy arrays per field
    w["qpos"][:] = ...; orc.step(w)
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
if ROOT not in sys.path:
  sys.path.insert(0, ROOT)

from asimov_mjlab_b200 import abi  # noqa: E402  (data format only: the public header's ctypes mirror)

BUILD = os.path.join(HERE, "_build")


def build(force: bool = False) -> None:
  """Compile the C restatement into oracle/_build/liboracle_{f32,f64}.so (gcc, seconds)."""
  os.makedirs(BUILD, exist_ok=True)
  src = os.path.join(HERE, "mjstep_oracle.c")
  deps = [src, os.path.join(HERE, "oracle.h"), os.path.join(ROOT, "include", "b200sim.h")]
  for name, flag in (("liboracle_f32.so", []), ("liboracle_f64.so", ["-DORC_DOUBLE"])):
    out = os.path.join(BUILD, name)
    if not force and os.path.exists(out) and all(os.path.getmtime(out) >= os.path.getmtime(d) for d in deps):
      continue
    subprocess.check_call(["gcc", "-O2", "-fPIC", "-shared", "-fopenmp", *flag, "-I", os.path.join(ROOT, "include"),
                           "-o", out, src, "-lm"])


with open(os.path.join(HERE, "oracle.h")) as _f:
  _TEXT = _f.read()


class World(dict):
  """Per-world numpy arrays + the C struct that points at them."""
  struct = None


class Oracle:
  def __init__(self, model, dtype: str = "f32", nconmax: int = 64, njmax: int = 300, overrides=None,
               ls_parallel=None, opt: dict | None = None):
    build()
    self.real = np.float32 if dtype == "f32" else np.float64
    self.lib = C.CDLL(os.path.join(BUILD, f"liboracle_{dtype}.so"))
    assert self.lib.orc_real_size() == np.dtype(self.real).itemsize
    self.members = abi.parse_struct(_TEXT, "orc_data", typemap={"real": "float" if dtype == "f32" else "double"})
    self.Struct = abi.make_structure("orc_data_" + dtype, self.members)
    self.model = model
    self.desc, self._keep = abi.make_model_desc(model, overrides)
    if ls_parallel is not None:
      self.desc.opt_ls_parallel = int(ls_parallel)
    for k, v in (opt or {}).items():
      setattr(self.desc, "opt_" + k, v)
    self.dims = abi.model_dims(model)
    self.dims.update(nconmax=nconmax, njmax=njmax)
    self.nconmax, self.njmax = nconmax, njmax
    for fn in ("orc_forward", "orc_step", "orc_kinematics", "orc_com_pos", "orc_crb", "orc_factor_m", "orc_collision",
               "orc_make_constraint", "orc_fwd_velocity", "orc_fwd_actuation", "orc_fwd_acceleration", "orc_solve",
               "orc_sensors", "orc_integrate"):
      getattr(self.lib, fn).argtypes = [C.c_void_p, C.c_void_p]
    self.lib.orc_rollout.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]

  def new_world(self, qpos=None, qvel=None, ctrl=None) -> World:
    w = World()
    st = self.Struct()
    st.nconmax, st.njmax = self.nconmax, self.njmax
    for mem, ct, ptr, alen, ext in self.members:
      if not ptr:
        continue
      n = abi.eval_extent(ext, self.dims)
      a = np.zeros(max(n, 1), dtype=np.int32 if ct == "int32_t" else self.real)
      w[mem] = a[:n] if n else a[:0]
      w["_" + mem] = a
      setattr(st, mem, a.ctypes.data)
    w.struct = st
    m = self.model
    w["qpos"][:] = m.qpos0 if qpos is None else qpos
    if qvel is not None:
      w["qvel"][:] = qvel
    if ctrl is not None:
      w["ctrl"][:] = ctrl
    if m.nmocap:
      ids = [i for i in range(m.nbody) if m.body_mocapid[i] >= 0]
      w["mocap_pos"][:] = np.asarray(m.body_pos)[ids].reshape(-1)
      w["mocap_quat"][:] = np.asarray(m.body_quat)[ids].reshape(-1)
    return w

  def _call(self, name, w):
    return getattr(self.lib, name)(C.byref(self.desc), C.byref(w.struct))

  def forward(self, w):
    rc = self._call("orc_forward", w)
    assert rc == 0
    return w

  def step(self, w, n: int = 1):
    for _ in range(n):
      rc = self._call("orc_step", w)
      assert rc == 0
    return w

  def stage(self, name, w):
    self._call("orc_" + name, w)
    return w

  def rollout(self, worlds, nstep: int, nthread: int = 1) -> None:
    arr = (self.Struct * len(worlds))(*[w.struct for w in worlds])
    rc = self.lib.orc_rollout(C.byref(self.desc), arr, len(worlds), nstep, nthread)
    assert rc == 0
