"""Python host side of the step backend: the five names mjlab uses from its
physics dependency (reference src/mjlab/sim/sim.py:116,120,143,146) on top of
the C ABI in include/b200sim.h.

    put_model(mjm)                         -> Model
    put_data(mjm, mjd, nworld, nconmax, njmax) -> Data
    step(m, d) / forward(m, d)

`Model` / `Data` expose their arrays as zero-copy `torch.Tensor`s that alias
the library's device memory (what `wp.to_torch` does in the reference,
src/mjlab/sim/sim_data.py:21); per-world-capable Model fields have leading
dim 1 expanded with stride 0 until `expand_model_fields` gives them storage
(sim_data.py:23-31, sim/randomization.py:20-54).  Field attributes cannot be
rebound (sim_data.py:217-223): write into them with indexing.

There is NO CPU fallback: if the CUDA library is missing or no GPU is
present, construction fails loudly.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import abi

_LIB = None
# B2S_LIB: development override (kernel experiments build variant libraries next to the shipped one)
_LIB_PATH = os.environ.get("B2S_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "libb200sim.so")


class BackendError(RuntimeError):
  pass


def lib():
  """Load libb200sim.so (built by build.sh / __graft_entry__.build)."""
  global _LIB
  if _LIB is None:
    if not os.path.exists(_LIB_PATH):
      raise BackendError(f"{_LIB_PATH} is missing: run ./build.sh (there is no CPU fallback)")
    L = C.CDLL(_LIB_PATH)
    L.b2s_last_error.restype = C.c_char_p
    L.b2s_launch_count.restype = C.c_int64
    L.b2s_data_bytes.restype = C.c_int64
    L.b2s_data_bytes.argtypes = [C.c_void_p]
    L.b2s_model_field_name.restype = C.c_char_p
    L.b2s_model_field_name.argtypes = [C.c_void_p, C.c_int]
    L.b2s_data_field_name.restype = C.c_char_p
    L.b2s_data_field_name.argtypes = [C.c_void_p, C.c_int]
    L.b2s_model_create.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p)]
    L.b2s_model_destroy.argtypes = [C.c_void_p]
    L.b2s_model_set_opt.argtypes = [C.c_void_p, C.c_char_p, C.c_double]
    L.b2s_model_get_opt.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(C.c_double)]
    L.b2s_model_expand.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
    L.b2s_model_field.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p]
    L.b2s_model_num_fields.argtypes = [C.c_void_p]
    L.b2s_data_create.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p)]
    L.b2s_data_destroy.argtypes = [C.c_void_p]
    L.b2s_data_field.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p]
    L.b2s_data_num_fields.argtypes = [C.c_void_p]
    L.b2s_forward.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    L.b2s_step.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    L.b2s_step_n.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    L.b2s_step_host.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.b2s_refresh_static.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    L.b2s_efc_jacobian.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.b2s_profile_step.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
    _LIB = L
  return _LIB


def _check(rc):
  if rc != 0:
    raise BackendError(lib().b2s_last_error().decode())


def launch_count() -> int:
  return int(lib().b2s_launch_count())


class _DevMem:
  """__cuda_array_interface__ holder; keeps the owning handle alive while torch views exist."""

  def __init__(self, owner, ptr, shape, typestr):
    self._owner = owner
    self.__cuda_array_interface__ = dict(shape=tuple(shape), typestr=typestr, data=(int(ptr), False), version=3, strides=None)


def _as_tensor(owner, info, device, nworld=None):
  import torch
  ndim = info.ndim
  shape = [int(info.shape[k]) for k in range(ndim)]
  bcast = ndim > 0 and info.stride[0] == 0 and shape[0] == 1
  if shape and shape[0] == -1:
    shape[0] = nworld
  typestr = "<f4" if info.dtype == 0 else "<i4"
  if any(s == 0 for s in shape):
    t = torch.empty(shape, dtype=torch.float32 if info.dtype == 0 else torch.int32, device=device)
  else:
    t = torch.as_tensor(_DevMem(owner, info.ptr, shape, typestr), device=device)
  return t, bcast


class _Handle:
  def __init__(self, ptr, destroy):
    self.ptr, self._destroy = ptr, destroy

  def __del__(self):
    try:
      if self.ptr:
        self._destroy(self.ptr)
        self.ptr = None
    except Exception:
      pass


_OPT_KEYS = ("ls_parallel", "contact_sensor_maxmatch", "iterations", "ls_iterations", "tolerance", "ls_tolerance",
             "impratio", "integrator", "cone", "solver", "disableflags", "ls_parallel_min_step", "stage_overlap", "solve_rowcap")


class _Option:
  """`Model.opt`: plain attribute reads/writes forwarded to the library (sim.py:117-118)."""

  def __init__(self, model, solver):
    object.__setattr__(self, "_m", model)      # `solver` is read back from the library like every other option

  def _get(self, key):
    v = C.c_double()
    _check(lib().b2s_model_get_opt(self._m._h.ptr, key.encode(), C.byref(v)))
    return v.value

  def __getattr__(self, k):
    import torch
    if k == "timestep":
      return torch.tensor([self._get("timestep")], dtype=torch.float32)
    if k == "gravity":
      return torch.tensor([[self._get("gravity_x"), self._get("gravity_y"), self._get("gravity_z")]], dtype=torch.float32)
    if k in _OPT_KEYS:
      v = self._get(k)
      if k == "ls_parallel":
        return bool(v)
      return v if k in ("tolerance", "ls_tolerance", "impratio", "ls_parallel_min_step") else int(v)
    raise AttributeError(k)

  def __setattr__(self, k, v):
    if k == "timestep":
      _check(lib().b2s_model_set_opt(self._m._h.ptr, b"timestep", float(np.asarray(v).reshape(-1)[0])))
    elif k == "gravity":
      g = np.asarray(v, float).reshape(-1)
      for a, x in zip("xyz", g):
        _check(lib().b2s_model_set_opt(self._m._h.ptr, f"gravity_{a}".encode(), float(x)))
    elif k in _OPT_KEYS:
      _check(lib().b2s_model_set_opt(self._m._h.ptr, k.encode(), float(v)))
    else:
      raise AttributeError(f"unknown option {k!r}")


class _Struct:
  """Common read-only-attribute behaviour of Model and Data."""
  _owned = ()

  def _field_info(self, name):
    raise NotImplementedError

  def __getattr__(self, name):
    if name.startswith("_"):
      raise AttributeError(name)
    cache = self.__dict__["_cache"]
    if name in cache:
      return cache[name]
    if name not in self.__dict__["_names"]:
      raise AttributeError(f"{type(self).__name__} has no field {name!r}")
    t = self._make(name)
    cache[name] = t
    return t

  def __setattr__(self, name, value):
    if name in self.__dict__.get("_names", ()):
      raise AttributeError(
        f"Cannot set attribute '{name}': rebinding would invalidate captured CUDA graphs. "
        f"Write in place instead: obj.{name}[:] = value")
    object.__setattr__(self, name, value)

  def clear_cache(self):
    self.__dict__["_cache"].clear()

  @property
  def field_names(self):
    return tuple(self.__dict__["_names"])

  @property
  def __dataclass_fields__(self):
    """Field-name mapping, as on the reference backend's dataclasses (sim/randomization.py:51-54 iterates it)."""
    return {n: None for n in sorted(self.__dict__["_names"])}


class Model(_Struct):
  def __init__(self, handle, device, dims, solver, host_model):
    d = self.__dict__
    d["_h"], d["_device"], d["_cache"], d["_dims"] = handle, device, {}, dims
    n = lib().b2s_model_num_fields(handle.ptr)
    d["_names"] = {lib().b2s_model_field_name(handle.ptr, i).decode() for i in range(n)}
    d["_nworld"] = None
    d["opt"] = _Option(self, solver)
    d["mj_model"] = host_model
    for k, v in dims.items():
      d[k] = v

  def _make(self, name):
    info = abi.FieldInfo()
    _check(lib().b2s_model_field(self._h.ptr, name.encode(), C.byref(info)))
    t, bcast = _as_tensor(self._h, info, self._device, self._nworld)
    if bcast and self._nworld:
      t = t.expand(self._nworld, *t.shape[1:])
    return t

  def expand(self, name, nworld):
    _check(lib().b2s_model_expand(self._h.ptr, name.encode(), int(nworld)))
    self.__dict__["_nworld"] = int(nworld)
    self.clear_cache()


class _Sub:
  def __init__(self, data, prefix, names):
    self._d, self._p, self._n = data, prefix, names

  def __getattr__(self, k):
    if k.startswith("_"):
      raise AttributeError(k)
    return getattr(self._d, self._p + k)   # efc.J -> Data.efc_J (materialised on access)


class SplitField:
  """Read view of a (nworld, nsite, ...) Data field whose static columns are stored ONCE (mjlab adds a
  visualisation site per env to the world body, terrain_importer.py:98-120, which makes `site_xpos` O(nworld^2) in the
  reference's dense layout -- 206 GB at 65 536 envs).  Indexing follows torch's rules for the patterns mjlab uses
  (`data.site_xpos[:, site_ids]`, entity/data.py:315-327; `[env_ids, ids]`, `[:, k, 2]`, ...): the selected columns are gathered
  from the per-world block `world` (nworld, nsite_world, ...) and the shared block `static` (1, nstatic, ...) and the rest of the
  key is applied to that.  `dense()` materialises the full array (small models / tests only)."""

  def __init__(self, world, static, slot):
    self.world, self.static, self.slot = world, static, slot
    self.shape = (world.shape[0], int(slot.numel()), *world.shape[2:])
    self.device, self.dtype = world.device, world.dtype

  def _columns(self, ids):
    import torch
    sl = self.slot[ids]
    mov = sl >= 0
    W = self.shape[0]
    if bool(mov.all()):
      return self.world[:, sl]
    if not bool(mov.any()):
      return self.static[:, ~sl].expand(W, *([-1] * (self.static.dim() - 1)))
    out = torch.empty((W, int(sl.numel()), *self.shape[2:]), dtype=self.dtype, device=self.device)
    out[:, mov] = self.world[:, sl[mov]]
    out[:, ~mov] = self.static[:, ~sl[~mov]]
    return out

  def __getitem__(self, key):
    import torch
    if not isinstance(key, tuple):
      key = (key,)
    if len(key) < 2:
      key = (*key, slice(None))
    k0, k1, rest = key[0], key[1], key[2:]
    n = self.shape[1]
    if isinstance(k1, slice):
      ids = torch.arange(n, device=self.device)[k1]
      return self._columns(ids)[(k0, slice(None), *rest)]
    if isinstance(k1, int):
      return self._columns(torch.tensor([k1 % n], device=self.device))[(k0, 0, *rest)]
    ids = torch.as_tensor(k1, device=self.device).long()
    pos = torch.arange(ids.numel(), device=self.device).view(ids.shape)
    return self._columns(ids.reshape(-1) % n)[(k0, pos, *rest)]

  def dense(self):
    return self[:, :]


class Data(_Struct):
  def __init__(self, handle, model, nworld, nconmax, njmax):
    d = self.__dict__
    d["_h"], d["_device"], d["_cache"], d["_model"] = handle, model._device, {}, model
    n = lib().b2s_data_num_fields(handle.ptr)
    d["_names"] = {lib().b2s_data_field_name(handle.ptr, i).decode() for i in range(n)}
    d["nworld"], d["nconmax"], d["njmax"], d["naconmax"] = nworld, nconmax, njmax, nworld * nconmax
    d["contact"] = _Sub(self, "contact_", None)
    d["efc"] = _Sub(self, "efc_", None)
    model.__dict__["_nworld"] = nworld
    if "site_xpos_world" in d["_names"]:          # split site storage (H3): the reference's names resolve to SplitField views
      d["_names"] |= {"site_xpos", "site_xmat"}

  def _make(self, name):
    if name in ("site_xpos", "site_xmat") and "site_xpos_world" in self.__dict__["_names"]:
      return SplitField(getattr(self, name + "_world"), getattr(self, name + "_static"), self._model.site_slot.long())
    info = abi.FieldInfo()
    _check(lib().b2s_data_field(self._h.ptr, name.encode(), C.byref(info)))
    t, _ = _as_tensor(self._h, info, self._device)
    return t

  @property
  def efc_J(self):
    """Dense constraint Jacobian (mjwarp's Data.efc.J), shape (nworld, njmax, roundup4(nv)).  The step keeps
    rows matrix-free (efc_g, efc_bp); every access re-materialises the dense matrix from the current rows."""
    info = abi.FieldInfo()
    _check(lib().b2s_efc_jacobian(self._model._h.ptr, self._h.ptr, _stream(self._model), C.byref(info)))
    cache = self.__dict__["_cache"]
    if "efc_J" not in cache:
      cache["efc_J"], _ = _as_tensor(self._h, info, self._device)
    return cache["efc_J"]

  @property
  def bytes(self):
    return int(lib().b2s_data_bytes(self._h.ptr))


def _device_index(device) -> int:
  import torch
  dev = torch.device(device)
  if dev.type != "cuda":
    raise BackendError(f"this backend runs on CUDA devices only (got {device!r}); there is no CPU fallback")
  if not torch.cuda.is_available():
    raise BackendError("no CUDA device is available; there is no CPU fallback")
  return dev.index if dev.index is not None else torch.cuda.current_device()


_LAST_MODEL: dict = {}      # id(host model) -> weakref to the Model put_model returned last (put_data's default)


def put_model(mjm, device="cuda:0") -> Model:
  """Upload a compiled model (HostModel or mujoco.MjModel-like).  Raises on unsupported features."""
  idx = _device_index(device)
  desc, keep = abi.make_model_desc(mjm)
  h = C.c_void_p()
  _check(lib().b2s_model_create(C.byref(desc), idx, C.byref(h)))
  handle = _Handle(h.value, lib().b2s_model_destroy)
  m = Model(handle, f"cuda:{idx}", abi.model_dims(mjm), int(mjm.opt.solver), mjm)
  rg = getattr(mjm, "geom_rgba", None)
  if rg is not None and len(rg):
    import torch
    m.geom_rgba[:] = torch.as_tensor(np.asarray(rg, np.float32).reshape(1, -1, 4), device=m._device)
  import weakref
  _LAST_MODEL[id(mjm)] = weakref.ref(m)
  return m


def put_data(mjm, mjd, nworld: int, nconmax: int | None = None, njmax: int | None = None, *, model: Model | None = None) -> Data:
  """Allocate per-world arrays and initialise every world from `mjd`.

  The reference signature is put_data(mjm, mjd, nworld, nconmax, njmax) (sim.py:120-126) and works as is: the device Model
  is the one the last put_model(mjm) returned (it carries the device and the shared-memory configuration); `model=` names
  another one explicitly (several devices / several uploads of one host model).
  """
  if model is None:
    ref = _LAST_MODEL.get(id(mjm))
    model = ref() if ref is not None else None
    if model is None:
      raise BackendError("put_data: call put_model(mjm) first (or pass model=<Model returned by put_model>)")
  st, keep = abi.make_state_desc(mjm, mjd)
  h = C.c_void_p()
  _check(lib().b2s_data_create(model._h.ptr, C.byref(st), int(nworld), int(nconmax or 0), int(njmax or 0), C.byref(h)))
  handle = _Handle(h.value, lib().b2s_data_destroy)
  # read back the sizes the library settled on
  info = abi.FieldInfo()
  _check(lib().b2s_data_field(handle.ptr, b"contact_dist", C.byref(info)))
  ncm = int(info.shape[1])
  _check(lib().b2s_data_field(handle.ptr, b"efc_D", C.byref(info)))
  njm = int(info.shape[1])
  return Data(handle, model, int(nworld), ncm, njm)


def _stream(m: Model):
  import torch
  return C.c_void_p(torch.cuda.current_stream(m._device).cuda_stream)


def forward(m: Model, d: Data) -> None:
  _check(lib().b2s_forward(m._h.ptr, d._h.ptr, _stream(m)))


def refresh_static(m: Model, d: Data) -> None:
  """Recompute geom/site frames of static bodies (they are not touched by step/forward; see include/b200sim.h)."""
  _check(lib().b2s_refresh_static(m._h.ptr, d._h.ptr, _stream(m)))


def step(m: Model, d: Data, nsub: int = 1) -> None:
  _check(lib().b2s_step_n(m._h.ptr, d._h.ptr, int(nsub), _stream(m)))


def step_host(m: Model, d: Data, ctrl_host, nsub, qpos_host, qvel_host, sensordata_host) -> None:
  """End-to-end call with HOST (pinned) buffers: H2D ctrl, nsub sub-steps, D2H qpos/qvel/sensordata, sync."""
  def p(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None
  _check(lib().b2s_step_host(m._h.ptr, d._h.ptr, p(ctrl_host), int(nsub), p(qpos_host), p(qvel_host), p(sensordata_host), _stream(m)))


def profile_step(m: Model, d: Data, reps: int = 10) -> dict:
  names = (C.c_char_p * 8)()
  ms = (C.c_float * 8)()
  n = lib().b2s_profile_step(m._h.ptr, d._h.ptr, _stream(m), int(reps), names, ms, 8)
  if n < 0:
    _check(n)
  return {names[i].decode(): float(ms[i]) for i in range(n)}


def expand_model_fields(m: Model, nworld: int, fields) -> None:
  """Give each named Model field per-world storage (reference sim/randomization.py:20-54)."""
  for f in fields:
    m.expand(f, nworld)


def solve_noalloc(m: Model, d: Data) -> None:
  """Allocation-free re-solve of the constraint problem left in `d` (b2s_solve_noalloc, include/b200sim.h: no counterpart in
  the reference; scheduling scratch in a fixed device array, nworld <= noalloc_max_worlds())."""
  L = lib()
  L.b2s_solve_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
  _check(L.b2s_solve_noalloc(m._h.ptr, d._h.ptr, _stream(m)))


def noalloc_max_worlds() -> int:
  return int(lib().b2s_noalloc_max_worlds())
