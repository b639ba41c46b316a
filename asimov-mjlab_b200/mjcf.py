"""MJCF-subset spec + compiler: the model on-ramp of the step backend.

The reference obtains its `MjModel` from MuJoCo C (`mujoco.MjSpec.compile()`,
reference src/mjlab/scene/scene.py:42-43) which is not installed here
(SURVEY.md §7.3 H1).  This module is a small stand-in for the part of that
compiler the hot path depends on: it parses the MJCF subset used by the
reference's robots and test fixtures (defaults classes, `fromto` capsules,
explicit/derived inertial frames, hinge/slide/ball/free joints, excludes,
sensors), supports the spec edits mjlab applies before compiling
(src/mjlab/utils/spec_config.py:206-237 collisions, :361-414 actuators,
src/mjlab/sensor/contact_sensor.py:472-533 contact sensors,
src/mjlab/scene/scene.py:38-39 prefix attach) and derives the compile-time
quantities the step needs (`body_subtreemass`, `dof_invweight0`,
`body_invweight0`, `geom_rbound`, tree tables, `stat.meaninertia`).

It produces a `HostModel` whose attribute names follow MuJoCo's `MjModel`
so `put_model` accepts either object.  Mesh geoms are carried as
non-colliding placeholders (all collision geoms of the benchmark robots are
primitives).
"""
from __future__ import annotations

import copy
import math
import os
import re
import xml.etree.ElementTree as ET
from dataclasses import dataclass, field

import numpy as np

from .hostmodel import HostModel, finalize_model

JNT = {"free": 0, "ball": 1, "slide": 2, "hinge": 3}
GEOM = {"plane": 0, "hfield": 1, "sphere": 2, "capsule": 3, "ellipsoid": 4,
        "cylinder": 5, "box": 6, "mesh": 7}
OBJ = {"body": 1, "xbody": 2, "joint": 3, "dof": 4, "geom": 5, "site": 6, "tendon": 18, "actuator": 19}
SENS = {"accelerometer": 1, "velocimeter": 2, "gyro": 3, "jointpos": 9, "jointvel": 10, "tendonpos": 11, "tendonvel": 12,
        "actuatorpos": 13, "actuatorvel": 14, "actuatorfrc": 15, "jointactuatorfrc": 16,
        "framepos": 25, "framequat": 26, "framexaxis": 27, "frameyaxis": 28, "framezaxis": 29,
        "framelinvel": 30, "frameangvel": 31, "framelinacc": 32, "frameangacc": 33, "subtreecom": 34, "subtreelinvel": 35,
        "subtreeangmom": 36, "contact": 42, "clock": 45, "magnetometer": 6, "tendonactuatorfrc": 17,
        "jointlimitpos": 20, "jointlimitvel": 21, "jointlimitfrc": 22, "e_potential": 43, "e_kinetic": 44, "force": 4, "torque": 5, "rangefinder": 7}
SENS_DIM = {"accelerometer": 3, "velocimeter": 3, "gyro": 3, "jointpos": 1, "jointvel": 1, "tendonpos": 1, "tendonvel": 1,
            "actuatorpos": 1, "actuatorvel": 1, "actuatorfrc": 1, "jointactuatorfrc": 1,
            "framepos": 3, "framequat": 4, "framexaxis": 3, "frameyaxis": 3, "framezaxis": 3,
            "framelinvel": 3, "frameangvel": 3, "framelinacc": 3, "frameangacc": 3, "subtreecom": 3, "subtreelinvel": 3,
            "subtreeangmom": 3, "clock": 1, "magnetometer": 3, "tendonactuatorfrc": 1, "jointlimitpos": 1, "jointlimitvel": 1,
            "jointlimitfrc": 1, "e_potential": 1, "e_kinetic": 1, "force": 3, "torque": 3, "rangefinder": 1}
CONTACT_FIELD_DIM = {0: 1, 1: 3, 2: 3, 3: 1, 4: 3, 5: 3, 6: 3}


def _floats(s, n=None):
  v = [float(x) for x in s.split()]
  if n is not None and len(v) != n:
    raise ValueError(f"expected {n} numbers, got {s!r}")
  return v


def quat_mul(a, b):
  aw, ax, ay, az = a
  bw, bx, by, bz = b
  return np.array([aw*bw - ax*bx - ay*by - az*bz, aw*bx + ax*bw + ay*bz - az*by,
                   aw*by - ax*bz + ay*bw + az*bx, aw*bz + ax*by - ay*bx + az*bw])


def quat_z2vec(vec):
  """Quaternion rotating +z onto vec (MuJoCo's mjuu_z2quat convention)."""
  v = np.asarray(vec, float)
  n = np.linalg.norm(v)
  if n < 1e-15:
    return np.array([1.0, 0, 0, 0])
  v = v / n
  axis = np.cross([0.0, 0, 1.0], v)
  s = np.linalg.norm(axis)
  if s < 1e-10:
    axis = np.array([1.0, 0, 0])
  else:
    axis = axis / s
  ang = math.atan2(s, v[2])
  return np.array([math.cos(ang/2), *(axis*math.sin(ang/2))])


def euler2quat(e, seq="xyz"):
  q = np.array([1.0, 0, 0, 0])
  for ang, ax in zip(e, seq):
    a = {"x": [1, 0, 0], "y": [0, 1, 0], "z": [0, 0, 1]}[ax.lower()]
    r = np.array([math.cos(ang/2), *(np.array(a)*math.sin(ang/2))])
    q = quat_mul(q, r) if ax.islower() else quat_mul(r, q)
  return q


@dataclass
class GeomSpec:
  name: str = ""
  type: str = "sphere"
  size: list = field(default_factory=lambda: [0.0, 0.0, 0.0])
  pos: list = field(default_factory=lambda: [0.0, 0.0, 0.0])
  quat: list = field(default_factory=lambda: [1.0, 0.0, 0.0, 0.0])
  fromto: list | None = None
  contype: int = 1
  conaffinity: int = 1
  condim: int = 3
  priority: int = 0
  friction: list = field(default_factory=lambda: [1.0, 0.005, 0.0001])
  solmix: float = 1.0
  solref: list = field(default_factory=lambda: [0.02, 1.0])
  solimp: list = field(default_factory=lambda: [0.9, 0.95, 0.001, 0.5, 2.0])
  margin: float = 0.0
  gap: float = 0.0
  mass: float | None = None
  density: float = 1000.0
  rgba: list = field(default_factory=lambda: [0.5, 0.5, 0.5, 1.0])
  mesh: str = ""
  hfield: str = ""


@dataclass
class MeshSpec:
  """<asset><mesh>: vertices come from `vertex` (inline) or from a binary STL `file` (read lazily, at compile time,
  and only for meshes that a COLLIDING geom uses)."""
  name: str = ""
  file: str = ""
  vertex: object = None            # (n, 3) array or None
  scale: list = field(default_factory=lambda: [1.0, 1.0, 1.0])


@dataclass
class HFieldSpec:
  """<asset><hfield>: size = (x half extent, y half extent, z top, z base thickness); data[nrow, ncol] in [0, 1]
  (row <-> y, column <-> x; MuJoCo normalises elevation data the same way)."""
  name: str = ""
  nrow: int = 0
  ncol: int = 0
  size: list = field(default_factory=lambda: [1.0, 1.0, 1.0, 0.1])
  data: object = None


@dataclass
class JointSpec:
  name: str = ""
  type: str = "hinge"
  pos: list = field(default_factory=lambda: [0.0, 0.0, 0.0])
  axis: list = field(default_factory=lambda: [0.0, 0.0, 1.0])
  range: list = field(default_factory=lambda: [0.0, 0.0])
  limited: str = "auto"
  ref: float = 0.0
  springref: float = 0.0
  stiffness: float = 0.0
  damping: float = 0.0
  armature: float = 0.0
  frictionloss: float = 0.0
  margin: float = 0.0
  solreflimit: list = field(default_factory=lambda: [0.02, 1.0])
  solimplimit: list = field(default_factory=lambda: [0.9, 0.95, 0.001, 0.5, 2.0])


@dataclass
class SiteSpec:
  name: str = ""
  pos: list = field(default_factory=lambda: [0.0, 0.0, 0.0])
  quat: list = field(default_factory=lambda: [1.0, 0.0, 0.0, 0.0])


@dataclass
class InertialSpec:
  pos: list
  quat: list
  mass: float
  diaginertia: list | None = None
  fullinertia: list | None = None


@dataclass
class BodySpec:
  name: str = ""
  pos: list = field(default_factory=lambda: [0.0, 0.0, 0.0])
  quat: list = field(default_factory=lambda: [1.0, 0.0, 0.0, 0.0])
  mocap: bool = False
  inertial: InertialSpec | None = None
  joints: list = field(default_factory=list)
  geoms: list = field(default_factory=list)
  sites: list = field(default_factory=list)
  bodies: list = field(default_factory=list)

  def add_body(self, **kw):
    b = BodySpec(**kw)
    self.bodies.append(b)
    return b

  def add_geom(self, **kw):
    g = GeomSpec(**kw)
    self.geoms.append(g)
    return g

  def add_site(self, **kw):
    s = SiteSpec(**kw)
    self.sites.append(s)
    return s

  def add_joint(self, **kw):
    j = JointSpec(**kw)
    self.joints.append(j)
    return j

  def add_freejoint(self, name=""):
    return self.add_joint(name=name, type="free")


@dataclass
class TendonSpec:
  """<tendon><fixed>: length = sum_k coef_k * qpos[joint_k]."""
  name: str
  joints: list = field(default_factory=list)      # [(joint name, coef)]
  stiffness: float = 0.0
  damping: float = 0.0
  springlength: list = field(default_factory=lambda: [-1.0, -1.0])
  range: list | None = None
  limited: bool = False


@dataclass
class ActuatorSpec:
  name: str
  joint: str = ""
  tendon: str = ""
  gaintype: int = 0
  biastype: int = 0
  gainprm: list = field(default_factory=lambda: [1.0, 0.0, 0.0])
  biasprm: list = field(default_factory=lambda: [0.0, 0.0, 0.0])
  gear: float = 1.0
  ctrlrange: list | None = None
  forcerange: list | None = None
  inheritrange: float = 0.0


@dataclass
class SensorSpec:
  name: str
  type: str
  objtype: str = ""
  objname: str = ""
  reftype: str = ""
  refname: str = ""
  intprm: list = field(default_factory=lambda: [0, 0, 0])
  cutoff: float = 0.0


class Spec:
  """Editable model specification (a tiny analogue of `mujoco.MjSpec`)."""

  def __init__(self):
    self.modelname = ""
    self.worldbody = BodySpec(name="world")
    self.actuators: list[ActuatorSpec] = []
    self.tendons: list[TendonSpec] = []
    self.meshes: dict[str, MeshSpec] = {}
    self.hfields: dict[str, HFieldSpec] = {}
    self.sensors: list[SensorSpec] = []
    self.excludes: list[tuple[str, str]] = []
    self.keys: dict[str, dict] = {}
    self.option = dict(timestep=0.002, gravity=[0.0, 0.0, -9.81], integrator="euler",
                       cone="pyramidal", solver="newton", iterations=100, tolerance=1e-8,
                       ls_iterations=50, ls_tolerance=0.01, impratio=1.0, disableflags=0)

  # -- parsing --------------------------------------------------------------
  @classmethod
  def from_string(cls, xml: str) -> "Spec":
    return _Parser().parse(ET.fromstring(xml))

  @classmethod
  def from_file(cls, path) -> "Spec":
    return _Parser(basedir=os.path.dirname(os.path.abspath(str(path)))).parse(ET.parse(str(path)).getroot())

  # -- queries --------------------------------------------------------------
  def _walk(self, body=None):
    body = body or self.worldbody
    yield body
    for c in body.bodies:
      yield from self._walk(c)

  @property
  def bodies(self):
    return list(self._walk())

  @property
  def geoms(self):
    return [g for b in self._walk() for g in b.geoms]

  @property
  def joints(self):
    return [j for b in self._walk() for j in b.joints]

  @property
  def sites(self):
    return [s for b in self._walk() for s in b.sites]

  def body(self, name):
    for b in self._walk():
      if b.name == name:
        return b
    raise KeyError(name)

  def geom(self, name):
    for g in self.geoms:
      if g.name == name:
        return g
    raise KeyError(name)

  def joint(self, name):
    for j in self.joints:
      if j.name == name:
        return j
    raise KeyError(name)

  # -- edits ----------------------------------------------------------------
  def add_actuator(self, **kw):
    a = ActuatorSpec(**kw)
    self.actuators.append(a)
    return a

  def add_sensor(self, **kw):
    s = SensorSpec(**kw)
    self.sensors.append(s)
    return s

  def add_mesh(self, name, vertex=None, file="", scale=(1.0, 1.0, 1.0)):
    self.meshes[name] = MeshSpec(name=name, file=str(file), vertex=None if vertex is None else np.asarray(vertex, float).reshape(-1, 3),
                                 scale=list(scale))
    return self.meshes[name]

  def add_hfield(self, name, data, size):
    data = np.asarray(data, float)
    lo, hi = float(data.min()), float(data.max())
    norm = (data - lo)/(hi - lo) if hi > lo else np.zeros_like(data)
    self.hfields[name] = HFieldSpec(name=name, nrow=data.shape[0], ncol=data.shape[1], size=list(map(float, size)), data=norm)
    return self.hfields[name]

  def add_exclude(self, body1, body2):
    self.excludes.append((body1, body2))

  def attach(self, child: "Spec", prefix: str = "") -> None:
    """Attach `child`'s world children under this world body with a name prefix
    (what mjlab's Scene does per entity; reference scene/scene.py:38-39)."""
    child = copy.deepcopy(child)

    def ren(n):
      return prefix + n if n else n
    for b in child._walk():
      if b is not child.worldbody:
        b.name = ren(b.name)
      for lst in (b.joints, b.geoms, b.sites):
        for e in lst:
          e.name = ren(e.name)
      for g in b.geoms:
        g.mesh, g.hfield = ren(g.mesh), ren(g.hfield)
    for k, v in child.meshes.items():
      v.name = ren(k); self.meshes[ren(k)] = v
    for k, v in child.hfields.items():
      v.name = ren(k); self.hfields[ren(k)] = v
    self.worldbody.bodies.extend(child.worldbody.bodies)
    self.worldbody.geoms.extend(child.worldbody.geoms)
    self.worldbody.sites.extend(child.worldbody.sites)
    for t in child.tendons:
      t.name = ren(t.name); t.joints = [(ren(j), c) for j, c in t.joints]
      self.tendons.append(t)
    for a in child.actuators:
      a.name, a.joint, a.tendon = ren(a.name), ren(a.joint), ren(a.tendon)
      self.actuators.append(a)
    for s in child.sensors:
      s.name, s.objname, s.refname = ren(s.name), ren(s.objname), ren(s.refname)
      self.sensors.append(s)
    for e in child.excludes:
      self.excludes.append((ren(e[0]), ren(e[1])))
    for k, v in child.keys.items():
      self.keys[ren(k)] = v

  def compile(self) -> HostModel:
    return _compile(self)


class _Parser:
  def __init__(self, basedir=""):
    self.basedir = basedir
    self.meshdir = ""
    self.defaults: dict[str, dict[str, dict[str, str]]] = {"main": {}}
    self.angle_deg = True
    self.autolimits = True
    self.eulerseq = "xyz"

  def parse(self, root) -> Spec:
    spec = Spec()
    spec.modelname = root.get("model", "")
    for comp in root.findall("compiler"):
      if comp.get("angle"):
        self.angle_deg = comp.get("angle") == "degree"
      if comp.get("autolimits"):
        self.autolimits = comp.get("autolimits") == "true"
      if comp.get("eulerseq"):
        self.eulerseq = comp.get("eulerseq")
      if comp.get("meshdir"):
        self.meshdir = comp.get("meshdir")
    for opt in root.findall("option"):
      for k in ("timestep", "tolerance", "ls_tolerance", "impratio"):
        if opt.get(k):
          spec.option[k] = float(opt.get(k))
      for k in ("iterations", "ls_iterations"):
        if opt.get(k):
          spec.option[k] = int(opt.get(k))
      for k in ("integrator", "cone", "solver"):
        if opt.get(k):
          spec.option[k] = opt.get(k).lower()
      if opt.get("gravity"):
        spec.option["gravity"] = _floats(opt.get("gravity"), 3)
      if opt.get("magnetic"):
        spec.option["magnetic"] = _floats(opt.get("magnetic"), 3)
    for d in root.findall("default"):
      self._parse_default(d, "main", top=True)
    for sec in root.findall("asset"):
      for el in sec:
        if el.tag == "mesh":
          f = el.get("file", "")
          name = el.get("name") or os.path.splitext(os.path.basename(f))[0]
          vert = np.array(_floats(el.get("vertex"))).reshape(-1, 3) if el.get("vertex") else None
          # asset search order of the reference's Entity (update_assets puts <robot>/meshes beside <robot>/xmls)
          cands = [os.path.join(self.basedir, self.meshdir, f), os.path.join(self.basedir, f),
                   os.path.join(self.basedir, "assets", f), os.path.join(os.path.dirname(self.basedir), "meshes", f)]
          path = next((c for c in cands if f and os.path.exists(c)), cands[0] if f else "")
          spec.meshes[name] = MeshSpec(name=name, file=path, vertex=vert, scale=_floats(el.get("scale", "1 1 1"), 3))
        elif el.tag == "hfield":
          if el.get("file"):
            # MuJoCo's own binary height-field format: int32 nrow, int32 ncol, then nrow * ncol float32 (PNG files need an
            # image decoder that is not available here)
            f = el.get("file")
            path = next((c for c in (os.path.join(self.basedir, self.meshdir, f), os.path.join(self.basedir, f)) if os.path.exists(c)), f)
            if path.lower().endswith(".png"):
              raise NotImplementedError("<hfield file=*.png> is not supported (binary .bin height fields or inline elevation)")
            with open(path, "rb") as fh:
              raw = fh.read()
            nrow, ncol = (int(x) for x in np.frombuffer(raw, "<i4", 2))
            if len(raw) < 8 + 4*nrow*ncol:
              raise ValueError(f"{path}: truncated height field")
            data = np.frombuffer(raw, "<f4", nrow*ncol, 8).astype(float).reshape(nrow, ncol)
          else:
            nrow, ncol = int(el.get("nrow")), int(el.get("ncol"))
            data = np.array(_floats(el.get("elevation"))).reshape(nrow, ncol) if el.get("elevation") else np.zeros((nrow, ncol))
          spec.add_hfield(el.get("name", ""), data, _floats(el.get("size"), 4))
    wb = root.find("worldbody")
    if wb is not None:
      self._parse_body_children(wb, spec.worldbody, None)
    for con in root.findall("contact"):
      for ex in con.findall("exclude"):
        spec.add_exclude(ex.get("body1"), ex.get("body2"))
      if con.findall("pair"):
        raise NotImplementedError("<contact><pair> is not supported")
    for sec in root.findall("sensor"):
      for el in sec:
        self._parse_sensor(el, spec)
    for sec in root.findall("actuator"):
      for el in sec:
        self._parse_actuator(el, spec)
    for sec in root.findall("keyframe"):
      for k in sec.findall("key"):
        spec.keys[k.get("name", "")] = {a: _floats(k.get(a)) for a in ("qpos", "qvel", "ctrl") if k.get(a)}
    for sec in root.findall("tendon"):
      for el in sec:
        if el.tag != "fixed":
          raise NotImplementedError("only <fixed> tendons are supported (no spatial tendons)")
        t = TendonSpec(name=el.get("name", ""))
        t.joints = [(j.get("joint"), float(j.get("coef"))) for j in el.findall("joint")]
        t.stiffness = float(el.get("stiffness", 0)); t.damping = float(el.get("damping", 0))
        if el.get("springlength"):
          v = _floats(el.get("springlength")); t.springlength = [v[0], v[-1]]
        if el.get("range"):
          t.range = _floats(el.get("range"), 2)
          t.limited = (el.get("limited", "auto") == "true") or (el.get("limited", "auto") == "auto" and self.autolimits)
        if float(el.get("frictionloss", 0)) != 0 or float(el.get("armature", 0)) != 0:
          raise NotImplementedError("tendon frictionloss / armature are not supported")
        spec.tendons.append(t)
    for bad in ("equality",):
      if root.find(bad) is not None and len(root.find(bad)):
        raise NotImplementedError(f"<{bad}> is not supported by this backend")
    return spec

  def _parse_default(self, el, parent, top=False):
    name = el.get("class", "main" if top else None)
    if name is None:
      raise ValueError("nested default needs a class")
    base = copy.deepcopy(self.defaults.get(parent, {})) if name != parent else self.defaults.get(name, {})
    if name == "main" and top:
      base = self.defaults["main"]
    for ch in el:
      if ch.tag == "default":
        continue
      base.setdefault(ch.tag, {}).update(ch.attrib)
    self.defaults[name] = base
    for ch in el.findall("default"):
      self._parse_default(ch, name)

  def _attrs(self, el, childclass):
    cls = el.get("class") or childclass or "main"
    if cls not in self.defaults:
      raise ValueError(f"unknown default class {cls!r}")
    out = dict(self.defaults[cls].get(el.tag, {}))
    out.update(el.attrib)
    out.pop("class", None)
    return out

  def _orient(self, a):
    if "quat" in a:
      q = np.array(_floats(a["quat"], 4))
      return list(q / np.linalg.norm(q))
    if "euler" in a:
      e = np.array(_floats(a["euler"], 3))
      if self.angle_deg:
        e = np.deg2rad(e)
      return list(euler2quat(e, self.eulerseq))
    if "axisangle" in a:
      v = _floats(a["axisangle"], 4)
      ang = math.radians(v[3]) if self.angle_deg else v[3]
      ax = np.array(v[:3]) / np.linalg.norm(v[:3])
      return [math.cos(ang/2), *(ax*math.sin(ang/2))]
    if "zaxis" in a:
      return list(quat_z2vec(_floats(a["zaxis"], 3)))
    if "xyaxes" in a:
      v = _floats(a["xyaxes"], 6)
      x = np.array(v[:3]); x /= np.linalg.norm(x)
      y = np.array(v[3:]); y -= x*np.dot(x, y); y /= np.linalg.norm(y)
      return list(mat2quat(np.stack([x, y, np.cross(x, y)], axis=1)))
    return [1.0, 0.0, 0.0, 0.0]

  def _parse_body_children(self, el, body: BodySpec, childclass):
    for ch in el:
      t = ch.tag
      if t == "body":
        cc = ch.get("childclass", childclass)
        b = body.add_body(name=ch.get("name", ""), pos=_floats(ch.get("pos", "0 0 0"), 3),
                          quat=self._orient(ch.attrib), mocap=ch.get("mocap") == "true")
        self._parse_body_children(ch, b, cc)
      elif t == "inertial":
        body.inertial = InertialSpec(
          pos=_floats(ch.get("pos", "0 0 0"), 3), quat=self._orient(ch.attrib), mass=float(ch.get("mass")),
          diaginertia=_floats(ch.get("diaginertia"), 3) if ch.get("diaginertia") else None,
          fullinertia=_floats(ch.get("fullinertia"), 6) if ch.get("fullinertia") else None)
      elif t == "freejoint":
        body.add_freejoint(ch.get("name", ""))
      elif t == "joint":
        a = self._attrs(ch, childclass)
        j = body.add_joint(name=a.get("name", ""), type=a.get("type", "hinge"))
        ang = j.type in ("hinge", "ball") and self.angle_deg
        if "pos" in a: j.pos = _floats(a["pos"], 3)
        if "axis" in a: j.axis = _floats(a["axis"], 3)
        if "range" in a:
          j.range = _floats(a["range"], 2)
          if ang: j.range = [math.radians(x) for x in j.range]
        if "limited" in a: j.limited = a["limited"]
        elif not self.autolimits: j.limited = "false"
        for k in ("ref", "springref"):
          if k in a:
            setattr(j, k, math.radians(float(a[k])) if ang else float(a[k]))
        for k in ("stiffness", "damping", "armature", "frictionloss", "margin"):
          if k in a: setattr(j, k, float(a[k]))
        if "solreffriction" in a or "solimpfriction" in a:
          raise NotImplementedError("joint solreffriction / solimpfriction are not supported (friction-loss rows use MuJoCo's defaults)")
        if "solreflimit" in a: j.solreflimit = _floats(a["solreflimit"], 2)
        if "solimplimit" in a:
          v = _floats(a["solimplimit"]); j.solimplimit[:len(v)] = v
      elif t == "geom":
        a = self._attrs(ch, childclass)
        g = body.add_geom(name=a.get("name", ""), type=a.get("type", "sphere"))
        if "size" in a:
          v = _floats(a["size"]); g.size[:len(v)] = v
        if "pos" in a: g.pos = _floats(a["pos"], 3)
        g.quat = self._orient(a)
        if "fromto" in a: g.fromto = _floats(a["fromto"], 6)
        for k in ("contype", "conaffinity", "condim", "priority"):
          if k in a: setattr(g, k, int(a[k]))
        if "friction" in a:
          v = _floats(a["friction"]); g.friction[:len(v)] = v
        for k in ("solmix", "margin", "gap", "density"):
          if k in a: setattr(g, k, float(a[k]))
        if "mass" in a: g.mass = float(a["mass"])
        if "solref" in a: g.solref = _floats(a["solref"], 2)
        if "solimp" in a:
          v = _floats(a["solimp"]); g.solimp[:len(v)] = v
        if "rgba" in a: g.rgba = _floats(a["rgba"], 4)
        g.mesh = a.get("mesh", "")
        g.hfield = a.get("hfield", "")
      elif t == "site":
        a = self._attrs(ch, childclass)
        body.add_site(name=a.get("name", ""), pos=_floats(a.get("pos", "0 0 0"), 3), quat=self._orient(a))
      elif t in ("light", "camera"):
        pass  # rendering only (unavailable on this box)
      else:
        raise NotImplementedError(f"<{t}> inside <body> is not supported")

  def _parse_sensor(self, el, spec):
    t = el.tag
    if t not in SENS or t == "contact":
      raise NotImplementedError(f"sensor <{t}> is not supported")
    s = SensorSpec(name=el.get("name", ""), type=t, cutoff=float(el.get("cutoff", 0)))
    if t in ("accelerometer", "velocimeter", "gyro", "magnetometer", "force", "torque", "rangefinder"):
      s.objtype, s.objname = "site", el.get("site")
    elif t in ("jointpos", "jointvel", "jointactuatorfrc", "jointlimitpos", "jointlimitvel", "jointlimitfrc"):
      s.objtype, s.objname = "joint", el.get("joint")
    elif t in ("actuatorpos", "actuatorvel", "actuatorfrc"):
      s.objtype, s.objname = "actuator", el.get("actuator")
    elif t in ("tendonpos", "tendonvel", "tendonactuatorfrc"):
      s.objtype, s.objname = "tendon", el.get("tendon")
    elif t.startswith("subtree"):
      s.objtype, s.objname = "body", el.get("body")
    elif t.startswith("frame"):
      s.objtype, s.objname = el.get("objtype"), el.get("objname")
      if el.get("reftype"):
        s.reftype, s.refname = el.get("reftype"), el.get("refname")
    spec.sensors.append(s)

  def _parse_actuator(self, el, spec):
    t = el.tag
    a = ActuatorSpec(name=el.get("name", ""), joint=el.get("joint", ""), tendon=el.get("tendon", ""))
    if not a.joint and not a.tendon:
      raise NotImplementedError("only joint and (fixed) tendon transmissions are supported")
    a.gear = _floats(el.get("gear", "1"))[0]
    if el.get("ctrlrange"): a.ctrlrange = _floats(el.get("ctrlrange"), 2)
    if el.get("forcerange"): a.forcerange = _floats(el.get("forcerange"), 2)
    if t == "motor":
      pass
    elif t == "position":
      kp, kv = float(el.get("kp", 1)), float(el.get("kv", 0))
      a.gainprm = [kp, 0, 0]; a.biastype = 1; a.biasprm = [0, -kp, -kv]
      a.inheritrange = float(el.get("inheritrange", 0))
    elif t == "velocity":
      kv = float(el.get("kv", 1))
      a.gainprm = [kv, 0, 0]; a.biastype = 1; a.biasprm = [0, 0, -kv]
    elif t == "general":
      if el.get("gainprm"):
        v = _floats(el.get("gainprm")); a.gainprm[:min(3, len(v))] = v[:3]
      if el.get("biasprm"):
        v = _floats(el.get("biasprm")); a.biasprm[:min(3, len(v))] = v[:3]
      a.biastype = {"none": 0, "affine": 1}[el.get("biastype", "none")]
      if el.get("gaintype", "fixed") != "fixed" or el.get("dyntype", "none") != "none":
        raise NotImplementedError("only fixed-gain, stateless actuators are supported")
    else:
      raise NotImplementedError(f"actuator <{t}> is not supported")
    spec.actuators.append(a)


def mat2quat(R):
  """Rotation matrix -> unit quaternion (w,x,y,z), w >= 0 branch-stable."""
  m = np.asarray(R, float)
  tr = m[0, 0] + m[1, 1] + m[2, 2]
  if tr > 0:
    s = math.sqrt(tr + 1.0) * 2
    q = [0.25*s, (m[2, 1]-m[1, 2])/s, (m[0, 2]-m[2, 0])/s, (m[1, 0]-m[0, 1])/s]
  elif m[0, 0] > m[1, 1] and m[0, 0] > m[2, 2]:
    s = math.sqrt(1.0 + m[0, 0] - m[1, 1] - m[2, 2]) * 2
    q = [(m[2, 1]-m[1, 2])/s, 0.25*s, (m[0, 1]+m[1, 0])/s, (m[0, 2]+m[2, 0])/s]
  elif m[1, 1] > m[2, 2]:
    s = math.sqrt(1.0 + m[1, 1] - m[0, 0] - m[2, 2]) * 2
    q = [(m[0, 2]-m[2, 0])/s, (m[0, 1]+m[1, 0])/s, 0.25*s, (m[1, 2]+m[2, 1])/s]
  else:
    s = math.sqrt(1.0 + m[2, 2] - m[0, 0] - m[1, 1]) * 2
    q = [(m[1, 0]-m[0, 1])/s, (m[0, 2]+m[2, 0])/s, (m[1, 2]+m[2, 1])/s, 0.25*s]
  q = np.array(q)
  return q / np.linalg.norm(q)


def quat2mat(q):
  w, x, y, z = q
  return np.array([[w*w+x*x-y*y-z*z, 2*(x*y-w*z), 2*(x*z+w*y)],
                   [2*(x*y+w*z), w*w-x*x+y*y-z*z, 2*(y*z-w*x)],
                   [2*(x*z-w*y), 2*(y*z+w*x), w*w-x*x-y*y+z*z]])


def _geom_mass_inertia(g: GeomSpec, size):
  """Mass and diagonal inertia (geom frame) of a primitive; mirrors the
  closed forms MuJoCo's compiler uses for uniform-density primitives."""
  t = g.type
  r, h = size[0], size[1]
  if t == "sphere":
    vol = 4.0/3.0*math.pi*r**3
    unit = np.full(3, 0.4*r*r)
  elif t == "capsule":
    vc, vs = math.pi*r*r*2*h, 4.0/3.0*math.pi*r**3
    vol = vc + vs
    izz = (vc*r*r/2 + vs*0.4*r*r)/vol
    ixx = (vc*(r*r/4 + h*h/3) + vs*(0.4*r*r + h*h + 0.75*h*r))/vol
    unit = np.array([ixx, ixx, izz])
  elif t == "cylinder":
    vol = math.pi*r*r*2*h
    unit = np.array([(3*r*r + 4*h*h)/12, (3*r*r + 4*h*h)/12, r*r/2])
  elif t == "box":
    a, b, c = size
    vol = 8*a*b*c
    unit = np.array([(b*b+c*c)/3, (a*a+c*c)/3, (a*a+b*b)/3])
  elif t == "ellipsoid":
    a, b, c = size
    vol = 4.0/3.0*math.pi*a*b*c
    unit = np.array([(b*b+c*c)/5, (a*a+c*c)/5, (a*a+b*b)/5])
  else:
    return 0.0, np.zeros(3)
  mass = g.mass if g.mass is not None else g.density*vol
  return mass, unit*mass


def mesh_mass_properties(verts: np.ndarray):
  """Volume, centre of mass and inertia tensor about it (unit density) of the convex hull of `verts`, by summing the
  tetrahedra (c0, a, b, c) over the hull's triangles: integral x x^T over a tetrahedron with one vertex at the origin is
  det/120 (a a^T + b b^T + c c^T + (a+b+c)(a+b+c)^T).  (MuJoCo integrates the mesh itself; for convex meshes the two
  coincide, and collision uses the hull anyway.)"""
  from scipy.spatial import ConvexHull
  v = np.asarray(verts, float)
  hull = ConvexHull(v)
  c0 = v[hull.vertices].mean(axis=0)
  vol, com, C = 0.0, np.zeros(3), np.zeros((3, 3))
  for tri in hull.simplices:
    a, b, c = v[tri] - c0
    d = abs(float(np.linalg.det(np.stack([a, b, c]))))
    S = a + b + c
    vol += d/6; com += d/24*S
    C += d/120*(np.outer(a, a) + np.outer(b, b) + np.outer(c, c) + np.outer(S, S))
  com /= vol
  Cc = C - vol*np.outer(com, com)
  return vol, com + c0, np.trace(Cc)*np.eye(3) - Cc


def _mat2quat(R):
  """Unit quaternion (w, x, y, z) of a proper rotation matrix."""
  t = np.trace(R)
  if t > 0:
    s = math.sqrt(t + 1.0)*2; q = [0.25*s, (R[2, 1] - R[1, 2])/s, (R[0, 2] - R[2, 0])/s, (R[1, 0] - R[0, 1])/s]
  else:
    i = int(np.argmax(np.diag(R))); j, k = (i + 1) % 3, (i + 2) % 3
    s = math.sqrt(1.0 + R[i, i] - R[j, j] - R[k, k])*2
    q = [0.0, 0.0, 0.0, 0.0]
    q[0] = (R[k, j] - R[j, k])/s; q[1 + i] = 0.25*s; q[1 + j] = (R[j, i] + R[i, j])/s; q[1 + k] = (R[k, i] + R[i, k])/s
  q = np.array(q); return q/np.linalg.norm(q)


def load_stl(path) -> np.ndarray:
  """Vertices of a binary STL (80-byte header, uint32 count, 50 bytes per triangle)."""
  with open(path, "rb") as f:
    b = f.read()
  n = int.from_bytes(b[80:84], "little")
  if len(b) < 84 + 50*n:
    raise ValueError(f"{path}: not a binary STL")
  tri = np.frombuffer(b, dtype=np.dtype([("n", "<f4", 3), ("v", "<f4", (3, 3)), ("a", "<u2")]), count=n, offset=84)
  return tri["v"].reshape(-1, 3).astype(np.float64)


def convex_hull_vertices(v: np.ndarray) -> np.ndarray:
  """Hull vertices in first-occurrence order of the input (qhull via scipy; MuJoCo's compiler uses qhull too).
  Coordinates are rounded to float32 first: that is what the device holds, and it keeps the hull reproducible."""
  v = np.unique(np.asarray(v, np.float32).astype(np.float64), axis=0)
  if len(v) <= 4:
    return v
  from scipy.spatial import ConvexHull
  try:
    idx = np.sort(ConvexHull(v).vertices)
  except Exception:       # flat or degenerate input: keep every vertex
    return v
  return v[idx]


def _resolve_geom(g: GeomSpec):
  """size/pos/quat after `fromto` resolution (MuJoCo: z axis along from-to, pos = midpoint)."""
  size, pos, quat = list(g.size), np.array(g.pos, float), np.array(g.quat, float)
  if g.fromto is not None:
    ft = np.array(g.fromto, float)
    vec = ft[0:3] - ft[3:6]
    size[1] = np.linalg.norm(vec)/2
    pos = 0.5*(ft[0:3] + ft[3:6])
    quat = quat_z2vec(vec)
    if g.type == "box" or g.type == "ellipsoid":
      size[2] = size[1]; size[1] = size[0]
  return size, pos, quat


def _compile(spec: Spec) -> HostModel:
  bodies, parent = [], []

  def rec(b, p):
    i = len(bodies)
    bodies.append(b); parent.append(p)
    for c in b.bodies:
      rec(c, i)
  rec(spec.worldbody, 0)
  nbody = len(bodies)
  m = HostModel()
  m.names = dict(body=[b.name for b in bodies], joint=[], geom=[], site=[],
                 actuator=[a.name for a in spec.actuators], sensor=[s.name for s in spec.sensors])
  body_pos = np.zeros((nbody, 3)); body_quat = np.tile([1.0, 0, 0, 0], (nbody, 1))
  body_ipos = np.zeros((nbody, 3)); body_iquat = np.tile([1.0, 0, 0, 0], (nbody, 1))
  body_mass = np.zeros(nbody); body_inertia = np.zeros((nbody, 3))
  body_jntnum = np.zeros(nbody, np.int32); body_jntadr = -np.ones(nbody, np.int32)
  body_dofnum = np.zeros(nbody, np.int32); body_dofadr = -np.ones(nbody, np.int32)
  body_mocapid = -np.ones(nbody, np.int32)
  J = dict(type=[], qposadr=[], dofadr=[], bodyid=[], limited=[], pos=[], axis=[], stiffness=[],
           range=[], margin=[], solref=[], solimp=[])
  D = dict(bodyid=[], jntid=[], armature=[], damping=[], frictionloss=[])
  G = dict(dataid=[], type=[], bodyid=[], contype=[], conaffinity=[], condim=[], priority=[], size=[], pos=[],
           quat=[], friction=[], solmix=[], solref=[], solimp=[], margin=[], gap=[], rbound=[], rgba=[])
  mesh_index: dict[str, int] = {}
  mesh_hulls: list = []

  def mesh_id(name):
    """Index + convex-hull vertices (geom frame) of mesh `name`; MuJoCo collides mesh geoms by their convex hull."""
    if name in mesh_index:
      return mesh_index[name], mesh_hulls[mesh_index[name]]
    if name not in spec.meshes:
      raise ValueError(f"unknown mesh {name!r}")
    ms = spec.meshes[name]
    v = ms.vertex if ms.vertex is not None else load_stl(ms.file)
    v = np.asarray(v, float)*np.asarray(ms.scale, float)[None, :]
    hull = convex_hull_vertices(v)
    mesh_index[name] = len(mesh_hulls); mesh_hulls.append(hull)
    return mesh_index[name], hull
  S = dict(bodyid=[], pos=[], quat=[])
  qpos0, qpos_spring = [], []
  nq = nv = nmocap = 0
  is_static = [True]*nbody
  for i, b in enumerate(bodies):
    is_static[i] = (not b.joints) and (i == 0 or is_static[parent[i]])
    body_pos[i] = b.pos
    q = np.array(b.quat, float); body_quat[i] = q/np.linalg.norm(q)
    if b.mocap:
      body_mocapid[i] = nmocap; nmocap += 1
    # joints
    if b.joints:
      body_jntadr[i] = len(J["type"]); body_jntnum[i] = len(b.joints); body_dofadr[i] = nv
    for j in b.joints:
      jt = JNT[j.type]
      jid = len(J["type"])
      m.names["joint"].append(j.name)
      J["type"].append(jt); J["qposadr"].append(nq); J["dofadr"].append(nv); J["bodyid"].append(i)
      J["pos"].append(j.pos)
      ax = np.array(j.axis, float); n = np.linalg.norm(ax)
      J["axis"].append(list(ax/n) if n > 0 else [0.0, 0, 1])
      lim = j.limited
      limited = (j.range[0] < j.range[1]) if lim == "auto" else (lim == "true")
      if jt == 0:
        limited = False
      J["limited"].append(int(limited)); J["range"].append(j.range); J["margin"].append(j.margin)
      J["stiffness"].append(j.stiffness); J["solref"].append(j.solreflimit); J["solimp"].append(j.solimplimit)
      ndof = {0: 6, 1: 3, 2: 1, 3: 1}[jt]
      if jt == 0:
        qpos0 += list(body_pos[i]) + list(body_quat[i]); qpos_spring += list(body_pos[i]) + list(body_quat[i]); nq += 7
      elif jt == 1:
        qpos0 += [1.0, 0, 0, 0]; qpos_spring += [1.0, 0, 0, 0]; nq += 4
      else:
        qpos0.append(j.ref); qpos_spring.append(j.springref); nq += 1
      for _ in range(ndof):
        D["bodyid"].append(i); D["jntid"].append(jid); D["armature"].append(j.armature)
        D["damping"].append(j.damping); D["frictionloss"].append(j.frictionloss)
      nv += ndof
    body_dofnum[i] = nv - body_dofadr[i] if b.joints else 0
    # geoms
    acc = []
    for g in b.geoms:
      m.names["geom"].append(g.name)
      size, pos, quat = _resolve_geom(g)
      t = GEOM[g.type]
      G["type"].append(t); G["bodyid"].append(i)
      for k in ("contype", "conaffinity", "condim", "priority", "solmix", "margin", "gap"):
        G[k].append(getattr(g, k))
      G["size"].append(size); G["pos"].append(list(pos)); G["quat"].append(list(quat/np.linalg.norm(quat)))
      G["friction"].append(list(g.friction)); G["solref"].append(list(g.solref)); G["solimp"].append(list(g.solimp))
      G["rgba"].append(list(g.rgba))
      if t == 2: rb = size[0]
      elif t == 3: rb = size[0] + size[1]
      elif t == 5: rb = math.hypot(size[0], size[1])
      elif t in (4,): rb = max(size)
      elif t == 6: rb = float(np.linalg.norm(size))
      else: rb = 0.0
      dataid = -1
      if t == 7 and (g.contype or g.conaffinity):
        dataid, hull = mesh_id(g.mesh)
        rb = float(np.linalg.norm(hull, axis=1).max())
        G["size"][-1] = list(np.abs(hull).max(axis=0))
      if t == 1:
        if g.hfield not in spec.hfields:
          raise ValueError(f"geom {g.name!r}: unknown hfield {g.hfield!r}")
        dataid = list(spec.hfields).index(g.hfield)
        hs = spec.hfields[g.hfield].size
        rb = float(np.linalg.norm([hs[0], hs[1], max(hs[2], hs[3])]))
        G["size"][-1] = [hs[0], hs[1], max(hs[2], hs[3])]
      G["rbound"].append(rb); G["dataid"].append(dataid)
      if t == 7:
        if b.inertial is None and not is_static[i] and (g.mass or (g.mass is None and g.density > 0)):
          # mesh inertia: uniform density over the convex hull, principal axes found by diagonalising the tensor
          if g.mesh not in spec.meshes:
            raise ValueError(f"unknown mesh {g.mesh!r}")
          ms = spec.meshes[g.mesh]
          mv = np.asarray(ms.vertex if ms.vertex is not None else load_stl(ms.file), float)*np.asarray(ms.scale, float)[None, :]
          vol, mcom, I = mesh_mass_properties(mv)
          mass = g.mass if g.mass is not None else g.density*vol
          ev, R = np.linalg.eigh(I*(mass/vol))
          if np.linalg.det(R) < 0:
            R[:, 2] = -R[:, 2]
          gq = np.array(quat, float)
          acc.append((mass, ev, np.array(pos) + quat2mat(gq) @ mcom, quat_mul(gq, _mat2quat(R))))
        continue
      if t in (0, 1):
        continue
      gm, gi = _geom_mass_inertia(g, size)
      if gm > 0:
        acc.append((gm, gi, np.array(pos), np.array(quat)))
    # inertial frame
    if b.inertial is not None:
      it = b.inertial
      body_mass[i] = it.mass; body_ipos[i] = it.pos
      if it.fullinertia is not None:
        xx, yy, zz, xy, xz, yz = it.fullinertia
        w, V = np.linalg.eigh(np.array([[xx, xy, xz], [xy, yy, yz], [xz, yz, zz]]))
        order = np.argsort(-w); w, V = w[order], V[:, order]
        if np.linalg.det(V) < 0: V[:, 2] *= -1
        body_inertia[i] = w; body_iquat[i] = quat_mul(np.array(it.quat), mat2quat(V))
      else:
        body_inertia[i] = it.diaginertia
        q = np.array(it.quat, float); body_iquat[i] = q/np.linalg.norm(q)
    elif len(acc) == 1:
      gm, gi, gp, gq = acc[0]
      body_mass[i] = gm; body_inertia[i] = gi; body_ipos[i] = gp; body_iquat[i] = gq/np.linalg.norm(gq)
    elif len(acc) > 1:
      mt = sum(a[0] for a in acc)
      com = sum(a[0]*a[2] for a in acc)/mt
      I = np.zeros((3, 3))
      for gm, gi, gp, gq in acc:
        R = quat2mat(gq/np.linalg.norm(gq)); d = gp - com
        I += R @ np.diag(gi) @ R.T + gm*(np.dot(d, d)*np.eye(3) - np.outer(d, d))
      w, V = np.linalg.eigh(I)
      order = np.argsort(-w); w, V = w[order], V[:, order]
      if np.linalg.det(V) < 0: V[:, 2] *= -1
      body_mass[i] = mt; body_ipos[i] = com; body_inertia[i] = w; body_iquat[i] = mat2quat(V)
    for s in b.sites:
      m.names["site"].append(s.name)
      S["bodyid"].append(i); S["pos"].append(s.pos)
      q = np.array(s.quat, float); S["quat"].append(list(q/np.linalg.norm(q)))

  def arr(x, dt=float, shape=None):
    a = np.array(x, dtype=dt)
    if shape is not None:
      a = a.reshape(shape)
    return a
  njnt, ngeom, nsite = len(J["type"]), len(G["type"]), len(S["bodyid"])
  m.nq, m.nv, m.nbody, m.njnt, m.ngeom, m.nsite, m.nmocap = nq, nv, nbody, njnt, ngeom, nsite, nmocap
  m.na = 0
  m.body_parentid = arr(parent, np.int32)
  m.body_pos, m.body_quat, m.body_ipos, m.body_iquat = body_pos, body_quat, body_ipos, body_iquat
  m.body_mass, m.body_inertia = body_mass, body_inertia
  m.body_jntnum, m.body_jntadr, m.body_dofnum, m.body_dofadr = body_jntnum, body_jntadr, body_dofnum, body_dofadr
  m.body_mocapid = body_mocapid
  m.jnt_type = arr(J["type"], np.int32); m.jnt_qposadr = arr(J["qposadr"], np.int32)
  m.jnt_dofadr = arr(J["dofadr"], np.int32); m.jnt_bodyid = arr(J["bodyid"], np.int32)
  m.jnt_limited = arr(J["limited"], np.int32)
  m.jnt_pos = arr(J["pos"], float, (njnt, 3)); m.jnt_axis = arr(J["axis"], float, (njnt, 3))
  m.jnt_stiffness = arr(J["stiffness"]); m.jnt_range = arr(J["range"], float, (njnt, 2))
  m.jnt_margin = arr(J["margin"]); m.jnt_solref = arr(J["solref"], float, (njnt, 2))
  m.jnt_solimp = arr(J["solimp"], float, (njnt, 5))
  m.qpos0 = arr(qpos0); m.qpos_spring = arr(qpos_spring)
  m.dof_bodyid = arr(D["bodyid"], np.int32); m.dof_jntid = arr(D["jntid"], np.int32)
  m.dof_armature = arr(D["armature"]); m.dof_damping = arr(D["damping"]); m.dof_frictionloss = arr(D["frictionloss"])
  m.geom_type = arr(G["type"], np.int32); m.geom_bodyid = arr(G["bodyid"], np.int32)
  for k in ("contype", "conaffinity", "condim", "priority"):
    setattr(m, "geom_" + k, arr(G[k], np.int32))
  m.geom_size = arr(G["size"], float, (ngeom, 3)); m.geom_pos = arr(G["pos"], float, (ngeom, 3))
  m.geom_quat = arr(G["quat"], float, (ngeom, 4)); m.geom_friction = arr(G["friction"], float, (ngeom, 3))
  m.geom_solmix = arr(G["solmix"]); m.geom_solref = arr(G["solref"], float, (ngeom, 2))
  m.geom_solimp = arr(G["solimp"], float, (ngeom, 5)); m.geom_margin = arr(G["margin"]); m.geom_gap = arr(G["gap"])
  m.geom_rbound = arr(G["rbound"]); m.geom_rgba = arr(G["rgba"], float, (ngeom, 4))
  m.geom_dataid = arr(G["dataid"], np.int32)
  # convex meshes (hull vertices of the meshes that colliding geoms use) and height fields
  m.nmesh = len(mesh_hulls); m.names["mesh"] = list(mesh_index)
  m.mesh_vertnum = arr([len(h) for h in mesh_hulls], np.int32)
  m.mesh_vertadr = arr(np.concatenate([[0], np.cumsum(m.mesh_vertnum)[:-1]]) if m.nmesh else [], np.int32)
  m.mesh_vert = np.concatenate(mesh_hulls, axis=0) if m.nmesh else np.zeros((0, 3))
  m.nmeshvert = int(m.mesh_vert.shape[0])
  hf = list(spec.hfields.values())
  m.nhfield = len(hf); m.names["hfield"] = [h.name for h in hf]
  m.hfield_size = arr([h.size for h in hf], float, (len(hf), 4))
  m.hfield_nrow = arr([h.nrow for h in hf], np.int32); m.hfield_ncol = arr([h.ncol for h in hf], np.int32)
  m.hfield_adr = arr(np.concatenate([[0], np.cumsum([h.nrow*h.ncol for h in hf])[:-1]]) if hf else [], np.int32)
  m.hfield_data = (np.concatenate([np.asarray(h.data, np.float32).reshape(-1) for h in hf]) if hf else np.zeros(0, np.float32))
  m.nhfielddata = int(m.hfield_data.size)
  m.site_bodyid = arr(S["bodyid"], np.int32); m.site_pos = arr(S["pos"], float, (nsite, 3))
  m.site_quat = arr(S["quat"], float, (nsite, 4))
  # actuators
  nu = len(spec.actuators)
  m.nu = nu
  jn = m.names["joint"]
  m.actuator_trnid = np.zeros((nu, 2), np.int32); m.actuator_gaintype = np.zeros(nu, np.int32)
  m.actuator_biastype = np.zeros(nu, np.int32); m.actuator_gear = np.zeros((nu, 6))
  m.actuator_gainprm = np.zeros((nu, 10)); m.actuator_biasprm = np.zeros((nu, 10))
  m.actuator_ctrllimited = np.zeros(nu, np.int32); m.actuator_ctrlrange = np.zeros((nu, 2))
  m.actuator_forcelimited = np.zeros(nu, np.int32); m.actuator_forcerange = np.zeros((nu, 2))
  # fixed tendons (mjWRAP_JOINT = 1 wraps only)
  nt = len(spec.tendons)
  m.ntendon = nt; m.names["tendon"] = [t.name for t in spec.tendons]
  m.tendon_adr = np.zeros(nt, np.int32); m.tendon_num = np.zeros(nt, np.int32); m.tendon_limited = np.zeros(nt, np.int32)
  m.tendon_stiffness = np.zeros(nt); m.tendon_damping = np.zeros(nt); m.tendon_lengthspring = np.zeros((nt, 2))
  m.tendon_range = np.zeros((nt, 2)); m.tendon_length0 = np.zeros(nt)
  wt, wo, wp = [], [], []
  for k, t in enumerate(spec.tendons):
    m.tendon_adr[k] = len(wo); m.tendon_num[k] = len(t.joints)
    for jname, coef in t.joints:
      if jname not in jn:
        raise ValueError(f"tendon {t.name!r}: unknown joint {jname!r}")
      jid = jn.index(jname)
      if m.jnt_type[jid] in (0, 1):
        raise NotImplementedError("fixed tendons over free/ball joints are not supported")
      wt.append(1); wo.append(jid); wp.append(coef)
      m.tendon_length0[k] += coef * m.qpos0[m.jnt_qposadr[jid]]
    m.tendon_stiffness[k], m.tendon_damping[k] = t.stiffness, t.damping
    sl = list(t.springlength)
    if sl[0] == -1 and sl[1] == -1:  # springlength -1 (unset): use the length at qpos0 (MuJoCo compiler)
      sl = [m.tendon_length0[k], m.tendon_length0[k]]
    m.tendon_lengthspring[k] = sl
    if t.range is not None:
      m.tendon_range[k] = t.range
    m.tendon_limited[k] = int(bool(t.limited))
  m.nwrap = len(wo)
  m.wrap_type = np.asarray(wt, np.int32); m.wrap_objid = np.asarray(wo, np.int32); m.wrap_prm = np.asarray(wp, float)
  m.actuator_trntype = np.zeros(nu, np.int32)
  tn = m.names["tendon"]
  for k, a in enumerate(spec.actuators):
    if a.tendon:
      if a.tendon not in tn:
        raise ValueError(f"actuator {a.name!r}: unknown tendon {a.tendon!r}")
      jid = tn.index(a.tendon); m.actuator_trntype[k] = 3      # mjTRN_TENDON
    else:
      if a.joint not in jn:
        raise ValueError(f"actuator {a.name!r}: unknown joint {a.joint!r}")
      jid = jn.index(a.joint)
      if m.jnt_type[jid] in (0, 1):
        raise NotImplementedError("actuators on free/ball joints are not supported")
    m.actuator_trnid[k] = [jid, -1]; m.actuator_gaintype[k] = a.gaintype; m.actuator_biastype[k] = a.biastype
    m.actuator_gear[k, 0] = a.gear; m.actuator_gainprm[k, :3] = a.gainprm; m.actuator_biasprm[k, :3] = a.biasprm
    cr = a.ctrlrange
    if cr is None and a.inheritrange > 0 and not a.tendon:
      lo, hi = m.jnt_range[jid]
      mean, rad = 0.5*(lo + hi), 0.5*(hi - lo)*a.inheritrange
      cr = [mean - rad, mean + rad]
    if cr is not None and cr[0] < cr[1]:
      m.actuator_ctrllimited[k] = 1; m.actuator_ctrlrange[k] = cr
    if a.forcerange is not None and a.forcerange[0] < a.forcerange[1]:
      m.actuator_forcelimited[k] = 1; m.actuator_forcerange[k] = a.forcerange
  # sensors
  ns = len(spec.sensors)
  m.nsensor = ns
  m.sensor_type = np.zeros(ns, np.int32); m.sensor_objtype = np.zeros(ns, np.int32)
  m.sensor_objid = -np.ones(ns, np.int32); m.sensor_reftype = np.zeros(ns, np.int32)
  m.sensor_refid = -np.ones(ns, np.int32); m.sensor_intprm = np.zeros((ns, 3), np.int32)
  m.sensor_dim = np.zeros(ns, np.int32); m.sensor_adr = np.zeros(ns, np.int32); m.sensor_cutoff = np.zeros(ns)
  namekey = {"body": "body", "xbody": "body", "joint": "joint", "geom": "geom", "site": "site", "actuator": "actuator", "tendon": "tendon"}
  adr = 0
  for k, s in enumerate(spec.sensors):
    m.sensor_type[k] = SENS[s.type]; m.sensor_cutoff[k] = s.cutoff
    for typ, name, tf, idf in ((s.objtype, s.objname, "sensor_objtype", "sensor_objid"),
                               (s.reftype, s.refname, "sensor_reftype", "sensor_refid")):
      if typ:
        lst = m.names[namekey[typ]]
        if name not in lst:
          raise ValueError(f"sensor {s.name!r}: unknown {typ} {name!r}")
        getattr(m, tf)[k] = OBJ[typ]; getattr(m, idf)[k] = lst.index(name)
    if s.type == "contact":
      bits, _, slots = s.intprm
      dim = sum(CONTACT_FIELD_DIM[b] for b in range(7) if bits >> b & 1) * slots
      m.sensor_intprm[k] = s.intprm
    else:
      dim = SENS_DIM[s.type]
    m.sensor_dim[k] = dim; m.sensor_adr[k] = adr; adr += dim
  m.nsensordata = adr
  bn = m.names["body"]
  m.nexclude = len(spec.excludes)
  m.exclude_body1 = arr([bn.index(a) for a, _ in spec.excludes], np.int32)
  m.exclude_body2 = arr([bn.index(b) for _, b in spec.excludes], np.int32)
  m.exclude_signature = ((m.exclude_body1.astype(np.int64) << 16) + m.exclude_body2).astype(np.int32)
  o = spec.option
  m.opt.timestep = o["timestep"]; m.opt.gravity = np.array(o["gravity"], float)
  m.opt.magnetic = np.array(o.get("magnetic", (0.0, -0.5, 0.0)), float)
  m.opt.integrator = {"euler": 0, "rk4": 1, "implicit": 2, "implicitfast": 3}[o["integrator"]]
  m.opt.cone = {"pyramidal": 0, "elliptic": 1}[o["cone"]]
  m.opt.solver = {"pgs": 0, "cg": 1, "newton": 2}[o["solver"]]
  m.opt.iterations = o["iterations"]; m.opt.ls_iterations = o["ls_iterations"]
  m.opt.tolerance = o["tolerance"]; m.opt.ls_tolerance = o["ls_tolerance"]
  m.opt.impratio = o["impratio"]; m.opt.disableflags = o["disableflags"]
  m.keys = spec.keys
  finalize_model(m)
  return m


def filter_names(exprs, names):
  """Names matched by any regex (re.match semantics); mirrors reference utils/string.py:28-33."""
  pats = [re.compile(e) for e in exprs]
  return tuple(n for n in names if any(p.match(n) for p in pats))


def resolve_expr(pattern_map, names, default=0.0):
  """First matching pattern wins; mirrors reference utils/string.py:7-25."""
  pats = [(re.compile(p), v) for p, v in pattern_map.items()]
  out = []
  for n in names:
    for p, v in pats:
      if p.match(n):
        out.append(v); break
    else:
      out.append(default)
  return tuple(out)
