"""Scene assembly that mirrors what mjlab does to a robot MJCF before compiling.

Restates, against this package's `Spec`, the spec edits the reference applies:
  * terrain body + plane geom         (src/mjlab/terrains/terrain_importer.py:154-163)
  * entity attach with "<name>/" prefix (src/mjlab/scene/scene.py:147-152)
  * CollisionCfg.edit_spec            (src/mjlab/utils/spec_config.py:206-237)
  * ActuatorSetCfg.edit_spec          (src/mjlab/utils/spec_config.py:361-414)
  * ContactSensor -> one mjSENS_CONTACT per (primary, field)
                                      (src/mjlab/sensor/contact_sensor.py:164-190,472-533)
  * init_state keyframe               (src/mjlab/entity/entity.py:144-166)
and the task-level settings of Mjlab-Velocity-Flat-Unitree-G1
(src/mjlab/tasks/velocity/velocity_env_cfg.py:53-61, config/g1/env_cfgs.py:16-43).

The robot MJCF itself is read from the reference checkout when model fixtures
are (re)generated (`tools/export_models.py`); the runtime loads the exported
`assets/*.npz` model blobs, so nothing here needs /root/reference on the GPU box.
"""
from __future__ import annotations

import math
import os
from dataclasses import dataclass, field

import numpy as np

from .hostmodel import HostModel
from .mjcf import Spec, filter_names, resolve_expr

ASSET_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "assets")

CONTACT_DATA = {"found": 0, "force": 1, "torque": 2, "dist": 3, "pos": 4, "normal": 5, "tangent": 6}
CONTACT_REDUCE = {"none": 0, "mindist": 1, "maxforce": 2, "netforce": 3}


@dataclass
class CollisionCfg:
  geom_names_expr: tuple
  contype: object = 1
  conaffinity: object = 1
  condim: object = 3
  priority: object = 0
  friction: object = None
  solref: object = None
  solimp: object = None
  disable_other_geoms: bool = True


@dataclass
class ActuatorCfg:
  joint_names_expr: tuple
  effort_limit: float
  stiffness: float
  damping: float
  frictionloss: float = 0.0
  armature: float = 0.0


@dataclass
class ContactMatch:
  mode: str
  pattern: object
  entity: str | None = None
  exclude: tuple = ()


@dataclass
class ContactSensorCfg:
  name: str
  primary: ContactMatch
  secondary: ContactMatch | None = None
  fields: tuple = ("found", "force")
  reduce: str = "maxforce"
  num_slots: int = 1


def _resolve(fieldval, names, default):
  if isinstance(fieldval, dict):
    return resolve_expr(fieldval, names, default)
  return tuple([fieldval] * len(names))


def apply_collision_cfg(spec: Spec, cfg: CollisionCfg) -> None:
  names = tuple(g.name for g in spec.geoms)
  subset = filter_names(cfg.geom_names_expr, names)
  res = {k: _resolve(getattr(cfg, k), subset, d) for k, d in
         (("condim", 1), ("contype", 1), ("conaffinity", 1), ("priority", 0),
          ("friction", None), ("solref", None), ("solimp", None))}
  for i, n in enumerate(subset):
    g = spec.geom(n)
    g.condim, g.contype = res["condim"][i], res["contype"][i]
    g.conaffinity, g.priority = res["conaffinity"][i], res["priority"][i]
    for k in ("friction", "solref", "solimp"):
      v = res[k][i]
      if v is not None:
        getattr(g, k)[:len(v)] = list(v)
  if cfg.disable_other_geoms:
    for g in spec.geoms:
      if g.name not in subset:
        g.contype = g.conaffinity = 0


def apply_actuator_cfgs(spec: Spec, cfgs) -> None:
  jnts = [j for j in spec.joints if j.type != "free"]
  jnames = tuple(j.name for j in jnts)
  pairs = []
  for c in cfgs:
    for n in filter_names(c.joint_names_expr, jnames):
      pairs.append((c, n))
  pairs.sort(key=lambda p: jnames.index(p[1]))
  for c, n in pairs:
    j = spec.joint(n)
    if not (j.range[0] < j.range[1]):
      raise ValueError(f"Joint {n} must be limited for position control")
    j.armature, j.frictionloss = c.armature, c.frictionloss
    spec.add_actuator(name=n, joint=n, gaintype=0, biastype=1, gainprm=[c.stiffness, 0.0, 0.0],
                      biasprm=[0.0, -c.stiffness, -c.damping], inheritrange=1.0,
                      forcerange=[-c.effort_limit, c.effort_limit])


def add_init_state_key(spec: Spec, pos, rot, joint_pos: dict) -> None:
  jn = tuple(j.name for j in spec.joints if j.type != "free")
  jp = resolve_expr(joint_pos, jn, 0.0)
  has_free = any(j.type == "free" for j in spec.joints)
  qpos = (list(pos) + list(rot) if has_free else []) + list(jp)
  name2pos = dict(zip(jn, jp))
  ctrl = [name2pos.get(a.name, 0.0) for a in spec.actuators]
  spec.keys["init_state"] = dict(qpos=qpos, ctrl=ctrl)


def add_contact_sensor(scene: Spec, cfg: ContactSensorCfg, entity_bodies: dict) -> None:
  """entity_bodies: entity name -> tuple of un-prefixed body names (for regex expansion)."""
  prim = cfg.primary
  if prim.entity:
    pats = (prim.pattern,) if isinstance(prim.pattern, str) else tuple(prim.pattern)
    pool = entity_bodies[prim.entity] if prim.mode != "geom" else entity_bodies[prim.entity + "#geoms"]
    prim_names = tuple(n for n in filter_names_full(pats, pool) if n not in prim.exclude)
  else:
    prim_names = (prim.pattern,) if isinstance(prim.pattern, str) else tuple(prim.pattern)
  sec_name = None
  if cfg.secondary is not None:
    sec_name = cfg.secondary.pattern
    if cfg.secondary.entity and cfg.secondary.mode != "subtree":
      sec_name = filter_names_full((sec_name,), entity_bodies[cfg.secondary.entity])[0]
  objt = {"geom": "geom", "body": "body", "subtree": "xbody"}
  for p in prim_names:
    for f in cfg.fields:
      s = scene.add_sensor(name=f"{cfg.name}_{p}_{f}", type="contact", objtype=objt[prim.mode],
                           objname=(f"{prim.entity}/{p}" if prim.entity else p),
                           intprm=[1 << CONTACT_DATA[f], CONTACT_REDUCE[cfg.reduce], cfg.num_slots])
      if sec_name is not None:
        s.reftype = objt[cfg.secondary.mode]
        s.refname = f"{cfg.secondary.entity}/{sec_name}" if cfg.secondary.entity else sec_name


def filter_names_full(pats, names):
  import re
  out = []
  for n in names:
    if any(re.fullmatch(p, n) for p in pats):
      out.append(n)
  return tuple(out)


# ---------------------------------------------------------------------------
# Unitree G1 (values restated from src/mjlab/asset_zoo/robots/unitree_g1/g1_constants.py:44-283)
# ---------------------------------------------------------------------------
def _refl(rotor, gear):
  return rotor[0]*(gear[1]*gear[2])**2 + rotor[1]*gear[2]**2 + rotor[2]


G1_ARM_5020 = _refl((0.139e-4, 0.017e-4, 0.169e-4), (1, 1 + 46/18, 1 + 56/16))
G1_ARM_7520_14 = _refl((0.489e-4, 0.098e-4, 0.533e-4), (1, 4.5, 1 + 48/22))
G1_ARM_7520_22 = _refl((0.489e-4, 0.109e-4, 0.738e-4), (1, 4.5, 5))
G1_ARM_4010 = _refl((0.068e-4, 0.0, 0.0), (1, 5, 5))
_WN = 10 * 2.0 * 3.1415926535
_ZETA = 2.0


def _act(exprs, arm, effort, scale=1.0):
  a = arm*scale
  return ActuatorCfg(joint_names_expr=exprs, effort_limit=effort*scale, armature=a,
                     stiffness=a*_WN**2, damping=2.0*_ZETA*a*_WN)


G1_ACTUATORS = (
  _act((".*_elbow_joint", ".*_shoulder_pitch_joint", ".*_shoulder_roll_joint",
        ".*_shoulder_yaw_joint", ".*_wrist_roll_joint"), G1_ARM_5020, 25.0),
  _act((".*_hip_pitch_joint", ".*_hip_yaw_joint", "waist_yaw_joint"), G1_ARM_7520_14, 88.0),
  _act((".*_hip_roll_joint", ".*_knee_joint"), G1_ARM_7520_22, 139.0),
  _act((".*_wrist_pitch_joint", ".*_wrist_yaw_joint"), G1_ARM_4010, 5.0),
  _act(("waist_pitch_joint", "waist_roll_joint"), G1_ARM_5020, 25.0, 2.0),
  _act((".*_ankle_pitch_joint", ".*_ankle_roll_joint"), G1_ARM_5020, 25.0, 2.0),
)
_FOOT = r"^(left|right)_foot[1-7]_collision$"
G1_FULL_COLLISION = CollisionCfg(
  geom_names_expr=(".*_collision",), condim={_FOOT: 3, ".*_collision": 1},
  priority={_FOOT: 1}, friction={_FOOT: (0.6,)})
G1_KNEES_BENT = dict(pos=(0.0, 0.0, 0.76), rot=(1.0, 0.0, 0.0, 0.0), joint_pos={
  ".*_hip_pitch_joint": -0.312, ".*_knee_joint": 0.669, ".*_ankle_pitch_joint": -0.363,
  ".*_elbow_joint": 0.6, "left_shoulder_roll_joint": 0.2, "left_shoulder_pitch_joint": 0.2,
  "right_shoulder_roll_joint": -0.2, "right_shoulder_pitch_joint": 0.2})


def g1_action_scale(joint_names) -> np.ndarray:
  """0.25 * effort / stiffness per actuated joint (g1_constants.py:286-297)."""
  out = np.zeros(len(joint_names))
  for c in G1_ACTUATORS:
    for n in filter_names(c.joint_names_expr, tuple(joint_names)):
      out[list(joint_names).index(n)] = 0.25 * c.effort_limit / c.stiffness
  return out


def velocity_sim_options(spec: Spec) -> None:
  """SIM_CFG of the velocity task (velocity_env_cfg.py:53-61) + MujocoCfg defaults (sim.py:41-62)."""
  spec.option.update(timestep=0.005, iterations=10, ls_iterations=20, integrator="implicitfast",
                     solver="newton", cone="pyramidal", tolerance=1e-8, ls_tolerance=0.01,
                     impratio=1.0, gravity=[0.0, 0.0, -9.81])


def build_robot_entity(xml_path, collision: CollisionCfg | None, actuators, init_state) -> Spec:
  spec = Spec.from_file(xml_path)
  if collision is not None:
    apply_collision_cfg(spec, collision)
  if actuators:
    apply_actuator_cfgs(spec, actuators)
  if init_state is not None:
    add_init_state_key(spec, **init_state)
  return spec


def build_scene(entities: dict, sensors=(), plane=True, num_origin_sites=0, env_spacing=2.5) -> Spec:
  scene = Spec()
  if plane:
    t = scene.worldbody.add_body(name="terrain")
    t.add_geom(name="terrain", type="plane", size=[0.0, 0.0, 0.01])
  origins = env_origins_grid(num_origin_sites, env_spacing) if num_origin_sites else ()
  for e in range(num_origin_sites):  # SURVEY.md §0 fact 10 (one site per env on the world body)
    scene.worldbody.add_site(name=f"env_origin_{e}", pos=list(origins[e]))
  ent_bodies = {}
  for name, espec in entities.items():
    ent_bodies[name] = tuple(b.name for b in espec.bodies[1:])
    ent_bodies[name + "#geoms"] = tuple(g.name for g in espec.geoms if g.name)
    scene.attach(espec, prefix=f"{name}/")
  for s in sensors:
    add_contact_sensor(scene, s, ent_bodies)
  return scene


def env_origins_grid(num_envs: int, spacing: float) -> np.ndarray:
  """Grid origins (src/mjlab/terrains/terrain_importer.py:225-240)."""
  rows = math.ceil(num_envs / int(math.sqrt(num_envs)))
  cols = math.ceil(num_envs / rows)
  ii, jj = np.meshgrid(np.arange(rows), np.arange(cols), indexing="ij")
  o = np.zeros((num_envs, 3))
  o[:, 0] = -(ii.flatten()[:num_envs] - (rows - 1) / 2) * spacing
  o[:, 1] = (jj.flatten()[:num_envs] - (cols - 1) / 2) * spacing
  return o


def with_env_origin_sites(model: HostModel, num_envs: int, env_spacing: float = 2.5) -> HostModel:
  """Copy of a compiled scene with the visualisation sites mjlab's terrain importer adds to the world body: one `env_origin_<i>`
  per environment at its grid origin (terrains/terrain_importer.py:95,98-120).  World-body sites precede the entities' sites in
  MuJoCo's numbering, so every site reference (site ids of sensors) moves up by `num_envs` -- as in a scene the reference
  compiles itself.  This is the O(num_envs) model array that makes `site_xpos` O(num_envs^2) in the reference (SURVEY.md §0
  fact 10); the engine stores such static frames once (Data.site_xpos becomes a SplitField)."""
  import copy
  m = copy.copy(model)
  m.names = dict(model.names)
  n = int(num_envs)
  o = env_origins_grid(n, env_spacing)
  m.site_bodyid = np.concatenate([np.zeros(n, np.int32), np.asarray(model.site_bodyid, np.int32)])
  m.site_pos = np.concatenate([o, np.asarray(model.site_pos, float).reshape(-1, 3)])
  q = np.zeros((n, 4)); q[:, 0] = 1.0
  m.site_quat = np.concatenate([q, np.asarray(model.site_quat, float).reshape(-1, 4)])
  m.names["site"] = [f"env_origin_{e}" for e in range(n)] + list(model.names["site"])
  m.nsite = model.nsite + n
  OBJ_SITE = 6
  oid = np.asarray(model.sensor_objid, np.int32).copy(); rid = np.asarray(model.sensor_refid, np.int32).copy()
  oid[np.asarray(model.sensor_objtype) == OBJ_SITE] += n
  rid[(np.asarray(model.sensor_reftype) == OBJ_SITE) & (rid >= 0)] += n
  m.sensor_objid, m.sensor_refid = oid, rid
  return m


def build_g1_flat(xml_path) -> HostModel:
  """Compiled model of Mjlab-Velocity-Flat-Unitree-G1 (one world)."""
  robot = build_robot_entity(xml_path, G1_FULL_COLLISION, G1_ACTUATORS, G1_KNEES_BENT)
  feet = ContactSensorCfg(
    name="feet_ground_contact",
    primary=ContactMatch("subtree", r"^(left_ankle_roll_link|right_ankle_roll_link)$", "robot"),
    secondary=ContactMatch("body", "terrain"), fields=("found", "force"), reduce="netforce", num_slots=1)
  selfc = ContactSensorCfg(
    name="self_collision", primary=ContactMatch("subtree", "pelvis", "robot"),
    secondary=ContactMatch("subtree", "pelvis", "robot"), fields=("found",), reduce="none", num_slots=1)
  scene = build_scene({"robot": robot}, sensors=(feet, selfc))
  velocity_sim_options(scene)
  m = scene.compile()
  m.meta = dict(task="Mjlab-Velocity-Flat-Unitree-G1", nconmax=50, njmax=300, decimation=4)
  return m


def build_g1_tracking(xml_path) -> HostModel:
  """Compiled model of Mjlab-Tracking-Flat-Unitree-G1 (tasks/tracking/config/g1/env_cfgs.py:19-82,
  tracking_env_cfg.py:47-62): the same robot entity, only the self-collision sensor, nconmax 35 / njmax 250."""
  robot = build_robot_entity(xml_path, G1_FULL_COLLISION, G1_ACTUATORS, G1_KNEES_BENT)
  selfc = ContactSensorCfg(
    name="self_collision", primary=ContactMatch("subtree", "pelvis", "robot"),
    secondary=ContactMatch("subtree", "pelvis", "robot"), fields=("found",), reduce="none", num_slots=1)
  scene = build_scene({"robot": robot}, sensors=(selfc,))
  velocity_sim_options(scene)
  m = scene.compile()
  m.meta = dict(task="Mjlab-Tracking-Flat-Unitree-G1", nconmax=35, njmax=250, decimation=4)
  return m


# ---------------------------------------------------------------------------
# Unitree Go1 (values restated from src/mjlab/asset_zoo/robots/unitree_go1/go1_constants.py:33-153)
# ---------------------------------------------------------------------------
_GO1_ROTOR = 0.000111842
_GO1_HIP_ARM = _GO1_ROTOR * 6.0 ** 2
_GO1_KNEE_ARM = _GO1_ROTOR * 9.0 ** 2
GO1_ACTUATORS = (
  ActuatorCfg((".*_hip_joint", ".*_thigh_joint"), effort_limit=23.7, stiffness=_GO1_HIP_ARM * _WN ** 2,
              damping=2.0 * _ZETA * _GO1_HIP_ARM * _WN, armature=_GO1_HIP_ARM),
  ActuatorCfg((".*_calf_joint",), effort_limit=35.55, stiffness=_GO1_KNEE_ARM * _WN ** 2,
              damping=2.0 * _ZETA * _GO1_KNEE_ARM * _WN, armature=_GO1_KNEE_ARM),
)
_GO1_FOOT = "^[FR][LR]_foot_collision$"
GO1_FULL_COLLISION = CollisionCfg(
  geom_names_expr=(".*_collision",), condim={_GO1_FOOT: 3, ".*_collision": 1}, priority={_GO1_FOOT: 1},
  friction={_GO1_FOOT: (0.6,)}, solimp={_GO1_FOOT: (0.9, 0.95, 0.023)}, contype=1, conaffinity=0)
GO1_INIT = dict(pos=(0.0, 0.0, 0.278), rot=(1.0, 0.0, 0.0, 0.0), joint_pos={
  ".*thigh_joint": 0.9, ".*calf_joint": -1.8, ".*R_hip_joint": 0.1, ".*L_hip_joint": -0.1})


def go1_action_scale(joint_names) -> np.ndarray:
  out = np.zeros(len(joint_names))
  for c in GO1_ACTUATORS:
    for n in filter_names(c.joint_names_expr, tuple(joint_names)):
      out[list(joint_names).index(n)] = 0.25 * c.effort_limit / c.stiffness
  return out


def _go1_sensors():
  """Contact sensors of Mjlab-Velocity-*-Unitree-Go1 (tasks/velocity/config/go1/env_cfgs.py:24-47)."""
  geom_names = tuple(f"{n}_foot_collision" for n in ("FR", "FL", "RR", "RL"))
  feet = ContactSensorCfg(
    name="feet_ground_contact", primary=ContactMatch("geom", geom_names, "robot"),
    secondary=ContactMatch("body", "terrain"), fields=("found", "force"), reduce="netforce", num_slots=1)
  nonfoot = ContactSensorCfg(
    name="nonfoot_ground_touch", primary=ContactMatch("geom", r".*_collision\d*$", "robot", exclude=geom_names),
    secondary=ContactMatch("body", "terrain"), fields=("found",), reduce="none", num_slots=1)
  return feet, nonfoot


def build_go1(xml_path, terrain="plane", terrain_seed=0) -> HostModel:
  """Compiled model of Mjlab-Velocity-Flat-Unitree-Go1 (terrain="plane") or of the Rough task with the reference's
  default box terrain (terrain="rough": terrains/config.py:7-27, generated by terrains.py)."""
  robot = build_robot_entity(xml_path, GO1_FULL_COLLISION, GO1_ACTUATORS, GO1_INIT)
  if terrain == "plane":
    scene = build_scene({"robot": robot}, sensors=_go1_sensors())
  else:
    from . import terrains
    scene = Spec()
    # "rough_hf": the height-field sub-terrains that terrains/config.py:28-53 keeps commented out, enabled
    terrains.add_rough_terrain(scene, seed=terrain_seed, cfg=terrains.rough_terrains_hf_cfg() if terrain == "rough_hf" else None)
    ent_bodies = {"robot": tuple(b.name for b in robot.bodies[1:]), "robot#geoms": tuple(g.name for g in robot.geoms if g.name)}
    scene.attach(robot, prefix="robot/")
    for sc in _go1_sensors():
      add_contact_sensor(scene, sc, ent_bodies)
  velocity_sim_options(scene)
  m = scene.compile()
  m.meta = dict(task="Mjlab-Velocity-%s-Unitree-Go1" % ("Flat" if terrain == "plane" else "Rough"), nconmax=50, njmax=300, decimation=4,
                terrain=terrain)
  if terrain != "plane":
    m.meta["terrain_origins"] = np.asarray(scene.terrain_origins).tolist()   # [num_rows][num_cols][3] spawn points
    m.meta["terrain_seed"] = terrain_seed
  return m


# ---------------------------------------------------------------------------
# Asimov-toe (values restated from src/mjlab/asset_zoo/robots/asimov/asimov_toe_constants.py:40-271)
# ---------------------------------------------------------------------------
_ASV_WN = 8 * 2.0 * 3.1415926535      # NATURAL_FREQ, 8 Hz (asimov_toe_constants.py:106)
_ASV_ZETA = 1.8                        # DAMPING_RATIO (:107)


def _asv_act(exprs, arm, effort):
  return ActuatorCfg(tuple(exprs), effort_limit=effort, stiffness=arm * _ASV_WN ** 2,
                     damping=2.0 * _ASV_ZETA * arm * _ASV_WN, armature=arm)


ASIMOV_TOE_ACTUATORS = (
  _asv_act((".*_hip_pitch_joint",), G1_ARM_7520_14, 55.0),
  _asv_act((".*_hip_roll_joint",), G1_ARM_7520_22, 90.0),
  _asv_act((".*_hip_yaw_joint",), G1_ARM_7520_14, 60.0),
  _asv_act((".*_knee_joint",), G1_ARM_7520_22, 50.0),
  ActuatorCfg(("left_toe_joint", "right_toe_joint"), effort_limit=5.0, stiffness=50.0, damping=0.8, armature=0.0001),
)
_ASV_FOOT = r"^(left|right)_(foot|toe)\d+_collision$"
ASIMOV_FEET_ONLY_COLLISION = CollisionCfg(
  geom_names_expr=(_ASV_FOOT,), contype=0, conaffinity=1, condim=3, priority=1, friction=(0.8,))
ASIMOV_FULL_COLLISION = CollisionCfg(
  geom_names_expr=(".*_collision",), condim={_ASV_FOOT: 3, ".*_collision": 1}, priority={_ASV_FOOT: 1},
  friction={_ASV_FOOT: (0.8,)})
ASIMOV_KNEES_BENT = dict(pos=(0.0, 0.0, 0.73), rot=(1.0, 0.0, 0.0, 0.0), joint_pos={
  "left_hip_pitch_joint": 0.2, "right_hip_pitch_joint": -0.2, ".*_hip_roll_joint": 0.0, ".*_hip_yaw_joint": 0.0,
  "left_knee_joint": 0.4, "right_knee_joint": -0.4, "left_ankle_pitch_joint": -0.25, "right_ankle_pitch_joint": 0.25,
  ".*_ankle_roll_joint": 0.0, "left_toe_joint": 0.0, "right_toe_joint": 0.0})
ASIMOV_ANKLE_L, ASIMOV_ANKLE_D = 0.09, 0.02   # tendon coefficients of asimov_toe.xml:176-196


def asimov_action_scale(joint_names) -> np.ndarray:
  """ASIMOV_ACTION_SCALE (asimov_toe_constants.py:283-297): 0.25 * effort / stiffness per actuated joint."""
  out = np.zeros(len(joint_names))
  for c in ASIMOV_TOE_ACTUATORS:
    for n in filter_names(c.joint_names_expr, tuple(joint_names)):
      out[list(joint_names).index(n)] = 0.25 * c.effort_limit / c.stiffness
  return out


def ankle_pr_to_tendon(pitch_l, roll_l, pitch_r, roll_r):
  """AnklePrToTendonAction's linear map (envs/mdp/actions/ankle_ab_action.py): ankle pitch/roll targets ->
  length targets of the four tendon position actuators [L_A, L_B, R_A, R_B]."""
  L, D = ASIMOV_ANKLE_L, ASIMOV_ANKLE_D
  return np.stack([-L * pitch_l - D * roll_l, -L * pitch_l + D * roll_l,
                   L * pitch_r - D * roll_r, L * pitch_r + D * roll_r], axis=-1)


def build_asimov_toe(xml_path, terrain="plane", terrain_seed=0, collision="feet_only") -> HostModel:
  """Compiled model of the fork's Asimov-toe velocity task (tasks/velocity/config/asimov_toe/env_cfgs.py:22-180):
  4 fixed ankle tendons with tendon position actuators from the XML, then the joint actuators of
  ASIMOV_ARTICULATION, KNEES_BENT keyframe.

  collision="feet_only" is the task's own FEET_ONLY_COLLISION (asimov_toe_constants.py:243-250): the 20 foot / toe
  capsules against the terrain; CollisionCfg.disable_other_geoms (spec_config.py:163,233-237) switches the link
  meshes' collision off.  collision="full" is FULL_COLLISION (:219-227): every *_collision geom, link meshes included,
  collides by its convex hull with the terrain and with the other links (K7: plane-mesh, GJK/EPA)."""
  robot = Spec.from_file(xml_path)
  if collision == "full":
    apply_collision_cfg(robot, ASIMOV_FULL_COLLISION)
  else:
    apply_collision_cfg(robot, ASIMOV_FEET_ONLY_COLLISION)
  apply_actuator_cfgs(robot, ASIMOV_TOE_ACTUATORS)
  add_init_state_key(robot, **ASIMOV_KNEES_BENT)
  # zero-action ctrl of the tendon actuators: AnklePrToTendonAction with the default ankle angles as offset
  jp = ASIMOV_KNEES_BENT["joint_pos"]
  ab = ankle_pr_to_tendon(jp["left_ankle_pitch_joint"], jp[".*_ankle_roll_joint"],
                          jp["right_ankle_pitch_joint"], jp[".*_ankle_roll_joint"])
  ctrl = robot.keys["init_state"]["ctrl"]
  for k, n in enumerate(("left_ankle_A", "left_ankle_B", "right_ankle_A", "right_ankle_B")):
    ctrl[[a.name for a in robot.actuators].index(n)] = float(ab[k])
  feet = ContactSensorCfg(
    name="feet_ground_contact",
    primary=ContactMatch("subtree", r"^(left_ankle_roll_link|right_ankle_roll_link)$", "robot"),
    secondary=ContactMatch("body", "terrain"), fields=("found", "force"), reduce="netforce", num_slots=1)
  selfc = ContactSensorCfg(
    name="self_collision", primary=ContactMatch("subtree", "pelvis_link", "robot"),
    secondary=ContactMatch("subtree", "pelvis_link", "robot"), fields=("found",), reduce="none", num_slots=1)
  if terrain == "plane":
    scene = build_scene({"robot": robot}, sensors=(feet, selfc))
  else:
    from . import terrains
    scene = Spec()
    terrains.add_rough_terrain(scene, seed=terrain_seed)
    ent_bodies = {"robot": tuple(b.name for b in robot.bodies[1:]), "robot#geoms": tuple(g.name for g in robot.geoms if g.name)}
    scene.attach(robot, prefix="robot/")
    for sc in (feet, selfc):
      add_contact_sensor(scene, sc, ent_bodies)
  velocity_sim_options(scene)
  m = scene.compile()
  m.meta = dict(task="Mjlab-Velocity-%s-Asimov-Toe" % ("Flat" if terrain == "plane" else "Rough"), nconmax=50, njmax=300,
                decimation=4, collision=collision)
  if terrain != "plane":
    m.meta["terrain_origins"] = np.asarray(scene.terrain_origins).tolist()
    m.meta["terrain_seed"] = terrain_seed
  return m


# ---------------------------------------------------------------------------
# Asimov, the 12-dof biped whose collision geometry is ALL link meshes (values restated from
# src/mjlab/asset_zoo/robots/asimov/asimov_constants.py:40-276; task: tasks/velocity/config/asimov/env_cfgs.py)
# ---------------------------------------------------------------------------
G1_ARM_5020_X2 = 2.0 * (0.139e-4 + 0.017e-4 * (1 + 46 / 18) ** 2 + 0.169e-4 * ((1 + 46 / 18) * (1 + 56 / 16)) ** 2)   # two 5020s on the parallel ankle linkage (:68-82,136-142)
ASIMOV_ACTUATORS = (
  _asv_act((".*_hip_pitch_joint", ".*_hip_yaw_joint"), G1_ARM_7520_14, 88.0),
  _asv_act((".*_hip_roll_joint", ".*_knee_joint"), G1_ARM_7520_22, 139.0),
  _asv_act((".*_ankle_pitch_joint", ".*_ankle_roll_joint"), G1_ARM_5020_X2, 50.0),
)
_ASM_FOOT = r"^(left|right)_ankle_roll_link_collision$"
ASIMOV12_FEET_ONLY_COLLISION = CollisionCfg(geom_names_expr=(_ASM_FOOT,), contype=0, conaffinity=1, condim=3, priority=1, friction=(0.8,))
ASIMOV12_FULL_COLLISION = CollisionCfg(geom_names_expr=(".*_collision",), condim={_ASM_FOOT: 3, ".*_collision": 1}, priority={_ASM_FOOT: 1},
                                       friction={_ASM_FOOT: (0.8,)})
ASIMOV12_KNEES_BENT = dict(pos=(0.0, 0.0, 0.73), rot=(1.0, 0.0, 0.0, 0.0), joint_pos={
  "left_hip_pitch_joint": 0.2, "right_hip_pitch_joint": -0.2, ".*_hip_roll_joint": 0.0, ".*_hip_yaw_joint": 0.0,
  "left_knee_joint": -0.4, "right_knee_joint": 0.4, "left_ankle_pitch_joint": -0.25, "right_ankle_pitch_joint": 0.25,
  ".*_ankle_roll_joint": 0.0})


def asimov12_action_scale(joint_names) -> np.ndarray:
  """ASIMOV_ACTION_SCALE (asimov_constants.py:262-276): 0.3 * effort / stiffness per actuated joint."""
  out = np.zeros(len(joint_names))
  for c in ASIMOV_ACTUATORS:
    for n in filter_names(c.joint_names_expr, tuple(joint_names)):
      out[list(joint_names).index(n)] = 0.3 * c.effort_limit / c.stiffness
  return out


def build_asimov(xml_path, collision="feet_only") -> HostModel:
  """Compiled model of Mjlab-Velocity-Flat-Asimov (tasks/velocity/config/asimov/env_cfgs.py:16-124): free base + 12 hinges,
  13 link meshes as the ONLY collision geometry (asimov.xml:50-110) -- each collides by its convex hull.
  collision="feet_only" is the task's own FEET_ONLY_COLLISION (asimov_constants.py:227-235): the two foot meshes against the
  terrain, a plane-mesh manifold of up to four contacts per foot.  collision="full" is FULL_COLLISION (:195-205): every link
  mesh against the terrain and the other links (12 <exclude> pairs, asimov.xml:140-153): GJK / EPA mesh-mesh."""
  robot = Spec.from_file(xml_path)
  apply_collision_cfg(robot, ASIMOV12_FULL_COLLISION if collision == "full" else ASIMOV12_FEET_ONLY_COLLISION)
  apply_actuator_cfgs(robot, ASIMOV_ACTUATORS)
  add_init_state_key(robot, **ASIMOV12_KNEES_BENT)
  feet = ContactSensorCfg(
    name="feet_ground_contact",
    primary=ContactMatch("subtree", r"^(left_ankle_roll_link|right_ankle_roll_link)$", "robot"),
    secondary=ContactMatch("body", "terrain"), fields=("found", "force"), reduce="netforce", num_slots=1)
  selfc = ContactSensorCfg(
    name="self_collision", primary=ContactMatch("subtree", "pelvis_link", "robot"),
    secondary=ContactMatch("subtree", "pelvis_link", "robot"), fields=("found",), reduce="none", num_slots=1)
  scene = build_scene({"robot": robot}, sensors=(feet, selfc))
  velocity_sim_options(scene)
  m = scene.compile()
  m.meta = dict(task="Mjlab-Velocity-Flat-Asimov", nconmax=50, njmax=300, decimation=4, collision=collision)
  return m


# ---------------------------------------------------------------------------
# model blob (.npz) save / load
# ---------------------------------------------------------------------------
_OPT_KEYS = ("timestep", "integrator", "cone", "solver", "iterations", "ls_iterations", "tolerance",
             "ls_tolerance", "impratio", "disableflags")


def save_model(m: HostModel, path) -> None:
  import json
  d = {k: v for k, v in m.__dict__.items() if isinstance(v, np.ndarray)}
  for k in ("nq", "nv", "nu", "na", "nbody", "njnt", "ngeom", "nsite", "nsensor", "nsensordata", "nexclude", "nmocap", "ntendon", "nwrap", "nmesh", "nmeshvert", "nhfield", "nhfielddata"):
    d["_" + k] = np.int64(getattr(m, k, 0))
  meta = dict(names=m.names, keys=m.keys, opt={k: getattr(m.opt, k) for k in _OPT_KEYS},
              gravity=list(map(float, m.opt.gravity)), meaninertia=m.stat.meaninertia,
              meta=getattr(m, "meta", {}))
  d["_json"] = np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8)
  np.savez_compressed(path, **d)


def load_model(path) -> HostModel:
  import json
  z = np.load(path)
  m = HostModel()
  for k in z.files:
    if k == "_json":
      continue
    if k.startswith("_"):
      setattr(m, k[1:], int(z[k]))
    else:
      setattr(m, k, z[k])
  meta = json.loads(bytes(z["_json"]).decode())
  m.names, m.keys, m.meta = meta["names"], meta["keys"], meta["meta"]
  for k, v in meta["opt"].items():
    setattr(m.opt, k, v)
  m.opt.gravity = np.array(meta["gravity"])
  m.stat.meaninertia = meta["meaninertia"]
  return m


def load_asset(name: str) -> HostModel:
  p = os.path.join(ASSET_DIR, name + ".npz")
  if not os.path.exists(p):
    raise FileNotFoundError(f"model blob {p} missing; regenerate with tools/export_models.py")
  return load_model(p)
