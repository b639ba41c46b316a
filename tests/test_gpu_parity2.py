"""Parity of the CUDA path against the CPU oracle, part 2: the inputs and regimes VERDICT round 1 found untested on the GPU --
xfrc_applied, mocap bodies, qM itself, refresh_static after a per-world terrain write, nconmax / njmax overflow in a subset of
the worlds, disable flags, the contact-sensor match cap, and FREE-RUNNING divergence curves (no teacher forcing).
"""
import copy

import numpy as np
import pytest

import fixtures as fx
from test_gpu_parity import TOL_FORWARD, _check_counts, _rel_err, _setup
from util import g1_states, legged_states

pytestmark = pytest.mark.gpu


def _push(D, worlds, name, arr):
  import torch
  getattr(D, name)[:] = torch.as_tensor(np.asarray(arr), dtype=torch.float32, device="cuda:0").reshape(getattr(D, name).shape)
  for w, a in zip(worlds, np.asarray(arr)):
    w[name][:] = np.asarray(a).reshape(-1)


# --------------------------------------------------------------------------------------------- K12: xfrc_applied
def test_xfrc_applied_on_several_bodies_matches_oracle(g1):
  """Sigma J^T xfrc_applied (k_smooth.cu RNE pass; writers in the reference: envs/mdp/events.py:190 apply_external_force_torque,
  entity/data.py:149-162).  Random wrenches on the torso, a hand and a foot of the G1 in 16 worlds: the smooth force, both
  accelerations and a 10-step horizon against the oracle; clearing the wrench restores the unforced result."""
  import torch
  W = 16
  qpos, qvel, ctrl = g1_states(g1, W, seed=31, spread=0.15, vel=0.2)
  engine, M, D, orc, worlds = _setup(g1, W, qpos, qvel, ctrl, True)
  rng = np.random.default_rng(5)
  names = g1.names["body"]
  bodies = [next(i for i, n in enumerate(names) if n.endswith(k)) for k in ("torso_link", "left_wrist_yaw_link", "right_ankle_roll_link")]
  x = np.zeros((W, g1.nbody, 6))
  x[:, bodies, 0:3] = rng.uniform(-40.0, 40.0, (W, 3, 3))
  x[:, bodies, 3:6] = rng.uniform(-5.0, 5.0, (W, 3, 3))
  x[0] = 0.0                                              # one world without a wrench
  engine.forward(M, D)
  qfs0 = D.qfrc_smooth.clone()
  _push(D, worlds, "xfrc_applied", x)
  engine.forward(M, D)
  for w in worlds:
    orc.forward(w)
  _check_counts(D, worlds)
  for f in ("qfrc_smooth", "qacc_smooth", "qacc", "qfrc_constraint"):
    assert _rel_err(D, worlds, f) < TOL_FORWARD[f], f
  d = (D.qfrc_smooth - qfs0).abs().amax(dim=1).cpu().numpy()
  assert d[0] == 0.0 and (d[1:] > 1.0).all()             # the wrench reached the generalized forces
  engine.step(M, D, 10)
  torch.cuda.synchronize()
  for w in worlds:
    orc.step(w, 10)
  _check_counts(D, worlds)
  assert _rel_err(D, worlds, "qpos") < 1e-4 and _rel_err(D, worlds, "qvel") < 1e-2
  _push(D, worlds, "xfrc_applied", np.zeros_like(x))
  D.qpos[:] = torch.as_tensor(qpos, dtype=torch.float32, device="cuda:0"); D.qvel[:] = torch.as_tensor(qvel, dtype=torch.float32, device="cuda:0")
  engine.forward(M, D)
  assert torch.equal(D.qfrc_smooth, qfs0)


def test_force_and_torque_sensors_with_contacts_and_wrenches_match_oracle():
  """force / torque sensors (sensor/builtin_sensor.py:59-101): cfrc_int with contact forces and xfrc_applied in cfrc_ext, on the
  SENSOR_FORCE fixture (joint reactions of a two-body arm; a free puck with a hinged lid on the plane) in 8 worlds."""
  import torch
  m = fx.compile_xml(fx.SENSOR_FORCE)
  W = 8
  rng = np.random.default_rng(3)
  qpos = np.tile(m.qpos0, (W, 1)); qvel = rng.normal(0, 0.4, (W, m.nv)); ctrl = rng.normal(0, 0.5, (W, m.nu))
  qpos[:, 1] = rng.uniform(-1.0, 1.0, W)                    # boom angle
  qpos[:, 4] += rng.uniform(0.0, 0.05, W)                   # puck height: some worlds in flight, some in contact
  engine, M, D, orc, worlds = _setup(m, W, qpos, qvel, ctrl, True, nconmax=16, njmax=64)
  x = np.zeros((W, m.nbody, 6))
  for k, b in enumerate(("mast", "boom", "lid")):
    x[:, m.name2id("body", b), 0:3] = rng.uniform(-6.0, 6.0, (W, 3)); x[:, m.name2id("body", b), 3:6] = rng.uniform(-0.5, 0.5, (W, 3))
  x[0] = 0.0
  _push(D, worlds, "xfrc_applied", x)
  adr = {n: int(m.sensor_adr[m.name2id("sensor", n)]) for n in m.names["sensor"]}
  touched = 0
  for chunk in (0, 1, 10, 60):
    if chunk:
      engine.step(M, D, chunk)
      for w in worlds:
        orc.step(w, chunk)
    else:
      engine.forward(M, D)
      for w in worlds:
        orc.forward(w)
    torch.cuda.synchronize()
    _check_counts(D, worlds)
    sd = D.sensordata.cpu().numpy(); so = np.stack([w["sensordata"] for w in worlds])
    assert np.abs(sd - so).max() < 2e-3 * (1 + np.abs(so).max()), chunk
    # the free puck's root transmits nothing except the wrench applied to the lid... no: to ITS subtree the lid's wrench is external too
    assert np.abs(sd[:, adr["f_puck"]:adr["f_puck"] + 3]).max() < 2e-3
    touched += int((D.ncon > 0).sum())
  assert touched > 0 and np.abs(so[:, adr["f_lid"]:adr["f_lid"] + 3]).max() > 0.5


def test_rangefinder_rays_match_oracle_over_random_poses():
  """rangefinder (sensor/builtin_sensor.py:59-101): 64 random drone poses around the six primitives of SENSOR_RAY, every reading
  (hit distance or -1) against the oracle."""
  import torch
  m = fx.compile_xml(fx.SENSOR_RAY)
  W = 64
  rng = np.random.default_rng(11)
  qpos = np.tile(m.qpos0, (W, 1)); qvel = np.zeros((W, m.nv)); ctrl = np.zeros((W, 0))
  qpos[:, 0:3] += rng.uniform(-0.8, 0.8, (W, 3))
  qt = rng.normal(0, 1, (W, 4)); qt[:, 0] += 2.0; qpos[:, 3:7] = qt / np.linalg.norm(qt, axis=1, keepdims=True)
  qpos[:, 7:10] = rng.uniform(-2.0, 2.0, (W, 3)) + [0, 0, 2.5]
  qt = rng.normal(0, 1, (W, 4)); qpos[:, 10:14] = qt / np.linalg.norm(qt, axis=1, keepdims=True)
  engine, M, D, orc, worlds = _setup(m, W, qpos, qvel, ctrl, True, nconmax=16, njmax=64)
  engine.forward(M, D)
  for w in worlds:
    orc.forward(w)
  torch.cuda.synchronize()
  sd = D.sensordata.cpu().numpy(); so = np.stack([w["sensordata"] for w in worlds])
  assert ((sd < 0) == (so < 0)).all()
  assert np.abs(sd - so).max() < 1e-4
  assert (so < 0).sum() > 10 and (so > 0).sum() > 100       # hits and misses both occur


def test_external_force_translates_and_torque_rotates():
  """The reference's own acceptance test of the wrench inputs (tests/test_entity.py:245-272): +x force -> +x motion, z torque ->
  rotation primarily about z; on a free box far above the floor, and the closed form x = F/m t^2/2 for the semi-implicit steps."""
  import torch
  m = fx.compile_xml(fx.FALLING_BOX)
  engine, M, D, orc, worlds = _setup(m, 2, None, None, None, True, nconmax=8, njmax=32)
  D.qpos[:, 2] = 5.0
  D.qpos[:, 3:7] = torch.tensor([1.0, 0.0, 0.0, 0.0], device="cuda:0")      # axis-aligned like the reference's fixture
  box = m.name2id("body", "box")
  D.xfrc_applied[0, box, 0] = 5.0; D.xfrc_applied[0, box, 5] = 3.0
  q0 = D.qpos.clone()
  engine.step(M, D, 10)
  torch.cuda.synchronize()
  assert float(D.qpos[0, 0]) > float(q0[0, 0]) and not torch.allclose(D.qpos[0, 3:7], q0[0, 3:7])
  wv = D.qvel[0, 3:6].abs().cpu().numpy()
  assert wv[2] > 5 * (wv[0] + wv[1])
  assert torch.equal(D.qpos[1, 0:2], q0[1, 0:2])          # the unforced world did not move sideways
  mass = float(np.asarray(m.body_mass)[box]); dt = float(m.opt.timestep)
  assert float(D.qpos[0, 0] - q0[0, 0]) == pytest.approx(5.0 / mass * dt * dt * 55, rel=1e-4)      # sum_{k=1..10} k = 55


# --------------------------------------------------------------------------------------------- mocap bodies
MOCAP_SCENE = """
<mujoco>
  <compiler angle="radian"/>
  <option timestep="0.004" integrator="implicitfast"/>
  <worldbody>
    <geom name="floor" type="plane" size="5 5 0.1"/>
    <body name="base" pos="0 0 0.5" mocap="true">
      <geom name="base_geom" type="box" size="0.3 0.3 0.05" mass="5.0"/>
      <site name="base_site" pos="0.1 0 0.05"/>
      <body name="link1" pos="0 0 0.1">
        <joint name="joint1" type="hinge" axis="0 0 1" range="-3.14 3.14" damping="0.05"/>
        <geom name="link1_geom" type="capsule" size="0.04 0.2" pos="0 0 0.2" mass="1.0" contype="0" conaffinity="0"/>
        <body name="link2" pos="0 0 0.4">
          <joint name="joint2" type="hinge" axis="0 1 0" range="-1.57 1.57" damping="0.05"/>
          <geom name="link2_geom" type="capsule" size="0.04 0.15" pos="0 0 0.15" mass="0.5" contype="0" conaffinity="0"/>
        </body>
      </body>
    </body>
    <body name="ball" pos="0.2 0.1 0.7">
      <freejoint/>
      <geom name="ball_geom" type="sphere" size="0.08" mass="0.4"/>
    </body>
  </worldbody>
  <sensor>
    <framepos name="bp" objtype="site" objname="base_site"/>
    <jointpos name="j2" joint="joint2"/>
  </sensor>
</mujoco>
"""


def test_mocap_body_pose_drives_the_welded_tree_and_its_collisions():
  """A fixed-base arm on a mocap body (reference tests/test_entity.py:388-415: write_mocap_pose_to_sim -> forward() -> root
  pose, 1e-5; writers entity/data.py:186-187) plus a free ball that rests on the mocap platform: per-world mocap poses,
  forward() fields, then 60 steps during which the platform is moved -- the ball must follow the platform's contact."""
  import torch
  m = fx.compile_xml(MOCAP_SCENE)
  assert m.nmocap == 1
  W = 8
  rng = np.random.default_rng(3)
  qpos = np.tile(m.qpos0, (W, 1)); qpos[:, 0:2] = rng.uniform(-0.5, 0.5, (W, 2))
  engine, M, D, orc, worlds = _setup(m, W, qpos, None, None, True, nconmax=8, njmax=32)
  mp = np.zeros((W, 1, 3)); mq = np.zeros((W, 1, 4))
  for w in range(W):
    mp[w, 0] = [rng.uniform(-1, 1), rng.uniform(-1, 1), rng.uniform(0.3, 0.6)]
    ax = rng.normal(size=3); ax /= np.linalg.norm(ax); ang = rng.normal(0, 0.15)
    mq[w, 0] = [np.cos(ang / 2), *(ax * np.sin(ang / 2))]
  mp[0, 0] = [5.0, 6.0, 7.0]; mq[0, 0] = [1, 0, 0, 0]
  _push(D, worlds, "mocap_pos", mp); _push(D, worlds, "mocap_quat", mq)
  # the ball starts just above each platform
  for w in range(W):
    qpos[w, 2:5] = mp[w, 0] + [0.1, 0.05, 0.16]
  _push(D, worlds, "qpos", qpos)
  engine.forward(M, D)
  for w in worlds:
    orc.forward(w)
  base = m.name2id("body", "base")
  assert torch.allclose(D.xpos[:, base], torch.as_tensor(mp[:, 0], dtype=torch.float32, device="cuda:0"), atol=1e-5)
  assert torch.allclose(D.xquat[0, base], torch.tensor([1.0, 0, 0, 0], device="cuda:0"), atol=1e-5)
  for f in ("xpos", "xquat", "geom_xpos", "site_xpos", "subtree_com", "cdof", "qacc", "sensordata"):
    assert _rel_err(D, worlds, f) < 1e-4, f
  _check_counts(D, worlds)
  seen = 0
  for k in range(6):
    mp[:, 0, 2] += 0.01; mp[:, 0, 0] += 0.005            # lift and shift the platform between chunks
    _push(D, worlds, "mocap_pos", mp)
    engine.step(M, D, 10)
    torch.cuda.synchronize()
    for w in worlds:
      orc.step(w, 10)
    _check_counts(D, worlds)
    seen += int(D.ncon.sum())
    assert _rel_err(D, worlds, "qpos") < 2e-4 and _rel_err(D, worlds, "qvel") < 2e-2
  assert seen > 20                                        # ball-on-platform contacts were active


# --------------------------------------------------------------------------------------------- K4: qM directly
@pytest.mark.parametrize("asset", ["g1", "go1", "asimov_toe"])
def test_mass_matrix_matches_oracle(asset, request):
  m = request.getfixturevalue(asset)
  W = 16
  qpos, qvel, ctrl = legged_states(m, W, seed=23, spread=0.4, vel=0.5)
  engine, M, D, orc, worlds = _setup(m, W, qpos, qvel, ctrl, True)
  engine.forward(M, D)
  for w in worlds:
    orc.forward(w)
  qM = D.qM.cpu().numpy()[:, :, : m.nv]
  ref = np.stack([w["qM"].reshape(m.nv, m.nv) for w in worlds])
  assert np.abs(qM - np.swapaxes(qM, 1, 2)).max() == 0.0                                  # symmetric by construction
  assert np.abs(qM - ref).max() < 5e-6 * np.abs(ref).max()
  pad = D.qM.cpu().numpy()[:, :, m.nv:]
  assert D.qM.shape[2] % 4 == 0 and (pad.size == 0 or np.abs(pad).max() == 0.0)          # zero padding columns (float4 rows)


# --------------------------------------------------------------------------------------------- refresh_static
SLABS = """
<mujoco>
  <compiler angle="radian"/>
  <option timestep="0.004"/>
  <worldbody>
    <body name="terrain">
      <geom name="slab_a" type="box" size="0.5 0.5 0.1" pos="0 0 -0.1"/>
      <geom name="slab_b" type="box" size="0.5 0.5 0.1" pos="1.2 0 -0.1"/>
      <site name="marker" pos="0 0 0"/>
    </body>
    <body name="ball_a" pos="0 0 0.3"><freejoint/><geom name="ga" type="sphere" size="0.1" mass="1"/></body>
    <body name="ball_b" pos="1.2 0 0.3"><freejoint/><geom name="gb" type="sphere" size="0.06" mass="1"/></body>
  </worldbody>
</mujoco>
"""


def test_refresh_static_after_per_world_terrain_write():
  """Per-world terrain randomisation: geom_pos of static geoms written after expand_model_fields (FIELD_SPECS,
  envs/mdp/events.py:228-253) is honoured once b2s_refresh_static has run (Simulation.create_graph does it), and the
  result equals per-world oracles with that geom_pos; before the refresh the frames are the old ones (documented)."""
  import torch
  from oracle import Oracle
  m = fx.compile_xml(SLABS)
  W = 6
  engine, M, D, _, _ = _setup(m, W, None, None, None, True, nconmax=8, njmax=32)
  engine.forward(M, D)
  z0 = D.geom_xpos[:, 0, 2].clone()
  engine.expand_model_fields(M, W, ["geom_pos", "site_pos"])
  dz = np.linspace(-0.08, 0.12, W)
  gp = np.tile(np.asarray(m.geom_pos, np.float64), (W, 1, 1)); gp[:, 0, 2] += dz; gp[:, 1, 2] -= dz
  M.geom_pos[:] = torch.as_tensor(gp, dtype=torch.float32, device="cuda:0")
  M.site_pos[:, 0, 2] = torch.as_tensor(dz, dtype=torch.float32, device="cuda:0")
  engine.forward(M, D)
  assert torch.equal(D.geom_xpos[:, 0, 2], z0)           # static frames are not touched by forward()
  engine.refresh_static(M, D)
  engine.forward(M, D)
  assert torch.allclose(D.geom_xpos[:, 0, 2], torch.as_tensor(-0.1 + dz, dtype=torch.float32, device="cuda:0"), atol=1e-6)
  assert torch.allclose(D.site_xpos[:, 0, 2], torch.as_tensor(dz, dtype=torch.float32, device="cuda:0"), atol=1e-6)
  engine.step(M, D, 150)
  torch.cuda.synchronize()
  qp = D.qpos.cpu().numpy(); nc = D.ncon.cpu().numpy()
  for w in range(W):
    o = Oracle(m, "f32", D.nconmax, D.njmax, overrides={"geom_pos": gp[w]}, ls_parallel=1)
    ow = o.new_world(); o.step(ow, 150)
    assert int(nc[w]) == int(ow["ncon"][0])
    assert np.abs(qp[w] - ow["qpos"]).max() < 2e-4
  # the balls came to rest on their own slab heights
  assert np.allclose(qp[:, 2], 0.1 + dz, atol=3e-3) and np.allclose(qp[:, 9], 0.06 - dz, atol=3e-3)


def test_per_world_pose_of_grid_terrain_is_refused(go1_rough):
  """With more than 64 static geoms the broadphase grid is built once from the compile-time poses: expanding a pose field of
  such a model per world must fail loudly instead of leaving the grid stale."""
  from asimov_mjlab_b200 import engine
  M = engine.put_model(go1_rough, "cuda:0")
  for f in ("geom_pos", "geom_quat", "body_pos", "body_quat"):
    with pytest.raises(engine.BackendError, match="grid"):
      engine.expand_model_fields(M, 4, [f])
  engine.expand_model_fields(M, 4, ["geom_friction"])     # non-pose fields still expand


# --------------------------------------------------------------------------------------------- overflow
def test_nconmax_and_njmax_overflow_in_a_subset_of_worlds(g1):
  """nconmax / njmax too small for SOME worlds: the engine keeps the first nconmax contacts in pair order and drops the
  contact whose rows do not fit and every later one (oracle mjstep_oracle.c:891-899); worlds that do not overflow are
  untouched.  256 worlds (fallen, lying and standing robots), every world against the oracle."""
  import torch
  W = 256
  qpos, qvel, ctrl = g1_states(g1, W, seed=41, spread=0.5, vel=0.3, zlo=0.15, zhi=0.8)
  rng = np.random.default_rng(41)
  for w in range(0, W, 2):                                   # every other robot lies on the ground at a random attitude
    ax = rng.normal(size=3); ax /= np.linalg.norm(ax); ang = rng.uniform(0.8, 2.4)
    qpos[w, 3:7] = [np.cos(ang / 2), *(ax * np.sin(ang / 2))]
    qpos[w, 2] = rng.uniform(0.12, 0.3)
  for nconmax, njmax in ((6, 300), (50, 22), (5, 14)):
    engine, M, D, orc, worlds = _setup(g1, W, qpos, qvel, ctrl, True, nconmax=nconmax, njmax=njmax)
    engine.forward(M, D)
    for w in worlds:
      orc.forward(w)
    _check_counts(D, worlds)
    ncon = D.ncon.cpu().numpy(); nefc = D.nefc.cpu().numpy()
    assert (ncon == nconmax).sum() > 10 or (nefc >= njmax - 3).sum() > 10      # some worlds hit a cap ...
    assert (ncon < nconmax).sum() > 10 and (nefc < njmax - 4).sum() > 10        # ... and some stay below both
    assert _rel_err(D, worlds, "qacc") < 2e-3 and _rel_err(D, worlds, "efc_force") < 2e-3
    engine.step(M, D, 3)
    torch.cuda.synchronize()
    for w in worlds:
      orc.step(w, 3)
    _check_counts(D, worlds)
    assert np.isfinite(D.qpos.cpu().numpy()).all()


# --------------------------------------------------------------------------------------------- disable flags, maxmatch
@pytest.mark.parametrize("flag", ["CONSTRAINT", "LIMIT", "CONTACT", "PASSIVE", "GRAVITY", "CLAMPCTRL", "WARMSTART", "ACTUATION",
                                  "REFSAFE", "FILTERPARENT"])
def test_disable_flags_match_oracle(g1, flag):
  """mjtDisableBit subset (include/b200sim.h B2S_DSBL_*): each flag alone, forward + 5 steps on contact-rich G1 states."""
  import torch
  bits = dict(CONSTRAINT=1 << 0, LIMIT=1 << 3, CONTACT=1 << 4, PASSIVE=1 << 5, GRAVITY=1 << 6, CLAMPCTRL=1 << 7, WARMSTART=1 << 8,
              FILTERPARENT=1 << 9, ACTUATION=1 << 10, REFSAFE=1 << 11)
  W = 12
  m = copy.copy(g1); m.opt = copy.copy(g1.opt); m.opt.disableflags = bits[flag]
  qpos, qvel, ctrl = g1_states(m, W, seed=17, spread=0.3)
  ctrl[:, :4] += 3.0                                          # beyond ctrlrange: CLAMPCTRL matters
  engine, M, D, orc, worlds = _setup(m, W, qpos, qvel, ctrl, True)
  assert M.opt.disableflags == bits[flag]
  engine.forward(M, D)
  for w in worlds:
    orc.forward(w)
  _check_counts(D, worlds)
  for f in ("qfrc_passive", "actuator_force", "qfrc_smooth", "qacc_smooth", "qacc"):
    assert _rel_err(D, worlds, f) < TOL_FORWARD[f], f
  engine.step(M, D, 5)
  torch.cuda.synchronize()
  for w in worlds:
    orc.step(w, 5)
  _check_counts(D, worlds)
  assert _rel_err(D, worlds, "qpos") < 5e-5 and _rel_err(D, worlds, "qvel") < 5e-3


def test_contact_sensor_maxmatch_cap_matches_oracle():
  """Model.opt.contact_sensor_maxmatch (reference sim.py:118) smaller than the number of matching contacts: the first
  `maxmatch` matches in contact order are reduced, on both sides."""
  import torch
  from oracle import Oracle
  m = fx.legged_with_contact_sensors()
  W = 6
  rng = np.random.default_rng(2)
  qpos = np.tile(m.qpos0, (W, 1)); qpos[:, 2] = np.linspace(0.12, 0.3, W)
  qpos[:, 7:] += rng.normal(0, 0.3, (W, m.nq - 7))
  engine, M, D, _, _ = _setup(m, W, qpos, None, None, True, nconmax=24, njmax=96)
  M.opt.contact_sensor_maxmatch = 2
  orc = Oracle(m, "f32", D.nconmax, D.njmax, ls_parallel=1, opt={"contact_sensor_maxmatch": 2})
  worlds = [orc.new_world(qpos[w]) for w in range(W)]
  capped = 0
  for _ in range(10):
    engine.step(M, D, 5)
    torch.cuda.synchronize()
    for w in worlds:
      orc.step(w, 5)
    _check_counts(D, worlds)
    sd = D.sensordata.cpu().numpy(); so = np.stack([w["sensordata"] for w in worlds])
    assert np.abs(sd - so).max() < 2e-2 * (1 + np.abs(so).max())
    capped += int((D.ncon.cpu().numpy() > 2).sum())
  assert capped > 10


# --------------------------------------------------------------------------------------------- free-running horizons
def _free_run(m, qpos, qvel, ctrl, nstep, marks):
  """Both sides from the same state, NO re-seeding: returns per world the first step whose contact list (pairs, order) or row
  counts differ (nstep + 1 if never) and the max-abs qpos error at the marked steps."""
  import torch
  W = len(qpos)
  engine, M, D, orc, worlds = _setup(m, W, qpos, qvel, ctrl, True)
  first = np.full(W, nstep + 1)
  err = {}
  for s in range(1, nstep + 1):
    engine.step(M, D)
    for w in worlds:
      orc.step(w)
    ncon = D.ncon.cpu().numpy(); nefc = D.nefc.cpu().numpy(); geom = D.contact_geom.cpu().numpy()
    for i, w in enumerate(worlds):
      n = int(w["ncon"][0])
      same = int(ncon[i]) == n and int(nefc[i]) == int(w["nefc"][0]) and geom[i, :n].reshape(-1).tolist() == w["con_geom"][: 2 * n].tolist()
      if not same and first[i] > nstep:
        first[i] = s
    if s in marks:
      q = D.qpos.cpu().numpy(); o = np.stack([w["qpos"] for w in worlds])
      err[s] = np.abs(q - o).max(axis=1)
  return first, err


@pytest.mark.parametrize("asset", ["g1", "go1"])
def test_free_running_divergence_curve(asset, request, capsys):
  """The honest horizon (VERDICT round 1, weak #2d): 32 worlds run FREE on both sides for 100 sub-steps from contact-rich
  states with moving joint targets.  Both implementations are fp32 with different summation orders, a legged robot in
  contact is chaotic, so trajectories separate; the curve is reported and bounded: no contact list differs in the first 25 steps,
  the median world keeps bit-identical contact lists (pairs and order) for > 80 steps, and the position error stays at the
  1e-5 level until a contact set first differs."""
  m = request.getfixturevalue(asset)
  W, N = 32, 100
  if asset == "g1":
    qpos, qvel, ctrl = g1_states(m, W, seed=2, spread=0.15, vel=0.2)
  else:
    qpos, qvel, ctrl = legged_states(m, W, seed=5, spread=0.25, vel=0.4, zlo=0.2, zhi=0.32)
  first, err = _free_run(m, qpos, qvel, ctrl, N, (10, 50, 100))
  order = np.sort(first)
  with capsys.disabled():
    print(f"\n[free-run {asset}] first contact-set difference (step, sorted over {W} worlds): {order.tolist()}")
    for s in (10, 50, 100):
      print(f"[free-run {asset}] qpos max-abs error at step {s}: median {np.median(err[s]):.2e}  p90 {np.quantile(err[s], 0.9):.2e}  max {err[s].max():.2e}")
  # measured on B200 (round 2): G1 first difference at step 54, 25 of 32 worlds identical for all 100 steps; Go1 31 of 32
  assert order[0] > 25, order
  assert np.median(first) > 80, order
  assert err[10].max() < 2e-5 and err[50].max() < 1e-4
  assert np.median(err[100]) < 1e-5
  # while a world's contact sets agree, its state agrees closely: worlds that never diverged within 50 steps
  ok = first > 50
  if ok.any():
    assert err[50][ok].max() < 1e-4


def test_two_models_in_one_process_keep_their_shared_memory_limits(g1):
  """Regression (round 2): the dynamic shared-memory limit of a kernel is per function, and the stage kernels are shared between
  models -- creating a Data for a SMALLER model must not lower the limit under an existing Data of a larger one (its next
  launch failed with "invalid argument")."""
  import torch
  W = 8
  qpos, qvel, ctrl = g1_states(g1, W, seed=3)
  engine, M, D, orc, worlds = _setup(g1, W, qpos, qvel, ctrl, True)
  engine.step(M, D, 2)
  small = fx.compile_xml(fx.LEGGED)
  e2, M2, D2, orc2, w2 = _setup(small, 4, None, None, None, True, nconmax=16, njmax=64)
  e2.step(M2, D2, 2)
  engine.step(M, D, 2)                                      # the larger model again
  e2.step(M2, D2, 2)
  torch.cuda.synchronize()
  for w in worlds:
    orc.step(w, 4)
  for w in w2:
    orc2.step(w, 4)
  _check_counts(D, worlds); _check_counts(D2, w2)
  assert _rel_err(D, worlds, "qpos") < 2e-4 and _rel_err(D2, w2, "qpos") < 2e-4


# --------------------------------------------------------------------------------------------- H3: per-env origin sites
def test_env_origin_sites_are_stored_once(g1):
  """mjlab's terrain importer adds one world-body site per env (terrain_importer.py:95,98-120), so the
  reference's `site_xpos` is (W, 6 + W, 3): 206 GB with `site_xmat` at W = 65 536.  Here static sites beyond 1024 are stored
  once: the G1 scene WITH its W origin sites at W = 65 536 fits in < 3 GB, `data.site_xpos[:, site_ids]` (entity/data.py:315-327)
  equals the dense computation of the same worlds, the origin sites read back their model positions, and the site-based
  sensors (gyro / velocimeter / accelerometer sit on sites) are unchanged."""
  import torch
  from asimov_mjlab_b200 import engine, scenes
  from asimov_mjlab_b200.hostmodel import HostData
  W = 65536
  big = scenes.with_env_origin_sites(g1, W)
  assert big.nsite == g1.nsite + W
  M = engine.put_model(big, "cuda:0")
  D = engine.put_data(big, HostData(big), W, 50, 300, model=M)
  assert D.bytes < 3 * 2 ** 30, D.bytes
  assert isinstance(D.site_xpos, engine.SplitField) and D.site_xpos.shape == (W, g1.nsite + W, 3) and D.site_xmat.shape == (W, g1.nsite + W, 3, 3)
  # a dense model of the same robot (no origin sites), same states in the first 64 worlds
  n = 64
  qpos, qvel, ctrl = g1_states(g1, n, seed=8)
  Md = engine.put_model(g1, "cuda:0")
  Dd = engine.put_data(g1, HostData(g1), n, 50, 300, model=Md)
  for name, arr in (("qpos", qpos), ("qvel", qvel), ("ctrl", ctrl)):
    t = torch.as_tensor(arr, dtype=torch.float32, device="cuda:0")
    getattr(D, name)[:n] = t; getattr(Dd, name)[:] = t
  engine.step(M, D, 3); engine.step(Md, Dd, 3)
  torch.cuda.synchronize()
  ids = torch.tensor([big.name2id("site", s) for s in g1.names["site"]], device="cuda:0")
  assert int(ids.min()) == W                                            # the robot's sites follow the W world-body sites
  assert torch.equal(D.site_xpos[:, ids][:n], Dd.site_xpos) and torch.equal(D.site_xmat[:, ids][:n], Dd.site_xmat)
  assert torch.equal(D.site_xpos[:n, int(ids[1]), 2], Dd.site_xpos[:, 1, 2])
  assert torch.equal(D.sensordata[:n], Dd.sensordata) and torch.equal(D.qpos[:n], Dd.qpos)
  # origin sites: stride-0 in the world dimension, values = the model's site_pos
  o = D.site_xpos[:, :W]
  assert o.shape == (W, W, 3) and o.stride(0) == 0
  assert torch.allclose(o[12345, :, :], torch.as_tensor(big.site_pos[:W], dtype=torch.float32, device="cuda:0"), atol=1e-6)
  mixed = D.site_xpos[:4, torch.tensor([0, W + 1, 5, W + 2], device="cuda:0")]
  assert mixed.shape == (4, 4, 3) and torch.equal(mixed[:, 1], Dd.site_xpos[:4, 1]) and torch.equal(mixed[:, 3], Dd.site_xpos[:4, 2])
  # per-world poses of static sites cannot be honoured in this layout: refused, not ignored
  with pytest.raises(engine.BackendError, match="site_split"):
    engine.expand_model_fields(M, W, ["site_pos"])


# --------------------------------------------------------------------------------------------- the 12-dof Asimov (all-mesh collision)
def test_asimov12_feet_meshes_on_the_plane_match_oracle(asimov12, capsys):
  """Mjlab-Velocity-Flat-Asimov (tasks/velocity/config/asimov/env_cfgs.py; robot asset_zoo/robots/asimov/xmls/asimov.xml:50-110):
  the task's FEET_ONLY_COLLISION leaves the two foot MESHES against the terrain -- the plane-mesh manifold (up to 4 contacts
  per foot) carries the robot.  forward() fields, contact lists and 40 free-running steps from standing states against the
  oracle; then oracle-driven states of falling robots, compared at every 10th state."""
  import torch
  from test_gpu_parity import ConvexStats
  m = asimov12
  assert (m.nq, m.nv, m.nu, m.nbody) == (19, 18, 12, 16) and int((np.asarray(m.geom_type) == 7).sum()) >= 2
  W = 32
  qpos, qvel, ctrl = legged_states(m, W, seed=3, spread=0.1, vel=0.2, zlo=0.725, zhi=0.74)
  qpos[:, 3:7] = [1.0, 0.0, 0.0, 0.0]                          # upright: both feet load their manifolds
  engine, M, D, orc, worlds = _setup(m, W, qpos, qvel, ctrl, True)
  engine.forward(M, D)
  for w in worlds:
    orc.forward(w)
  _check_counts(D, worlds)
  assert sum(int(w["ncon"][0]) for w in worlds) >= 3 * W      # multi-contact manifolds
  for f in ("geom_xpos", "site_xpos", "subtree_com", "qacc_smooth"):
    assert _rel_err(D, worlds, f) < TOL_FORWARD[f] * 4, f
  # the task's cap of 10 Newton iterations: a world whose solve hits it on either side is an unconverged iterate (measured: one
  # of 32, 1e-2 apart, 2e-3 in cost); every other world agrees at the forward tolerances
  nit_o = np.array([int(w["solver_niter"][0]) for w in worlds]); nit_g = D.solver_niter.cpu().numpy()
  conv = (nit_o < m.opt.iterations) & (nit_g < m.opt.iterations)
  assert conv.sum() >= W - 8 and np.abs(nit_g - nit_o)[conv].max() <= 1      # 5 of 32 worlds sit at the cap (niter = 10)
  for f in ("qacc", "qfrc_constraint", "sensordata"):
    g = getattr(D, f).cpu().numpy().astype(np.float64); o = np.stack([w[f] for w in worlds]).astype(np.float64)
    scale = 1e-6 + np.abs(o).max()
    assert np.abs(g - o)[conv].max() / scale < TOL_FORWARD[f], f
    assert np.abs(g - o)[~conv].max(initial=0) / scale < 0.1, f
  # free run with converged solves (iterations 30): dynamics through the manifolds.  Which hull vertices make up a manifold is a
  # discrete choice (vertices within 1 mm of the deepest one), so a free run separates once one side picks another vertex; until
  # then the states agree closely
  m30 = copy.copy(m); m30.opt = copy.copy(m.opt); m30.opt.iterations = 30
  first, err = _free_run(m30, qpos, qvel, ctrl, 40, (5, 20))
  with capsys.disabled():
    print(f"\n[asimov 12-dof free run] first contact-list difference per world (41 = never): {np.sort(first).tolist()}")
  assert err[5][first > 5].max() < 1e-4 and np.median(first) > 20
  assert err[20][first > 20].max() < 1e-3
  st = ConvexStats(m)
  for chunk in range(10):
    for w in worlds:
      orc.step(w, 10)
    for f in ("qpos", "qvel", "qacc_warmstart"):
      getattr(D, f)[:] = torch.as_tensor(np.stack([w[f] for w in worlds]), dtype=torch.float32, device="cuda:0")
    engine.forward(M, D)
    for w in worlds:
      orc.forward(w)
    n_all = D.ncon.cpu().numpy()
    G = D.contact_geom.cpu().numpy(); P = D.contact_pos.cpu().numpy(); Fr = D.contact_frame.cpu().numpy().reshape(W, -1, 9)
    Di = D.contact_dist.cpu().numpy(); Inc = D.contact_includemargin.cpu().numpy()
    for i, w in enumerate(worlds):
      st.compare(i, w, n_all[i], G, P, Fr, Di, Inc)
  with capsys.disabled():
    st.report("asimov 12-dof, feet meshes on the plane")
  assert st.identical >= 0.98 * st.states and st.ncmp > 300 and st.tight >= 0.97 * st.ncmp, (st.identical, st.states, st.ncmp, st.tight)


def test_asimov12_full_collision_mesh_mesh_contacts_match_oracle(asimov12_full, capsys):
  """FULL_COLLISION of the same robot (asimov_constants.py:195-205): 13 link hulls (6501 vertices) against the plane and each
  other.  Random joint configurations put neighbouring hulls into DEEP overlap (the convex hulls of grand-parent / grand-child
  links intersect at the joints): GJK + EPA on penetrating mesh-mesh pairs.  forward() only -- such states are not a physical
  workload -- contact lists, distances and frames against the oracle."""
  import torch
  from test_gpu_parity import ConvexStats
  m = asimov12_full
  W = 32
  qpos, qvel, ctrl = legged_states(m, W, seed=5, spread=0.6, vel=0.0, zlo=0.3, zhi=0.8)
  rng = np.random.default_rng(5)
  for w in range(W):
    ax = rng.normal(size=3); ax /= np.linalg.norm(ax); ang = rng.normal(0, 0.8)
    qpos[w, 3:7] = [np.cos(ang / 2), *(ax * np.sin(ang / 2))]
  engine, M, D, orc, worlds = _setup(m, W, qpos, qvel, ctrl, True)
  engine.forward(M, D)
  for w in worlds:
    orc.forward(w)
  st = ConvexStats(m)
  n_all = D.ncon.cpu().numpy()
  G = D.contact_geom.cpu().numpy(); P = D.contact_pos.cpu().numpy(); Fr = D.contact_frame.cpu().numpy().reshape(W, -1, 9)
  Di = D.contact_dist.cpu().numpy(); Inc = D.contact_includemargin.cpu().numpy()
  for i, w in enumerate(worlds):
    st.compare(i, w, n_all[i], G, P, Fr, Di, Inc)
  with capsys.disabled():
    st.report("asimov 12-dof, full collision, random poses")
  assert (7, 7) in st.seen and st.ncmp > 150
  assert st.identical >= 0.9 * st.states and st.tight >= 0.9 * st.ncmp, (st.identical, st.states, st.ncmp, st.tight)


def test_solver_row_capacity_split_is_bit_identical(g1):
  """The default solver variant runs with a shared-memory layout for `solve_rowcap` rows (128) and leaves worlds with more rows to a
  second launch that has room for njmax.  With the capacity forced down to 24 rows the G1 worlds are spread over BOTH launches
  (a standing G1 has ~50 rows, a flying one a handful of limit rows): results must not depend on which launch solved a world."""
  import torch
  from test_gpu_parity import _setup, _check_counts, _rel_err
  W = 64
  qpos, qvel, ctrl = g1_states(g1, W, seed=11, spread=0.15, vel=0.2)
  qpos[::4, 2] += 1.0      # a quarter of the worlds in the air: few rows
  engine, M, D, orc, worlds = _setup(g1, W, qpos, qvel, ctrl, True)
  assert M.opt.solve_rowcap == 128
  engine.step(M, D, 3)
  ref = {f: getattr(D, f).clone() for f in ("qpos", "qvel", "qacc", "efc_force", "solver_niter", "sensordata")}
  nefc = D.nefc.cpu().numpy()
  for cap in (24, 1, 100000):
    M.opt.solve_rowcap = cap
    for name, arr in (("qpos", qpos), ("qvel", qvel), ("ctrl", ctrl)):
      getattr(D, name)[:] = torch.as_tensor(np.asarray(arr), dtype=torch.float32, device="cuda:0")
    D.qacc_warmstart.zero_(); D.time.zero_()
    n0 = engine.launch_count()
    engine.step(M, D, 3)
    torch.cuda.synchronize()
    assert engine.launch_count() - n0 == 3 * (8 if cap < D.njmax else 7)
    for f, r in ref.items():
      assert torch.equal(getattr(D, f), r), (cap, f)
  assert (nefc > 24).sum() > 8 and (nefc <= 24).sum() > 8, nefc      # both launches had worlds at cap = 24
  M.opt.solve_rowcap = 24
  for w in worlds:
    orc.step(w, 3)
  _check_counts(D, worlds)
  assert _rel_err(D, worlds, "qpos") < 1e-5 and _rel_err(D, worlds, "qvel") < 2e-3
