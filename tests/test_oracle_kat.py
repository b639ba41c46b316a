"""Known-answer tests that pin the CPU oracle (parity is otherwise unpinned: SURVEY.md §8c lists these)."""
import numpy as np
import pytest
from scipy.optimize import brentq

import fixtures as fx
from asimov_mjlab_b200 import hostmodel
from oracle import Oracle


def test_free_fall_matches_semi_implicit_euler_closed_form():
  m = fx.compile_xml(fx.SPINNER)
  m.opt.gravity = np.array([0.0, 0.0, -9.81])
  orc = Oracle(m, "f64")
  w = orc.new_world()
  n, dt, g = 200, m.opt.timestep, 9.81
  orc.step(w, n)
  assert w["qvel"][2] == pytest.approx(-g * n * dt, rel=1e-12)
  assert w["qpos"][2] == pytest.approx(1.0 - g * dt * dt * n * (n + 1) / 2, rel=1e-12)
  assert w["time"][0] == pytest.approx(n * dt)


def test_quaternion_integration_of_constant_body_rate():
  m = fx.compile_xml(fx.SPINNER)
  orc = Oracle(m, "f64")
  omega = np.array([0.3, -0.5, 0.8])
  w = orc.new_world(qvel=np.concatenate([np.zeros(3), omega]))
  n = 500
  orc.step(w, n)
  ang = np.linalg.norm(omega) * n * m.opt.timestep
  expect = np.concatenate([[np.cos(ang / 2)], omega / np.linalg.norm(omega) * np.sin(ang / 2)])
  assert np.allclose(w["qpos"][3:7], expect, atol=1e-10)
  assert np.allclose(w["qvel"][3:], omega, atol=1e-12)   # sphere inertia: no precession


def test_pendulum_energy_and_period():
  m = fx.compile_xml(fx.PENDULUM)
  orc = Oracle(m, "f64")
  th0 = 0.2
  w = orc.new_world(qpos=[th0])
  orc.forward(w)
  M = w["qM"][0]
  L = 0.5
  # point mass at L plus sphere inertia 2/5 m r^2
  assert M == pytest.approx(1.0 * L * L + 0.4 * 0.05 ** 2, rel=1e-9)
  period = 2 * np.pi * np.sqrt(M / (9.81 * L)) * (1 + th0 ** 2 / 16)
  qs = []
  for _ in range(int(1.5 * period / m.opt.timestep)):
    orc.step(w)
    qs.append(w["qpos"][0])
  qs = np.array(qs)
  # first return to a maximum after t = 0
  k = np.argmax(qs[int(0.5 * period / m.opt.timestep):]) + int(0.5 * period / m.opt.timestep)
  assert k * m.opt.timestep == pytest.approx(period, rel=5e-3)
  assert qs[k] == pytest.approx(th0, rel=2e-2)


def test_pd_actuator_one_implicitfast_step_closed_form():
  m = fx.compile_xml(fx.SLIDER_PD)
  orc = Oracle(m, "f64")
  q0, v0, c = 0.1, -0.3, 0.5
  w = orc.new_world(qpos=[q0], qvel=[v0], ctrl=[c])
  orc.step(w)
  mass, kp, kv, dt = 3.0, 50.0, 4.0, 0.005
  f = kp * (c - q0) - kv * v0
  assert w["actuator_force"][0] == pytest.approx(f)
  assert w["qacc"][0] == pytest.approx(f / mass)
  a_impl = f / (mass + dt * kv)
  assert w["qvel"][0] == pytest.approx(v0 + dt * a_impl, rel=1e-12)
  assert w["qpos"][0] == pytest.approx(q0 + dt * (v0 + dt * a_impl), rel=1e-12)


def test_fixed_tendon_actuator_spring_damper_closed_form():
  """K3/K9: fixed tendon length/velocity, tendon position actuator, tendon spring + damper, and the implicitfast
  rank-1 velocity derivative c c^T (kv + b) -- one step against the closed form."""
  m = fx.compile_xml(fx.TENDON_SLIDERS)
  orc = Oracle(m, "f64")
  q = np.array([0.1, -0.2]); v = np.array([-0.3, 0.5]); u = 0.25
  w = orc.new_world(qpos=q, qvel=v, ctrl=[u])
  orc.step(w)
  c = np.array([0.7, -0.4]); M = np.diag([3.0, 0.5]); kp, kv, ks, kb, l0, dt = 50.0, 4.0, 20.0, 1.5, 0.05, 0.005
  length, vel = c @ q, c @ v
  assert w["ten_length"][0] == pytest.approx(length) and w["ten_velocity"][0] == pytest.approx(vel)
  assert w["actuator_length"][0] == pytest.approx(length) and w["actuator_velocity"][0] == pytest.approx(vel)
  f_act = kp * (u - length) - kv * vel
  f_pas = ks * (l0 - length) - kb * vel
  assert w["actuator_force"][0] == pytest.approx(f_act)
  assert np.allclose(w["qfrc_actuator"], c * f_act) and np.allclose(w["qfrc_passive"], c * f_pas)
  rhs = c * (f_act + f_pas)
  assert np.allclose(w["qacc"], np.linalg.solve(M, rhs))
  a_impl = np.linalg.solve(M + dt * (kv + kb) * np.outer(c, c), rhs)
  assert np.allclose(w["qvel"], v + dt * a_impl, rtol=1e-12, atol=0)
  assert np.allclose(w["qpos"], q + dt * (v + dt * a_impl), rtol=1e-12, atol=0)


def test_tendon_spring_dead_band_and_clamped_actuator():
  m = fx.compile_xml(fx.TENDON_ANKLE)
  assert np.allclose(m.tendon_lengthspring[2], [-0.01, 0.02])
  orc = Oracle(m, "f64")
  # tendon C = 0.05 roll + 0.1 toe: inside the dead band -> spring silent, damper only
  w = orc.forward(orc.new_world(qpos=[0.1, 0.2, 0.0], qvel=[0, 0, 1.0], ctrl=[0, 0, 0]))
  assert w["ten_length"][2] == pytest.approx(0.01)
  assert w["qfrc_passive"][1] == pytest.approx(0.05 * (-0.8 * 0.1))
  # above the band: pulls back towards the upper edge
  w = orc.forward(orc.new_world(qpos=[0.1, 0.4, 0.05], qvel=[0, 0, 0], ctrl=[0, 0, 0]))
  assert w["qfrc_passive"][1] == pytest.approx(0.05 * 40 * (0.02 - 0.025))
  # actuator A: ctrl clamped to 0.08, then force clamped to 12; moment = tendon coefficients
  w = orc.forward(orc.new_world(qpos=[0.3, 0.1, 0.0], qvel=[0, 0, 0], ctrl=[0.5, 0.0, 0.0]))
  la = -0.09 * 0.3 - 0.02 * 0.1
  assert 300 * (0.08 - la) > 12 and w["actuator_force"][0] == pytest.approx(12.0)
  lb = -0.09 * 0.3 + 0.02 * 0.1
  fb = 300 * (0.0 - lb)
  assert w["actuator_force"][1] == pytest.approx(fb)
  assert w["qfrc_actuator"][0] == pytest.approx(-0.09 * 12.0 - 0.09 * fb)
  assert w["qfrc_actuator"][1] == pytest.approx(-0.02 * 12.0 + 0.02 * fb)


def test_two_sphere_contact_geometry():
  m = fx.compile_xml(fx.TWO_SPHERES)
  orc = Oracle(m, "f64")
  w = orc.forward(orc.new_world())
  assert w["ncon"][0] == 1
  assert w["con_dist"][0] == pytest.approx(0.15 - 0.18)
  assert np.allclose(w["con_frame"][:3], [1, 0, 0])
  assert np.allclose(w["con_pos"][:3], [0.1 + (0.15 - 0.18) / 2, 0, 1.0])
  assert list(w["con_geom"][:2]) == [0, 1]
  # frame is right-handed orthonormal
  F = w["con_frame"][:9].reshape(3, 3)
  assert np.allclose(F @ F.T, np.eye(3), atol=1e-12) and np.linalg.det(F) == pytest.approx(1.0)


def test_capsule_on_plane_two_contacts():
  m = fx.compile_xml(fx.CAPSULE_ON_PLANE)
  orc = Oracle(m, "f64")
  w = orc.forward(orc.new_world())
  assert w["ncon"][0] == 2
  assert np.allclose(w["con_dist"][:2], [-0.005, -0.005])
  xs = sorted(w["con_pos"][:6].reshape(2, 3)[:, 0])
  assert np.allclose(xs, [-0.2, 0.2])
  assert np.allclose(w["con_pos"][:6].reshape(2, 3)[:, 2], -0.0025)
  # normal = plane z, tangent aligned with the capsule axis
  F = w["con_frame"][:9].reshape(3, 3)
  assert np.allclose(F[0], [0, 0, 1]) and abs(F[1][0]) == pytest.approx(1.0)
  # condim 3 pyramidal: 4 rows per contact
  assert w["nefc"][0] == 8


def test_sphere_rest_force_and_penetration_closed_form():
  m = fx.compile_xml(fx.SPHERE_DROP)
  orc = Oracle(m, "f64", ls_parallel=0)
  w = orc.new_world()
  w["qpos"][2] = 0.0999
  orc.step(w, 4000)
  assert abs(w["qvel"][2]) < 1e-9
  mass, g = 2.0, 9.81
  assert w["ncon"][0] == 1 and w["nefc"][0] == 1
  assert w["efc_force"][0] == pytest.approx(mass * g, rel=1e-7)
  # equilibrium: m g = D K imp(r) (-r), D = imp / ((1 - imp) * invweight), invweight = 1/m
  d0, dw, width, mid, power = 0.9, 0.95, 0.001, 0.5, 2.0
  tc = max(0.02, 2 * m.opt.timestep)
  K = 1.0 / (dw * dw * tc * tc)

  def imp(r):
    x = min(abs(r) / width, 1.0)
    y = x ** power / mid ** (power - 1) if x <= mid else 1 - (1 - x) ** power / (1 - mid) ** (power - 1)
    return d0 + y * (dw - d0)

  def balance(r):
    i = imp(r)
    return i / (1 - i) * mass * K * i * (-r) - mass * g
  r = brentq(balance, -0.01, -1e-9)
  assert w["con_dist"][0] == pytest.approx(r, rel=1e-6)
  assert m.body_invweight0[2, 0] == pytest.approx(1 / mass)


def test_limit_rows_sign_and_activation():
  m = fx.compile_xml(fx.LIMITED_HINGE)
  orc = Oracle(m, "f64")
  w = orc.forward(orc.new_world(qpos=[-0.35]))
  assert w["nefc"][0] == 1 and w["nl"][0] == 1
  assert w["efc_J"][0] == 1.0 and w["efc_pos"][0] == pytest.approx(0.05 - 0.1 + 0.0, abs=1) and w["efc_pos"][0] == pytest.approx(-0.05)
  w = orc.forward(orc.new_world(qpos=[0.45]))
  assert w["nefc"][0] == 1 and w["efc_J"][0] == -1.0 and w["efc_pos"][0] == pytest.approx(-0.05)
  assert w["efc_force"][0] > 0
  w = orc.forward(orc.new_world(qpos=[0.0]))
  assert w["nefc"][0] == 0


def test_kinematics_and_mass_matrix_match_independent_numpy(g1):
  """Oracle (C, tree recursion) vs hostmodel (numpy, float64) on random G1 poses."""
  from util import g1_states
  orc = Oracle(g1, "f64", 50, 300)
  qpos, _, _ = g1_states(g1, 4, seed=7)
  for q in qpos:
    w = orc.forward(orc.new_world(qpos=q))
    M, k, cdof = hostmodel.host_mass_matrix(g1, q / np.r_[np.ones(3), np.full(4, np.linalg.norm(q[3:7])), np.ones(g1.nq - 7)])
    assert np.allclose(w["xpos"].reshape(-1, 3), k["xpos"], atol=1e-10)
    assert np.allclose(w["xquat"].reshape(-1, 4), k["xquat"], atol=1e-10)
    assert np.allclose(w["subtree_com"].reshape(-1, 3), k["subtree_com"], atol=1e-10)
    assert np.allclose(w["cdof"].reshape(-1, 6), cdof, atol=1e-10)
    assert np.allclose(w["qM"].reshape(g1.nv, g1.nv), M, atol=1e-9)
    L = w["qL"].reshape(g1.nv, g1.nv)
    assert np.allclose(L @ L.T, M, atol=1e-8)


def test_g1_standing_contact_force_balances_weight(g1):
  key = g1.keys["robot/init_state"]
  orc = Oracle(g1, "f64", 50, 300, ls_parallel=0)
  w = orc.new_world(qpos=key["qpos"], ctrl=key["ctrl"])
  orc.step(w, 100)
  sl_l, sl_r = g1.sensor_slice("feet_ground_contact_left_ankle_roll_link_force"), g1.sensor_slice("feet_ground_contact_right_ankle_roll_link_force")
  fz = -(w["sensordata"][sl_l][2] + w["sensordata"][sl_r][2])
  weight = g1.body_subtreemass[2] * 9.81
  assert fz == pytest.approx(weight, rel=0.03)
  assert w["sensordata"][g1.sensor_slice("feet_ground_contact_left_ankle_roll_link_found")][0] == 14
  # accelerometer of a quasi-static robot reads +g along the site z axis
  assert w["sensordata"][g1.sensor_slice("robot/imu_lin_acc")][2] == pytest.approx(9.81, rel=0.05)


def test_solver_satisfies_kkt_for_contacts(g1):
  """At the Newton solution: M(qacc - qacc_smooth) = J^T f, f >= 0, f = -D min(J qacc - aref, 0)."""
  from util import g1_states
  orc = Oracle(g1, "f64", 50, 300, ls_parallel=0, opt=dict(iterations=100, tolerance=1e-12))
  qpos, qvel, ctrl = g1_states(g1, 6, seed=11)
  for i in range(6):
    w = orc.forward(orc.new_world(qpos[i], qvel[i], ctrl[i]))
    n, nv = w["nefc"][0], g1.nv
    if n == 0:
      continue
    J = w["efc_J"][: n * nv].reshape(n, nv); f = w["efc_force"][:n]
    M = w["qM"].reshape(nv, nv)
    assert np.all(f >= 0)
    jar = J @ w["qacc"] - w["efc_aref"][:n]
    assert np.allclose(f, -w["efc_D"][:n] * np.minimum(jar, 0), rtol=1e-9, atol=1e-9)
    res = M @ (w["qacc"] - w["qacc_smooth"]) - J.T @ f
    assert np.abs(res).max() < 1e-5 * (1 + np.abs(J.T @ f).max())


def test_box_and_capsule_rest_on_a_static_box():
  """box-box face manifold (4 clipped vertices) and capsule-box (two end contacts): resting forces equal the weights,
  contact normals point from geom1 to geom2, points lie on the slab's top face."""
  m = fx.compile_xml(fx.BOX_ON_BOX)
  orc = Oracle(m, "f64", ls_parallel=0)
  w = orc.new_world()
  orc.forward(w)
  n = int(w["ncon"][0])
  geoms = w["con_geom"][: 2 * n].reshape(n, 2).tolist()
  slab, crate, rod = (m.names["geom"].index(k) for k in ("slab", "crate_geom", "rod_geom"))
  assert geoms.count([slab, crate]) == 4 and geoms.count([rod, slab]) == 2
  F = w["con_frame"][: 9 * n].reshape(n, 3, 3)
  for i, (g1, g2) in enumerate(geoms):
    assert np.allclose(F[i][0], [0, 0, 1] if g1 == slab else [0, 0, -1], atol=1e-12)
    assert abs(w["con_pos"][3 * i + 2]) < 1e-3 and abs(w["con_dist"][i]) < 1e-3
  # crate corners clipped to themselves (the crate lies inside the slab's face)
  P = w["con_pos"][: 3 * n].reshape(n, 3)
  cr = sorted(map(tuple, np.round(P[[i for i, g in enumerate(geoms) if g == [slab, crate]], :2], 9).tolist()))
  assert np.allclose(cr, sorted([(0.1 + sx * 0.15, -0.05 + sy * 0.1) for sx in (-1, 1) for sy in (-1, 1)]), atol=1e-9)
  orc.step(w, 3000)
  assert np.abs(w["qvel"]).max() < 1e-6
  n = int(w["ncon"][0]); geoms = w["con_geom"][: 2 * n].reshape(n, 2).tolist()
  fz = {"crate": 0.0, "rod": 0.0}
  for i, (g1, g2) in enumerate(geoms):
    adr = int(w["con_efc_address"][i]); assert adr >= 0
    f = w["efc_force"][adr: adr + 4].sum()          # pyramidal condim 3: normal force = sum of the 4 edge forces
    fz["crate" if g2 == crate else "rod"] += f
  assert fz["crate"] == pytest.approx(3.0 * 9.81, rel=1e-6) and fz["rod"] == pytest.approx(1.5 * 9.81, rel=1e-6)


def test_box_box_edge_contact_and_separation():
  """Two boxes crossed at 45 deg about two axes touch edge-to-edge: one contact along the common perpendicular;
  moved apart beyond the margin they produce none."""
  from asimov_mjlab_b200.mjcf import Spec
  xml = """
  <mujoco><compiler angle="radian"/><worldbody>
    <body name="terrain"><geom name="a" type="box" size="0.5 0.1 0.1" euler="0.7853981633974483 0 0"/></body>
    <body name="b" pos="0 0 0.27"><freejoint/><geom name="bg" type="box" size="0.1 0.5 0.1" euler="0 0.7853981633974483 0" mass="1"/></body>
  </worldbody></mujoco>"""
  m = Spec.from_string(xml).compile()
  orc = Oracle(m, "f64")
  w = orc.new_world(); orc.forward(w)
  r = 0.1 * np.sqrt(2.0)
  assert w["ncon"][0] == 1
  assert w["con_dist"][0] == pytest.approx(0.27 - 2 * r, abs=1e-9)
  assert np.allclose(w["con_frame"][:3], [0, 0, 1], atol=1e-9) and np.allclose(w["con_pos"][:3], [0, 0, 0.135], atol=1e-9)
  w = orc.new_world(); w["qpos"][2] = 2 * r + 0.01; orc.forward(w)
  assert w["ncon"][0] == 0


# ---------------------------------------------------------------- K7: convex meshes and height fields
def _quat(ax, ang):
  ax = np.asarray(ax, float); ax /= np.linalg.norm(ax)
  return [np.cos(ang / 2), *(ax * np.sin(ang / 2))]


def _contacts(w, m=None):
  n = int(w["ncon"][0])
  return [dict(dist=w["con_dist"][i], pos=w["con_pos"][3 * i:3 * i + 3], n=w["con_frame"][9 * i:9 * i + 3],
               geom=tuple(w["con_geom"][2 * i:2 * i + 2])) for i in range(n)]


@pytest.mark.parametrize("dtype", ["f64", "f32"])
def test_gjk_epa_agrees_with_analytic_box_colliders(dtype):
  """A mesh whose hull is a box must collide like the box: EPA depth / normal against box-box SAT, sphere-box and
  capsule-box closed forms, on the same poses (mesh floor 'slab' = box 1 x 1 x 0.1)."""
  tol = 1e-7 if dtype == "f64" else 2e-5      # hull vertices are stored rounded to float32
  q_tilt = _quat([1, 0.3, 0], 0.05)
  for movers in (("mesh",), ("box",), ("sphere",), ("capsule",)):
    res = {}
    for ground in ("box", "meshbox"):
      m = fx.compile_xml(fx.convex_scene(ground, movers))
      orc = Oracle(m, dtype)
      q = np.array(m.qpos0); q[0:3] = [0.2, -0.1, {"mesh": 0.055, "box": 0.055, "sphere": 0.09, "capsule": 0.13}[movers[0]]]
      q[3:7] = q_tilt
      res[ground] = _contacts(orc.forward(orc.new_world(q)))
    a, b = res["box"], res["meshbox"]
    assert len(b) == 1 and len(a) >= 1
    deepest = min(a, key=lambda c: c["dist"])
    assert b[0]["dist"] == pytest.approx(deepest["dist"], abs=tol)
    # geom order: box floor / box mover -> the floor is geom 1; mesh floor / box mover -> the mover is geom 1 (lower
    # type first), so that one normal flips
    sign = -1.0 if movers[0] == "box" else 1.0
    assert np.allclose(sign * np.asarray(b[0]["n"]), deepest["n"], atol=1e-6 if dtype == "f64" else 1e-4)
    if movers[0] in ("sphere", "capsule"):
      assert np.allclose(b[0]["pos"], deepest["pos"], atol=tol * 10)


@pytest.mark.parametrize("dtype", ["f64", "f32"])
def test_gjk_distance_inside_margin(dtype):
  """Separated shapes within `margin` give a positive-distance contact (GJK branch, no EPA)."""
  m = fx.compile_xml(fx.convex_scene("meshbox", ("sphere", "mesh"), margin=0.02))
  orc = Oracle(m, dtype)
  q = np.array(m.qpos0); q[0:3] = [0.0, 0.0, 0.11]; q[7:10] = [0.5, 0.0, 0.075]
  cs = _contacts(orc.forward(orc.new_world(q)))
  assert len(cs) == 2
  tol = 1e-7 if dtype == "f64" else 1e-6
  assert cs[0]["dist"] == pytest.approx(0.01, abs=tol) and cs[1]["dist"] == pytest.approx(0.015, abs=tol)   # sphere, cube
  assert np.allclose(cs[0]["n"], [0, 0, -1], atol=1e-6) and np.allclose(cs[0]["pos"], [0, 0, 0.005], atol=1e-6)
  q[2] = 0.125; q[9] = 0.085        # outside the margin
  assert _contacts(orc.forward(orc.new_world(q))) == []


def test_plane_mesh_manifold_of_a_resting_cube():
  m = fx.compile_xml(fx.convex_scene("plane", ("mesh",)))
  orc = Oracle(m, "f64")
  q = np.array(m.qpos0); q[0:3] = [0.3, 0.2, 0.059]
  cs = _contacts(orc.forward(orc.new_world(q)))
  assert len(cs) == 4 and all(c["dist"] == pytest.approx(-0.001, abs=1e-7) for c in cs)
  xy = sorted((round(c["pos"][0] - 0.3, 6), round(c["pos"][1] - 0.2, 6)) for c in cs)
  assert xy == [(-0.1, -0.08), (-0.1, 0.08), (0.1, -0.08), (0.1, 0.08)]
  # tilted by more than the 1 mm band: only the deepest corner remains
  q[3:7] = _quat([1, 0, 0], 0.2)
  cs = _contacts(orc.forward(orc.new_world(q)))
  assert 1 <= len(cs) <= 2


def test_height_field_flat_and_sloped():
  """Flat field == plane at z = 0 for a sphere; a field that rises linearly in x gives the slope's normal."""
  m = fx.compile_xml(fx.convex_scene("hfield", ("sphere",)))
  orc = Oracle(m, "f64")
  q = np.array(m.qpos0); q[0:3] = [0.13, -0.21, 0.095]
  cs = _contacts(orc.forward(orc.new_world(q)))
  assert len(cs) >= 1
  for c in cs:
    assert c["dist"] == pytest.approx(-0.005, abs=1e-9) and np.allclose(c["n"], [0, 0, 1], atol=1e-9)
    assert np.allclose(c["pos"], [0.13, -0.21, -0.0025], atol=1e-9)
  nrow = ncol = 9
  elev = [c / (ncol - 1) for r in range(nrow) for c in range(ncol)]        # 0 .. 1 across x = -2 .. 2, top = 0.4
  m = fx.compile_xml(fx.convex_scene("hfield", ("sphere",), hfield_elev=elev))
  orc = Oracle(m, "f64")
  slope = 0.4 / 4.0
  nrm = np.array([-slope, 0, 1]) / np.hypot(slope, 1)
  x, y = 0.13, -0.21
  surf = 0.2 + slope * x
  q[0:3] = np.array([x, y, surf]) + nrm * 0.097
  cs = _contacts(orc.forward(orc.new_world(q)))
  assert len(cs) >= 1
  for c in cs:
    assert c["dist"] == pytest.approx(-0.003, abs=1e-9) and np.allclose(c["n"], nrm, atol=1e-9)


def test_mesh_cube_settles_on_plane_and_height_field():
  for ground in ("plane", "hfield"):
    m = fx.compile_xml(fx.convex_scene(ground, ("mesh", "sphere")))
    orc = Oracle(m, "f64")
    q = np.array(m.qpos0); q[2] = 0.1; q[3:7] = _quat([1, 1, 0], 0.3); q[7:10] = [0.6, 0.1, 0.15]
    w = orc.new_world(q)
    orc.step(w, 1500)
    assert np.abs(w["qvel"]).max() < 5e-2
    assert w["qpos"][2] == pytest.approx(0.06, abs=4e-3) and w["qpos"][9] == pytest.approx(0.1, abs=2e-3)


def test_cg_solver_converges_to_the_newton_solution(g1):
  """mj_solPrimal's CG variant (M^-1 preconditioner, Polak-Ribiere) minimises the same cost: with the exact line search and
  enough iterations its qacc equals the Newton solver's on contact-rich G1 states."""
  import copy
  from util import g1_states
  qpos, qvel, ctrl = g1_states(g1, 6, seed=7)
  out = {}
  for name, solver, iters in (("newton", 2, 50), ("cg", 1, 300)):
    m = copy.copy(g1); m.opt = copy.copy(g1.opt)
    m.opt.solver, m.opt.iterations, m.opt.tolerance = solver, iters, 1e-12
    orc = Oracle(m, "f64", 50, 300, ls_parallel=0)
    ws = [orc.forward(orc.new_world(qpos[i], qvel[i], ctrl[i])) for i in range(6)]
    out[name] = (np.stack([w["qacc"] for w in ws]), np.array([int(w["solver_niter"][0]) for w in ws]), np.array([int(w["nefc"][0]) for w in ws]))
  assert out["newton"][2].sum() > 50
  assert np.abs(out["cg"][0] - out["newton"][0]).max() < 1e-5 * (1 + np.abs(out["newton"][0]).max())
  assert out["cg"][1].max() > out["newton"][1].max()        # first-order directions need more iterations


def test_scalar_and_frame_sensor_readouts():
  """Tendon / actuator / joint-actuator-force / frame-axis / frame-acceleration / clock sensors against their definitions
  (engine_sensor.c): tendon length = sum coef q, actuator length = gear q, joint actuator force = moment^T force, frame
  axes = columns of the frame, accelerometer = R^T framelinacc (gravity included), a pendulum held still reads +g."""
  m = fx.compile_xml(fx.SENSOR_ARM)
  orc = Oracle(m, "f64", 8, 32)
  q, v = np.array([0.3, -0.7]), np.array([0.8, -1.1])
  w = orc.new_world(q, v, np.array([0.2, -0.1]))
  orc.forward(w)
  sd = {n: w["sensordata"][a:a + d] for n, a, d in zip(m.names["sensor"], m.sensor_adr, m.sensor_dim)}
  assert np.isclose(sd["tp"][0], 0.5 * q[0] - 1.5 * q[1]) and np.isclose(sd["tv"][0], 0.5 * v[0] - 1.5 * v[1])
  assert np.isclose(sd["ap"][0], 2.0 * q[0]) and np.isclose(sd["av"][0], 2.0 * v[0])
  f_sh, f_t = w["actuator_force"]
  assert np.isclose(sd["af"][0], f_t) and np.isclose(sd["jf"][0], -1.5 * f_t)        # only the tendon actuator reaches the elbow
  assert np.isclose(f_sh, 30 * (0.2 - 2.0 * q[0]) - 2 * 2.0 * v[0])
  smat = w["site_xmat"].reshape(-1, 3, 3); names = m.names["site"]
  mid, tip = smat[names.index("mid")], smat[names.index("tip")]
  assert np.allclose(sd["fx"], mid[:, 0]) and np.allclose(sd["fy"], mid.T @ tip[:, 1])
  assert np.allclose(sd["fz"], w["xmat"].reshape(-1, 3, 3)[m.names["body"].index("fore")][:, 2])
  assert np.allclose(sd["acc"], tip.T @ sd["la"], atol=1e-10)
  assert sd["clk"][0] == 0.0
  orc.step(w); orc.forward(w)
  assert np.isclose(w["sensordata"][m.sensor_adr[m.names["sensor"].index("clk")]], 0.005)
  # held still by a stiff actuator-free setup: zero velocity and zero acceleration -> the frame "accelerates" at +g
  w2 = orc.new_world(np.zeros(2), np.zeros(2), np.zeros(2))
  w2["qacc_warmstart"][:] = 0
  orc.forward(w2)
  # not in equilibrium (gravity swings the arm), so only check the angular part against qacc: alpha = axis * qacc summed
  sd2 = {n: w2["sensordata"][a:a + d] for n, a, d in zip(m.names["sensor"], m.sensor_adr, m.sensor_dim)}
  qa = w2["qacc"]
  assert np.allclose(sd2["aa"], np.array([0.0, 1.0, 0.0]) * qa[0] + np.array([0.0, 0.0, 1.0]) * qa[1], atol=1e-9)


def test_joint_friction_loss_closed_forms():
  """Dry joint friction (mj_instantiateFriction + the Huber-shaped cost of mj_constraintUpdate).  A 1 kg slider with
  frictionloss f = 2 N and default solreffriction / solimpfriction: R = (1 - 0.9) / 0.9 * invweight0 = 1/9, D = 9.
  Pushed with F < f it sits in the quadratic zone, qacc = F / (m + D); pushed with F > f the force saturates at -f and
  qacc = (F - f) / m; moving and unforced it decelerates with f / m (aref = -B v pushes jar past the zone)."""
  m = fx.compile_xml(fx.SLIDER_FRICTION)
  orc = Oracle(m, "f64", 4, 16)
  def solve(F, v=0.0):
    w = orc.new_world(None, np.array([v, 0, 0, 0.0]), None)
    w["qfrc_applied"][0] = F
    orc.forward(w)
    return w
  w = solve(1.0)
  assert int(w["nf"][0]) == 4 and int(w["nl"][0]) == 4 and int(w["nefc"][0]) == 4
  assert np.isclose(w["efc_D"][0], 9.0, rtol=1e-9) and np.isclose(w["efc_frictionloss"][0], 2.0)
  assert np.isclose(w["qacc"][0], 1.0 / (1.0 + 9.0), rtol=1e-8) and np.isclose(w["efc_force"][0], -0.9, rtol=1e-8)
  w = solve(5.0)
  assert np.isclose(w["qacc"][0], 3.0, rtol=1e-8) and np.isclose(w["efc_force"][0], -2.0)
  w = solve(-5.0)
  assert np.isclose(w["qacc"][0], -3.0, rtol=1e-8) and np.isclose(w["efc_force"][0], 2.0)
  w = solve(0.0, v=1.5)
  assert np.isclose(w["qacc"][0], -2.0, rtol=1e-8)
  # a resting arm under gravity: the elbow's 0.1 N m of friction cannot hold the forearm's gravity torque (0.4 * 9.81 * 0.15)
  assert abs(w["efc_force"][3]) == pytest.approx(0.1) and abs(w["qacc"][3]) > 1.0
  # disabling the feature removes the rows
  import copy
  m2 = copy.copy(m); m2.opt = copy.copy(m.opt); m2.opt.disableflags = 1 << 2
  w2 = Oracle(m2, "f64", 4, 16).forward(Oracle(m2, "f64", 4, 16).new_world())
  assert int(w2["nf"][0]) == 0 and int(w2["nefc"][0]) == 0
  # a sliding block stops and stays: 1.5 m/s decelerating at 2 m/s^2 is at rest after 0.75 s, within a few mm of v0^2 / (2 a)
  w = orc.new_world(None, np.array([1.5, 0, 0, 0.0]), None)
  orc.step(w, 400)
  assert abs(w["qvel"][0]) < 1e-6 and abs(w["qpos"][0] - 1.5 ** 2 / 4.0) < 5e-3


def _round_world(orc, m, cyl=None, ell=None, ball=None):
  q = np.array(m.qpos0, float).copy()
  for i, v in enumerate((cyl, ell, ball)):
    if v is not None:
      q[7 * i:7 * i + len(v)] = v
  w = orc.new_world(q)
  orc.stage("kinematics", w); orc.stage("com_pos", w); orc.stage("collision", w)
  return w


def test_cylinder_and_ellipsoid_contact_geometry():
  """Closed forms for the round shapes: a cylinder standing on the plane touches at three rim points 120 degrees apart,
  lying on its side at one rim point per cap; an ellipsoid touches at the support point of -n; a sphere pressed into the
  cylinder's wall (GJK on the cores) gets dist = gap and the radial normal."""
  m = fx.compile_xml(fx.ROUND_SHAPES)
  orc = Oracle(m, "f64", 16, 64)
  gi = {n: i for i, n in enumerate(m.names["geom"])}
  def contacts(w, g):
    n = int(w["ncon"][0])
    return [(w["con_dist"][c], w["con_pos"][3 * c:3 * c + 3].copy(), w["con_frame"][9 * c:9 * c + 3].copy())
            for c in range(n) if g in (w["con_geom"][2 * c], w["con_geom"][2 * c + 1])]
  far = [5, 5, 3.0]
  # standing cylinder, 1 mm deep
  w = _round_world(orc, m, cyl=[0, 0, 0.149], ell=far, ball=[-5, 5, 3.0])
  cs = contacts(w, gi["cyl_geom"])
  assert len(cs) == 3 and all(abs(c[0] + 0.001) < 1e-12 and np.allclose(c[2], [0, 0, 1]) for c in cs)
  assert all(abs(np.hypot(c[1][0], c[1][1]) - 0.1) < 1e-12 and abs(c[1][2] + 0.0005) < 1e-12 for c in cs)
  ang = sorted(np.degrees(np.arctan2(c[1][1], c[1][0])) % 360 for c in cs)
  assert np.allclose(np.diff(ang), 120.0)
  # lying on its side (axis along x): one contact under each cap
  w = _round_world(orc, m, cyl=[0, 0, 0.099, np.sqrt(0.5), 0, np.sqrt(0.5), 0], ell=far, ball=[-5, 5, 3.0])
  cs = contacts(w, gi["cyl_geom"])
  assert len(cs) == 2 and all(abs(c[0] + 0.001) < 1e-9 for c in cs)
  assert sorted(round(c[1][0], 6) for c in cs) == [-0.15, 0.15]
  # ellipsoid resting on its long axis, 2 mm deep; then tilted: the contact is where the surface normal is -n
  w = _round_world(orc, m, cyl=far, ell=[0, 1, 0.198], ball=[-5, 5, 3.0])
  (dist, pos, nrm), = contacts(w, gi["ell_geom"])
  assert abs(dist + 0.002) < 1e-12 and np.allclose(pos, [0, 1, -0.001]) and np.allclose(nrm, [0, 0, 1])
  th = 0.6
  w = _round_world(orc, m, cyl=far, ell=[0, 1, 0.3, np.cos(th / 2), np.sin(th / 2), 0, 0], ball=[-5, 5, 3.0])
  R = w["geom_xmat"].reshape(-1, 3, 3)[gi["ell_geom"]]
  S = np.diag([0.1, 0.15, 0.2]); dl = R.T @ np.array([0, 0, -1.0])
  sup = R @ (S @ S @ dl / np.linalg.norm(S @ dl)) + np.array([0, 1, 0.3])
  cs = contacts(w, gi["ell_geom"])
  assert (len(cs) == 1 and abs(cs[0][0] - sup[2]) < 1e-12) if sup[2] < 0 else len(cs) == 0
  assert sup[2] > 0 or np.allclose(cs[0][1][:2], sup[:2])
  # sphere pushed 3 mm into the cylinder's wall at mid height: distance between the cores minus the radius
  w = _round_world(orc, m, cyl=[0, 0, 1.0], ell=far, ball=[0.1 + 0.05 - 0.003, 0, 1.0])
  (dist, pos, nrm), = contacts(w, gi["ball_geom"])
  assert abs(dist + 0.003) < 1e-6 and np.allclose(nrm, [-1, 0, 0], atol=1e-5) and np.allclose(pos, [0.0985, 0, 1.0], atol=1e-5)
  # ... and onto its cap from above
  w = _round_world(orc, m, cyl=[0, 0, 1.0], ell=far, ball=[0.03, 0.02, 1.15 + 0.05 - 0.002])
  (dist, pos, nrm), = contacts(w, gi["ball_geom"])
  assert abs(dist + 0.002) < 1e-6 and np.allclose(nrm, [0, 0, -1], atol=1e-5)


def test_round_shapes_settle_with_their_weight():
  """Dropped from 2 cm, the cylinder (on its cap), the ellipsoid and the sphere come to rest on the plane: contact forces
  carry the weight, resting heights are the shape's extent along z minus a sub-millimetre penetration."""
  m = fx.compile_xml(fx.ROUND_SHAPES)
  orc = Oracle(m, "f64", 16, 64)
  w = _round_world(orc, m, cyl=[0, 0, 0.17], ell=[0, 1, 0.22], ball=[0, 2, 0.07])
  orc.step(w, 600)
  assert np.abs(w["qvel"]).max() < 1e-3
  assert abs(w["qpos"][2] - 0.15) < 1e-3 and abs(w["qpos"][9] - 0.2) < 1e-3 and abs(w["qpos"][16] - 0.05) < 1e-3
  n = int(w["ncon"][0])
  fz = {}
  for c in range(n):
    adr, dim = int(w["con_efc_address"][c]), int(w["con_dim"][c])
    f = sum(w["efc_force"][adr:adr + 2 * (dim - 1)]) if dim > 1 else w["efc_force"][adr]
    fz[int(w["con_geom"][2 * c + 1])] = fz.get(int(w["con_geom"][2 * c + 1]), 0.0) + f
  gi = {nm: i for i, nm in enumerate(m.names["geom"])}
  assert abs(fz[gi["cyl_geom"]] - 1.0 * 9.81) < 0.05 and abs(fz[gi["ell_geom"]] - 2.0 * 9.81) < 0.05 and abs(fz[gi["ball_geom"]] - 0.2 * 9.81) < 0.01


def _cone_contacts(w, n):
  """(first row, dim) of every elliptic contact among the n rows"""
  out, r = [], 0
  while r < n:
    if int(w["efc_type"][r]) == 6:
      dim = int(w["con_dim"][int(w["efc_id"][r])]); out.append((r, dim)); r += dim
    else:
      r += 1
  return out


def test_elliptic_cone_solution_properties(g1):
  """cone="elliptic" (reference sim.py:51): three rows per condim-3 contact (normal, two tangents) instead of four pyramid
  edges; at the Newton solution the stationarity condition M (qacc - qacc_smooth) = J^T f holds, every contact force lies
  in the friction cone with the UNREGULARISED coefficient (|f_t| <= mu f_n), sliding contacts sit exactly on it, and a
  resting box carries its weight on the normal rows alone.  Both line searches reach the same solution."""
  import copy
  from util import g1_states
  mg = copy.copy(g1); mg.opt = copy.copy(g1.opt); mg.opt.cone = 1
  qpos, qvel, ctrl = g1_states(g1, 6, seed=11)
  sliding = sticking = 0
  sols = {}
  for lsp in (0, 1):
    orc = Oracle(mg, "f64", 50, 300, ls_parallel=lsp, opt=dict(iterations=100, tolerance=1e-12, ls_iterations=50))
    for i in range(6):
      w = orc.forward(orc.new_world(qpos[i], qvel[i], ctrl[i]))
      n, nv = int(w["nefc"][0]), g1.nv
      sols[(lsp, i)] = w["qacc"].copy()
      if n == 0 or lsp == 1:
        continue
      J = w["efc_J"][: n * nv].reshape(n, nv); f = w["efc_force"][:n]; M = w["qM"].reshape(nv, nv)
      res = M @ (w["qacc"] - w["qacc_smooth"]) - J.T @ f
      assert np.abs(res).max() < 1e-5 * (1 + np.abs(J.T @ f).max())
      for r, dim in _cone_contacts(w, n):
        assert dim == 3
        mu = w["con_friction"][5 * int(w["efc_id"][r])]
        fn, ft = f[r], np.hypot(f[r + 1], f[r + 2])
        assert fn >= 0 and ft <= mu * fn * (1 + 1e-9) + 1e-12
        if fn > 1e-9:
          sliding += int(abs(ft - mu * fn) < 1e-9 * (1 + fn)); sticking += int(ft < mu * fn * (1 - 1e-6))
  assert sliding > 3 and sticking > 3
  for i in range(6):     # parallel search: discrete steps, converges to the same point more loosely
    assert np.abs(sols[(0, i)] - sols[(1, i)]).max() < 2e-3 * (1 + np.abs(sols[(0, i)]).max())
  # resting box: 4 contacts x 3 rows, friction rows idle, weight on the normals; pushed below / above mu m g it sticks / slides
  m = fx.compile_xml(fx.FALLING_BOX)
  me = copy.copy(m); me.opt = copy.copy(m.opt); me.opt.cone = 1; me.opt.iterations = 50
  orc = Oracle(me, "f64", 16, 64, ls_parallel=0)
  w = orc.new_world(); orc.step(w, 500)
  n = int(w["nefc"][0])
  assert int(w["ncon"][0]) == 4 and n == 12
  f = w["efc_force"][:n].reshape(4, 3)
  assert abs(f[:, 0].sum() - 9.81) < 1e-6 and np.abs(f[:, 1:]).max() < 1e-6
  def slide(F):
    w2 = orc.new_world(w["qpos"].copy(), np.zeros(m.nv)); w2["xfrc_applied"][6 * 2] = F
    orc.step(w2, 150); v1 = w2["qvel"][0]; orc.step(w2, 150)
    return (w2["qvel"][0] - v1) / (150 * 0.002), w2
  a, w2 = slide(5.0)
  assert abs(a) < 1e-3 and abs(w2["qvel"][0]) < 0.01
  a, w2 = slide(12.0)
  assert abs(a - (12.0 - 9.81)) < 0.1 * (12.0 - 9.81) + 0.1


def test_misc_scalar_sensors_closed_forms():
  """magnetometer / joint-limit / tendon-actuator-force / energy sensors (sensor/builtin_sensor.py:59-101) on SENSOR_MISC."""
  import fixtures as fx
  from oracle import Oracle
  m = fx.compile_xml(fx.SENSOR_MISC)
  assert m.nsensordata == 3 + 5 + 1 + 2
  o = Oracle(m, "f64", 4, 16)
  adr = {n: int(m.sensor_adr[m.name2id("sensor", n)]) for n in m.names["sensor"]}
  # state: slider 0.25 (inside its range), hinge 0.7 (0.1 beyond its upper limit), some velocities, controls
  w = o.new_world([0.25, 0.7], [0.3, 0.8], [0.05, 1.5, 0.0])
  o.forward(w)
  sd = w["sensordata"]
  # magnetometer: B in the site frame; the site frame = hinge rotation (about y by 0.7) times the site's own 45 deg about y
  ang = 0.7 + np.pi / 4
  R = np.array([[np.cos(ang), 0, np.sin(ang)], [0, 1, 0], [-np.sin(ang), 0, np.cos(ang)]])
  assert np.allclose(sd[adr["mag"]:adr["mag"] + 3], R.T @ np.array([0.2, -0.5, 0.1]), atol=1e-9)
  # hinge upper limit active: pos - margin = hi - q = -0.1, velocity of the row = -qvel, force >= 0; the slider's limit is inactive
  assert sd[adr["lp"]] == pytest.approx(0.6 - 0.7) and sd[adr["lv"]] == pytest.approx(-0.8) and sd[adr["lf"]] > 0.0
  assert sd[adr["lps"]] == 0.0 and sd[adr["lfs"]] == 0.0
  assert sd[adr["lf"]] == pytest.approx(float(w["efc_force"][0]))
  # tendon actuators: position servo + motor (gear 2 enters the moment, not the actuator force)
  L = 1.0 * 0.25 + 0.3 * 0.7; V = 1.0 * 0.3 + 0.3 * 0.8
  f1 = 20.0 * (0.05 - L) - 1.0 * V; f2 = 1.5
  assert sd[adr["taf"]] == pytest.approx(f1 + f2, rel=1e-9)
  # energies
  z_cart = 1.0 + 0.25
  xi_arm = np.array([0.15 * np.cos(0.7), 0.0, z_cart + 0.1 - 0.15 * np.sin(0.7)])     # capsule COM 0.15 along the rotated x axis
  ep = 9.81 * (2.0 * z_cart + 0.7 * xi_arm[2]) + 0.5 * 40.0 * (0.25 - 0.1) ** 2 + 0.5 * 3.0 * 0.7 ** 2 + 0.5 * 15.0 * (L - 0.05) ** 2
  assert sd[adr["ep"]] == pytest.approx(ep, rel=1e-6)
  qv = np.array([0.3, 0.8]); M = w["qM"].reshape(2, 2)
  assert sd[adr["ek"]] == pytest.approx(0.5 * qv @ M @ qv, rel=1e-9) and sd[adr["ek"]] > 0.0


def test_force_and_torque_sensors_closed_forms():
  """force / torque sensors (sensor/builtin_sensor.py:59-101) on SENSOR_FORCE: the interaction force between a body and its
  parent, cfrc_int of mj_rnePostConstraint, at the site and in its frame."""
  import fixtures as fx
  from oracle import Oracle
  m = fx.compile_xml(fx.SENSOR_FORCE)
  assert m.nsensordata == 24
  o = Oracle(m, "f64", 16, 64)
  adr = {n: int(m.sensor_adr[m.name2id("sensor", n)]) for n in m.names["sensor"]}
  g = 9.81
  # (1) boom horizontal and held by the motor torque m g l/2: nothing accelerates; the slider carries no x force
  q0 = np.array(m.qpos0, float)
  w = o.new_world(q0, np.zeros(m.nv), [-0.7 * g * 0.2, 0.0])
  o.forward(w)
  sd = w["sensordata"]
  assert np.abs(w["qacc"][:2]).max() < 1e-9
  # boom: force = its weight (up), torque about the hinge = -(site - com) x F = (0, -m g l/2, 0), which is what the motor supplies
  assert np.allclose(sd[adr["f_boom"]:adr["f_boom"] + 3], [0, 0, 0.7 * g], atol=1e-9)
  assert np.allclose(sd[adr["t_boom"]:adr["t_boom"] + 3], [0, -0.7 * g * 0.2, 0], atol=1e-9)
  # mast: carries both bodies; the site is rotated 45 deg about y and offset from the COM
  R = np.array([[np.cos(np.pi / 4), 0, np.sin(np.pi / 4)], [0, 1, 0], [-np.sin(np.pi / 4), 0, np.cos(np.pi / 4)]])
  F = np.array([0, 0, 2.7 * g])
  assert np.allclose(sd[adr["f_mast"]:adr["f_mast"] + 3], R.T @ F, atol=1e-9)
  site = np.array([0.05, 0.02, 0.97]); com = (2.0 * np.array([0, 0, 1.0]) + 0.7 * np.array([0.2, 0, 1.1])) / 2.7
  assert np.allclose(sd[adr["t_mast"]:adr["t_mast"] + 3], R.T @ (-np.cross(site - com, F)), atol=1e-9)
  # (2) an external force on the mast: the slider accelerates along x (no x reaction), the z part is taken off the joint load
  w = o.new_world(q0, np.zeros(m.nv), [-0.7 * g * 0.2, 0.0])
  xf = np.zeros((m.nbody, 6)); xf[m.name2id("body", "mast"), :3] = [3.0, 0.0, 4.0]
  w["xfrc_applied"][:] = xf.reshape(-1)
  o.forward(w)
  Fm = R @ w["sensordata"][adr["f_mast"]:adr["f_mast"] + 3]
  # (the site's quaternion has 7 digits in the XML: its matrix differs from the exact 45 degrees by 3e-8)
  assert abs(Fm[2] - (2.7 * g - 4.0)) < 1e-5 and abs(Fm[1]) < 1e-9
  # x: the slider transmits nothing along its axis: the subtree's inertial force sum(m a_x) equals the applied 3 N
  assert abs(Fm[0]) < 1e-5 and abs(w["qacc"][0] - 3.0 / 2.7) < 1e-12
  # (3) the puck is free: its root joint transmits nothing, whatever the contacts do (equation of motion) -- in flight and on the ground
  w = o.new_world(q0, np.zeros(m.nv), [0.0, 0.3])
  for n in (1, 60, 400):
    o.step(w, n)
    sd = w["sensordata"]
    assert np.abs(sd[adr["f_puck"]:adr["f_puck"] + 3]).max() < 1e-6 and np.abs(sd[adr["t_puck"]:adr["t_puck"] + 3]).max() < 1e-6
  assert w["ncon"][0] >= 4
  # the lid at rest on its servo / limit: the hinge carries its weight (site frame rotated 30 deg about x)
  c, s_ = np.cos(np.pi / 6), np.sin(np.pi / 6)
  Rl = np.array([[1, 0, 0], [0, c, -s_], [0, s_, c]])
  xq = w["xquat"].reshape(-1, 4)[m.name2id("body", "lid")]
  from asimov_mjlab_b200.mjcf import quat2mat as quat_to_mat
  Fl = quat_to_mat(xq) @ Rl @ sd[adr["f_lid"]:adr["f_lid"] + 3]
  assert np.allclose(Fl, [0, 0, 0.4 * g], atol=2e-3)
  # cutoff clamps the torque reading
  assert np.abs(sd[adr["t_lid"]:adr["t_lid"] + 3]).max() <= 0.5 + 1e-12


def test_rangefinder_closed_forms():
  """rangefinder (sensor/builtin_sensor.py:59-101) on SENSOR_RAY: one ray per primitive kind with a known distance, the
  site's own body excluded, -1 without a hit, the cutoff caps hits only."""
  import fixtures as fx
  from oracle import Oracle
  m = fx.compile_xml(fx.SENSOR_RAY)
  o = Oracle(m, "f64", 8, 32)
  adr = {n: int(m.sensor_adr[m.name2id("sensor", n)]) for n in m.names["sensor"]}
  q = np.array(m.qpos0, float)
  q[7:10] = [1.5, 0.0, 9.0]                  # the rod far above: out of every ray
  w = o.new_world(q, np.zeros(m.nv), [])
  o.forward(w)
  sd = w["sensordata"]
  want = {"r_down": 2.0, "r_px": 3.0 - 0.5, "r_nx": 2.0 - 0.3, "r_py": 3.0 - 0.2, "r_ny": 2.5 - 0.4, "r_up": 2.0 - 0.3, "r_down_cut": 1.0}
  for n, v in want.items():
    assert sd[adr[n]] == pytest.approx(v, abs=1e-6), n
  # the rod across the +x ray: its surface is 0.05 in front of its axis
  q[7:10] = [1.5, 0.0, 2.0]
  w = o.new_world(q, np.zeros(m.nv), []); o.forward(w)
  assert w["sensordata"][adr["r_px"]] == pytest.approx(1.5 - 0.05, abs=1e-6)
  # drone moved and yawed by 90 deg about z: +x site now looks along +y (the post), +y site along -x (the crate); below the floor's
  # back side nothing is seen; above the egg the up ray misses
  q[0:3] = [0.0, 0.0, 2.0]; q[3:7] = [np.cos(np.pi / 4), 0, 0, np.sin(np.pi / 4)]; q[7:10] = [1.5, 0.0, 9.0]
  w = o.new_world(q, np.zeros(m.nv), []); o.forward(w)
  sd = w["sensordata"]
  assert sd[adr["r_px"]] == pytest.approx(2.8, abs=1e-6) and sd[adr["r_py"]] == pytest.approx(1.7, abs=1e-6)
  q[0:3] = [0.0, 0.0, 5.0]; q[3:7] = [1, 0, 0, 0]
  w = o.new_world(q, np.zeros(m.nv), []); o.forward(w)
  sd = w["sensordata"]
  assert sd[adr["r_up"]] == -1.0 and sd[adr["r_down"]] == pytest.approx(5.0 - 4.3, abs=1e-6) and sd[adr["r_px"]] == -1.0
  # tilted ray onto the floor: distance = height / cos(tilt)
  th = 0.3
  q[0:3] = [1.0, 1.0, 1.5]; q[3:7] = [np.cos(th / 2), np.sin(th / 2), 0, 0]
  w = o.new_world(q, np.zeros(m.nv), []); o.forward(w)
  assert w["sensordata"][adr["r_down"]] == pytest.approx(1.5 / np.cos(th), rel=1e-9)
  assert w["sensordata"][adr["r_down_cut"]] == 1.0
