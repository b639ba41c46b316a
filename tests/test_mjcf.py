"""Model on-ramp: MJCF subset compiler and the mjlab scene assembly (host logic, CPU)."""
import numpy as np
import pytest

import fixtures as fx
from asimov_mjlab_b200 import hostmodel, scenes
from asimov_mjlab_b200.mjcf import Spec


def test_g1_flat_dimensions_match_survey_table(g1):
  # SURVEY.md §8.0 (hand-derived from the MJCF and cfg code)
  assert (g1.nq, g1.nv, g1.nu, g1.na) == (36, 35, 29, 0)
  assert g1.nbody == 32 and g1.ngeom == 69 and g1.nsite == 6 and g1.nsensordata == 21
  col = (g1.geom_contype != 0) | (g1.geom_conaffinity != 0)
  assert int(col.sum()) == 34                      # 33 robot collision geoms + the plane
  types = g1.geom_type[col]
  assert int((types == 3).sum()) == 31 and int((types == 2).sum()) == 2 and int((types == 0).sum()) == 1
  assert g1.names["body"][:3] == ["world", "terrain", "robot/pelvis"]
  assert g1.opt.timestep == 0.005 and g1.opt.integrator == 3 and g1.opt.iterations == 10 and g1.opt.ls_iterations == 20
  assert g1.nexclude == 4 and int(g1.jnt_limited.sum()) == 29


def test_g1_collision_and_actuator_edits(g1):
  names = g1.names["geom"]
  foot = [i for i, n in enumerate(names) if "foot" in n and n.endswith("_collision")]
  assert len(foot) == 14
  assert set(g1.geom_condim[foot]) == {3} and set(g1.geom_priority[foot]) == {1}
  assert np.allclose(g1.geom_friction[foot, 0], 0.6) and np.allclose(g1.geom_friction[foot, 1], 0.005)
  other = [i for i, n in enumerate(names) if n.endswith("_collision") and "foot" not in n]
  assert set(g1.geom_condim[other]) == {1}
  vis = [i for i, n in enumerate(names) if not n.endswith("_collision") and n != "terrain"]
  assert not g1.geom_contype[vis].any() and not g1.geom_conaffinity[vis].any()
  # position actuators: kp on ctrl, affine bias (0, -kp, -kd), ctrlrange inherited from the joint range
  kp = g1.actuator_gainprm[:, 0]
  assert np.allclose(g1.actuator_biasprm[:, 1], -kp) and (g1.actuator_biasprm[:, 2] < 0).all()
  j = g1.actuator_trnid[:, 0]
  assert np.allclose(g1.actuator_ctrlrange, g1.jnt_range[j]) and g1.actuator_ctrllimited.all() and g1.actuator_forcelimited.all()
  assert (g1.dof_armature[6:] > 0).all() and np.allclose(g1.dof_armature[:6], 0)
  hip = g1.names["actuator"].index("robot/left_hip_pitch_joint")
  assert g1.actuator_forcerange[hip].tolist() == [-88.0, 88.0]
  assert kp[hip] == pytest.approx(scenes.G1_ARM_7520_14 * (20 * 3.1415926535) ** 2)


def test_g1_contact_sensor_layout(g1):
  names = g1.names["sensor"]
  assert names[:4] == ["robot/imu_ang_vel", "robot/imu_lin_vel", "robot/imu_lin_acc", "robot/root_angmom"]
  assert names[4:] == ["feet_ground_contact_left_ankle_roll_link_found", "feet_ground_contact_left_ankle_roll_link_force",
                       "feet_ground_contact_right_ankle_roll_link_found", "feet_ground_contact_right_ankle_roll_link_force",
                       "self_collision_pelvis_found"]
  assert g1.sensor_dim.tolist() == [3, 3, 3, 3, 1, 3, 1, 3, 1]
  assert g1.sensor_intprm[4].tolist() == [1, 3, 1] and g1.sensor_intprm[5].tolist() == [2, 3, 1] and g1.sensor_intprm[8].tolist() == [1, 0, 1]
  assert g1.sensor_objtype[4] == 2 and g1.sensor_reftype[4] == 1 and g1.names["body"][g1.sensor_refid[4]] == "terrain"


def test_inertia_from_geoms_closed_forms():
  m = fx.compile_xml(fx.FALLING_BOX)
  b = m.name2id("body", "box")
  assert m.body_mass[b] == pytest.approx(1.0)
  a, bb, c = 0.1, 0.08, 0.06
  assert np.allclose(m.body_inertia[b], [(bb*bb + c*c) / 3, (a*a + c*c) / 3, (a*a + bb*bb) / 3])
  m = fx.compile_xml(fx.CAPSULE_ON_PLANE)
  r, h = 0.05, 0.2
  vc, vs = np.pi * r * r * 2 * h, 4 / 3 * np.pi * r ** 3
  izz = (vc * r * r / 2 + vs * 0.4 * r * r) / (vc + vs)
  assert m.body_inertia[2, 2] == pytest.approx(izz)
  assert m.geom_size[1, 1] == pytest.approx(0.2) and np.allclose(m.geom_pos[1], 0)
  # fromto puts the capsule z axis along the segment
  R = hostmodel._q2m(m.geom_quat[1])
  assert abs(R[0, 2]) == pytest.approx(1.0)
  assert m.geom_rbound[1] == pytest.approx(0.25)


def test_multi_geom_body_principal_axes():
  m = fx.compile_xml(fx.LEGGED)
  b = m.name2id("body", "trunk")
  assert m.body_mass[b] == pytest.approx(4.4)
  assert np.all(np.diff(m.body_inertia[b]) <= 1e-12)        # sorted descending
  R = hostmodel._q2m(m.body_iquat[b])
  assert np.allclose(R @ R.T, np.eye(3), atol=1e-12) and np.linalg.det(R) == pytest.approx(1.0)
  assert m.body_subtreemass[b] == pytest.approx(4.4 + 2 * (0.3 + 0.2 + 0.05) + 0.4)
  assert m.dof_parentid.tolist() == [-1, 0, 1, 2, 3, 4, 5, 6, 5, 8, 5]
  assert m.body_weldid[1] == 0 and m.body_rootid[b] == b
  # invweight0 of a free body chain is positive and finite
  assert (m.dof_invweight0 > 0).all() and np.isfinite(m.body_invweight0).all()


def test_defaults_classes_degrees_and_limits():
  xml = """
  <mujoco>
    <default><joint damping="0.5"/><default class="a"><joint armature="0.1" range="-90 90"/></default></default>
    <worldbody><body name="b" pos="0 0 1">
      <joint name="j0" axis="1 0 0"/>
      <joint name="j1" class="a" axis="0 1 0"/>
      <geom type="sphere" size="0.1"/>
    </body></worldbody>
  </mujoco>"""
  m = Spec.from_string(xml).compile()
  assert m.dof_damping.tolist() == [0.5, 0.5] and m.dof_armature.tolist() == [0.0, 0.1]
  assert m.jnt_limited.tolist() == [0, 1]
  assert np.allclose(m.jnt_range[1], [-np.pi / 2, np.pi / 2])     # default angle unit is degree
  assert m.body_mass[1] == pytest.approx(1000 * 4 / 3 * np.pi * 1e-3)


def test_unsupported_mjcf_raises():
  with pytest.raises(NotImplementedError):
    Spec.from_string("<mujoco><tendon><spatial name='t'/></tendon></mujoco>")
  with pytest.raises(ValueError):           # colliding mesh geom whose mesh asset is missing
    Spec.from_string("<mujoco><worldbody><body><geom type='mesh' mesh='m'/></body></worldbody></mujoco>").compile()
  # mesh inertia is computed from the hull: the unit tetrahedron at density 1000 weighs 1000 / 6
  mt = Spec.from_string("<mujoco><asset><mesh name='m' vertex='0 0 0 1 0 0 0 1 0 0 0 1'/></asset><worldbody><body><joint/>"
                        "<geom type='mesh' mesh='m'/></body></worldbody></mujoco>").compile()
  assert mt.body_mass[1] == pytest.approx(1000.0 / 6.0)
  with pytest.raises(NotImplementedError):  # friction-loss rows use MuJoCo's default solver parameters
    Spec.from_string("<mujoco><worldbody><body><joint frictionloss='1' solreffriction='0.01 1'/><geom size='0.1'/></body></worldbody></mujoco>")


def test_mesh_hull_and_hfield_assets_compile():
  """K7 inputs: a mesh collides by its convex hull (interior vertices dropped), an hfield keeps normalised elevation."""
  xml = """<mujoco><asset>
      <mesh name="tet" vertex="0 0 0  1 0 0  0 1 0  0 0 1  0.2 0.2 0.2" scale="2 2 2"/>
      <hfield name="hf" nrow="3" ncol="4" size="2 1.5 0.5 0.1" elevation="0 0 0 0  0 2 4 0  0 0 0 0"/></asset>
    <worldbody><geom type="hfield" hfield="hf" name="ground"/>
      <body pos="0 0 1"><freejoint/><inertial pos="0 0 0" mass="1" diaginertia="0.1 0.1 0.1"/>
        <geom type="mesh" mesh="tet" name="g"/><geom type="mesh" mesh="tet" name="vis" contype="0" conaffinity="0"/></body>
    </worldbody></mujoco>"""
  m = Spec.from_string(xml).compile()
  assert (m.nmesh, m.nmeshvert, m.nhfield, m.nhfielddata) == (1, 4, 1, 12)
  assert sorted(map(tuple, m.mesh_vert.tolist())) == [(0, 0, 0), (0, 0, 2), (0, 2, 0), (2, 0, 0)]
  assert m.geom_dataid.tolist() == [0, 0, -1] and m.geom_rbound[1] == pytest.approx(2.0)
  assert m.hfield_data.max() == 1.0 and m.hfield_data.reshape(3, 4)[1, 1] == pytest.approx(0.5)
  assert m.hfield_size.tolist() == [[2.0, 1.5, 0.5, 0.1]]


def test_fixed_tendon_and_tendon_actuator_compile():
  """asimov_toe.xml:175-204: <fixed> tendons over two hinges, <position tendon=...> actuators."""
  xml = """<mujoco><compiler angle="radian"/>
    <worldbody><body name="a" pos="0 0 1"><joint name="p" axis="0 1 0"/><geom type="sphere" size="0.1"/>
      <body name="b" pos="0 0 -0.2"><joint name="r" axis="1 0 0"/><geom type="sphere" size="0.05"/></body></body></worldbody>
    <tendon>
      <fixed name="A"><joint joint="p" coef="-0.09"/><joint joint="r" coef="-0.02"/></fixed>
      <fixed name="B" stiffness="3" damping="0.5"><joint joint="p" coef="-0.09"/><joint joint="r" coef="0.02"/></fixed>
    </tendon>
    <actuator>
      <position name="A" tendon="A" kp="300" ctrlrange="-0.08 0.08" forcerange="-72 72"/>
      <position name="B" tendon="B" kp="300" kv="2"/>
      <motor name="m" joint="p"/>
    </actuator></mujoco>"""
  m = Spec.from_string(xml).compile()
  assert (m.ntendon, m.nwrap) == (2, 4)
  assert m.tendon_adr.tolist() == [0, 2] and m.tendon_num.tolist() == [2, 2]
  assert m.wrap_objid.tolist() == [0, 1, 0, 1] and np.allclose(m.wrap_prm, [-0.09, -0.02, -0.09, 0.02])
  assert m.tendon_stiffness.tolist() == [0.0, 3.0] and m.tendon_damping.tolist() == [0.0, 0.5]
  assert m.actuator_trntype.tolist() == [3, 3, 0] and m.actuator_trnid[:, 0].tolist() == [0, 1, 0]
  assert np.allclose(m.actuator_biasprm[1][:3], [0, -300, -2])
  assert m.actuator_ctrllimited.tolist() == [1, 0, 0] and m.actuator_forcelimited.tolist() == [1, 0, 0]


def test_model_blob_round_trip(tmp_path, g1):
  p = tmp_path / "g1.npz"
  scenes.save_model(g1, p)
  m2 = scenes.load_model(p)
  for k, v in g1.__dict__.items():
    if isinstance(v, np.ndarray):
      assert np.array_equal(v, getattr(m2, k)), k
  assert m2.names == g1.names and m2.opt.timestep == g1.opt.timestep and m2.stat.meaninertia == g1.stat.meaninertia


def test_env_origin_grid_matches_reference_formula():
  o = scenes.env_origins_grid(16, 2.5)
  assert o.shape == (16, 3) and np.allclose(o.mean(axis=0), 0)
  assert sorted(set(np.round(o[:, 0], 6))) == [-3.75, -1.25, 1.25, 3.75]


def test_mesh_inertia_from_the_convex_hull():
  """A body whose only geom is a mesh gets its mass, centre of mass and principal inertia from the mesh's convex hull
  (uniform density): a box-shaped mesh reproduces the box closed forms, wherever it sits in the geom frame; a tetrahedron
  matches the textbook volume / centroid."""
  from asimov_mjlab_b200.mjcf import mesh_mass_properties
  hx, hy, hz = 0.1, 0.2, 0.05
  off = np.array([0.03, -0.02, 0.01])
  v = np.array([[sx * hx, sy * hy, sz * hz] for sx in (-1, 1) for sy in (-1, 1) for sz in (-1, 1)], float) + off
  vol, com, I = mesh_mass_properties(v)
  assert np.isclose(vol, 8 * hx * hy * hz) and np.allclose(com, off)
  assert np.allclose(I, np.diag([(hy**2 + hz**2) / 3, (hx**2 + hz**2) / 3, (hx**2 + hy**2) / 3]) * vol, atol=1e-12)
  tet = np.array([[0, 0, 0], [1, 0, 0], [0, 2, 0], [0, 0, 3.0]])
  vol, com, _ = mesh_mass_properties(tet)
  assert np.isclose(vol, 1.0) and np.allclose(com, [0.25, 0.5, 0.75])
  verts = " ".join(f"{x:.6f}" for x in v.reshape(-1))
  xml = f"""
  <mujoco>
    <asset><mesh name="brick" vertex="{verts}"/></asset>
    <worldbody>
      <body name="a" pos="0 0 1"><freejoint/><geom type="mesh" mesh="brick" density="500" pos="0.1 0 0" quat="0.7071068 0 0 0.7071068"/></body>
      <body name="b" pos="0 1 1"><freejoint/><geom type="box" size="{hx} {hy} {hz}" density="500" pos="{0.1 - off[1]} {off[0]} {off[2]}" quat="0.7071068 0 0 0.7071068"/></body>
    </worldbody>
  </mujoco>"""
  m = fx.compile_xml(xml)
  ia, ib = m.names["body"].index("a"), m.names["body"].index("b")
  assert np.isclose(m.body_mass[ia], m.body_mass[ib]) and np.isclose(m.body_mass[ia], 500 * 8 * hx * hy * hz)
  assert np.allclose(m.body_ipos[ia], m.body_ipos[ib], atol=1e-9)
  assert np.allclose(np.sort(m.body_inertia[ia]), np.sort(m.body_inertia[ib]), rtol=1e-9)


def test_hfield_binary_file(tmp_path):
  """<hfield file=...> in MuJoCo's binary format (int32 nrow, int32 ncol, float32 data) equals the inline elevation."""
  elev = np.arange(12, dtype=np.float32).reshape(3, 4)
  p = tmp_path / "field.bin"
  p.write_bytes(np.array([3, 4], "<i4").tobytes() + elev.astype("<f4").tobytes())
  a = Spec.from_string(f"<mujoco><asset><hfield name='h' file='{p}' size='2 1.5 0.5 0.1'/></asset></mujoco>")
  b = Spec.from_string("<mujoco><asset><hfield name='h' nrow='3' ncol='4' size='2 1.5 0.5 0.1' elevation='" + " ".join(str(float(x)) for x in elev.reshape(-1)) + "'/></asset></mujoco>")
  assert a.hfields["h"].nrow == 3 and a.hfields["h"].ncol == 4 and np.allclose(a.hfields["h"].data, b.hfields["h"].data)
  with pytest.raises(NotImplementedError):
    Spec.from_string("<mujoco><asset><hfield name='h' file='x.png' size='1 1 1 1'/></asset></mujoco>")
