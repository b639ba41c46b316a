"""Small MJCF fixtures (own models; same spirit as the fixture strings in the reference's tests,
tests/test_contact_sensor.py:19-86, tests/test_entity_data.py:11-20, tests/test_sim.py:18-40)."""
from asimov_mjlab_b200.mjcf import Spec

SPHERE_DROP = """
<mujoco>
  <compiler angle="radian"/>
  <option timestep="0.002" integrator="euler"/>
  <worldbody>
    <body name="ground"><geom name="floor" type="plane" size="5 5 0.1" condim="1"/></body>
    <body name="ball" pos="0 0 0.5">
      <freejoint name="ball_joint"/>
      <geom name="ball_geom" type="sphere" size="0.1" mass="2.0" condim="1"/>
    </body>
  </worldbody>
</mujoco>
"""

FALLING_BOX = """
<mujoco>
  <compiler angle="radian"/>
  <option timestep="0.002"/>
  <worldbody>
    <body name="ground"><geom name="floor" type="plane" size="5 5 0.1"/></body>
    <body name="box" pos="0 0 0.3" quat="0.98 0.1 0.15 0.05">
      <freejoint name="box_joint"/>
      <geom name="box_geom" type="box" size="0.1 0.08 0.06" mass="1.0"/>
      <site name="box_site" pos="0.05 0 0.02"/>
    </body>
  </worldbody>
  <sensor>
    <gyro name="g" site="box_site"/>
    <velocimeter name="v" site="box_site"/>
    <accelerometer name="a" site="box_site"/>
    <framepos name="p" objtype="site" objname="box_site"/>
    <framequat name="q" objtype="body" objname="box"/>
    <subtreelinvel name="slv" body="box"/>
    <subtreeangmom name="sam" body="box"/>
  </sensor>
</mujoco>
"""

PENDULUM = """
<mujoco>
  <compiler angle="radian"/>
  <option timestep="0.001" integrator="euler"/>
  <worldbody>
    <body name="arm" pos="0 0 1">
      <joint name="hinge" type="hinge" axis="0 1 0"/>
      <geom name="rod" type="capsule" fromto="0 0 0 0 0 -0.5" size="0.02" mass="0" contype="0" conaffinity="0"/>
      <geom name="bob" type="sphere" pos="0 0 -0.5" size="0.05" mass="1.0" contype="0" conaffinity="0"/>
    </body>
  </worldbody>
</mujoco>
"""

SLIDER_PD = """
<mujoco>
  <compiler angle="radian"/>
  <option timestep="0.005" integrator="implicitfast" gravity="0 0 0"/>
  <worldbody>
    <body name="cart" pos="0 0 1">
      <joint name="slide" type="slide" axis="1 0 0" range="-10 10"/>
      <geom name="cart_geom" type="box" size="0.1 0.1 0.1" mass="3.0" contype="0" conaffinity="0"/>
    </body>
  </worldbody>
  <actuator>
    <position name="pd" joint="slide" kp="50" kv="4"/>
  </actuator>
</mujoco>
"""

# two sliders with dry friction (joint frictionloss), one hinge arm with friction on both joints under gravity
SLIDER_FRICTION = """
<mujoco>
  <compiler angle="radian"/>
  <option timestep="0.005" integrator="implicitfast" gravity="0 0 -9.81"/>
  <worldbody>
    <body name="cart" pos="0 0 1">
      <joint name="s1" type="slide" axis="1 0 0" frictionloss="2.0"/>
      <geom type="box" size="0.1 0.1 0.1" mass="1.0" contype="0" conaffinity="0"/>
    </body>
    <body name="cart2" pos="0 1 1">
      <joint name="s2" type="slide" axis="1 0 0" frictionloss="0.5" damping="0.3"/>
      <geom type="box" size="0.1 0.1 0.1" mass="4.0" contype="0" conaffinity="0"/>
    </body>
    <body name="arm" pos="0 2 1">
      <joint name="h1" type="hinge" axis="0 1 0" frictionloss="0.4" range="-2 2"/>
      <geom type="capsule" fromto="0 0 0 0.4 0 0" size="0.03" mass="1.0" contype="0" conaffinity="0"/>
      <body name="fore" pos="0.4 0 0">
        <joint name="h2" type="hinge" axis="0 1 0" frictionloss="0.1"/>
        <geom type="capsule" fromto="0 0 0 0.3 0 0" size="0.02" mass="0.4" contype="0" conaffinity="0"/>
      </body>
    </body>
  </worldbody>
</mujoco>
"""

# free cylinder, ellipsoid and sphere over a plane and a static box (cylinder / ellipsoid colliders: analytic against the
# plane, GJK / EPA against everything else)
ROUND_SHAPES = """
<mujoco>
  <compiler angle="radian"/>
  <option timestep="0.005" integrator="implicitfast"/>
  <worldbody>
    <geom name="floor" type="plane" size="5 5 0.1"/>
    <geom name="block" type="box" pos="1.0 0 0.1" size="0.3 0.3 0.1"/>
    <body name="cyl" pos="0 0 0.5">
      <freejoint name="cyl_free"/>
      <geom name="cyl_geom" type="cylinder" size="0.1 0.15" mass="1.0"/>
    </body>
    <body name="ell" pos="0 1 0.5">
      <freejoint name="ell_free"/>
      <geom name="ell_geom" type="ellipsoid" size="0.1 0.15 0.2" mass="2.0"/>
    </body>
    <body name="ball" pos="0 2 0.5">
      <freejoint name="ball_free"/>
      <geom name="ball_geom" type="sphere" size="0.05" mass="0.2"/>
    </body>
  </worldbody>
</mujoco>
"""

SPINNER = """
<mujoco>
  <compiler angle="radian"/>
  <option timestep="0.002" integrator="euler" gravity="0 0 0"/>
  <worldbody>
    <body name="ball" pos="0 0 1">
      <freejoint name="free"/>
      <geom name="ball_geom" type="sphere" size="0.1" mass="1.0" contype="0" conaffinity="0"/>
    </body>
  </worldbody>
</mujoco>
"""

TWO_SPHERES = """
<mujoco>
  <compiler angle="radian"/>
  <option gravity="0 0 0"/>
  <worldbody>
    <body name="a" pos="0 0 1"><freejoint/><geom name="ga" type="sphere" size="0.1" mass="1" condim="1"/></body>
    <body name="b" pos="0.15 0.0 1.0"><freejoint/><geom name="gb" type="sphere" size="0.08" mass="1" condim="1"/></body>
  </worldbody>
</mujoco>
"""

CAPSULE_ON_PLANE = """
<mujoco>
  <compiler angle="radian"/>
  <worldbody>
    <body name="ground"><geom name="floor" type="plane" size="5 5 0.1"/></body>
    <body name="cap" pos="0 0 0.045">
      <freejoint/>
      <geom name="cap_geom" type="capsule" fromto="-0.2 0 0 0.2 0 0" size="0.05" mass="1"/>
    </body>
  </worldbody>
</mujoco>
"""

LIMITED_HINGE = """
<mujoco>
  <compiler angle="radian"/>
  <option timestep="0.002"/>
  <worldbody>
    <body name="arm" pos="0 0 1">
      <joint name="hinge" type="hinge" axis="0 1 0" range="-0.3 0.4" damping="0.1"/>
      <geom name="bob" type="sphere" pos="0.3 0 0" size="0.05" mass="1.0" contype="0" conaffinity="0"/>
    </body>
  </worldbody>
</mujoco>
"""

# free-floating multi-geom robot over a plane: spheres + capsules + a box trunk (plane-box, plane-capsule, plane-sphere)
LEGGED = """
<mujoco>
  <compiler angle="radian"/>
  <option timestep="0.004" integrator="implicitfast"/>
  <default>
    <default class="leg"><geom type="capsule" size="0.02" friction="0.8 0.01 0.001" conaffinity="0"/></default>
  </default>
  <worldbody>
    <body name="ground"><geom name="floor" type="plane" size="5 5 0.1"/></body>
    <body name="trunk" pos="0 0 0.35" childclass="leg">
      <freejoint name="root"/>
      <geom name="trunk_geom" type="box" size="0.2 0.1 0.05" mass="4.0"/>
      <geom name="head_geom" type="sphere" size="0.07" pos="0.25 0 0.05" mass="0.4"/>
      <site name="imu" pos="0.02 0 0.01"/>
      <body name="fl" pos="0.15 0.1 -0.05">
        <joint name="fl_hip" axis="0 1 0" range="-1 1" armature="0.01" damping="0.2"/>
        <geom name="fl_thigh" fromto="0 0 0 0 0 -0.15" mass="0.3"/>
        <body name="fl2" pos="0 0 -0.15">
          <joint name="fl_knee" axis="0 1 0" range="-2 0" armature="0.01"/>
          <geom name="fl_shin" fromto="0 0 0 0 0 -0.15" mass="0.2"/>
          <geom name="fl_foot" type="sphere" size="0.03" pos="0 0 -0.15" mass="0.05" condim="3" priority="1" friction="0.6 0.005 0.0001"/>
        </body>
      </body>
      <body name="fr" pos="0.15 -0.1 -0.05">
        <joint name="fr_hip" axis="0 1 0" range="-1 1" armature="0.01" damping="0.2"/>
        <geom name="fr_thigh" fromto="0 0 0 0 0 -0.15" mass="0.3"/>
        <body name="fr2" pos="0 0 -0.15">
          <joint name="fr_knee" axis="0 1 0" range="-2 0" armature="0.01"/>
          <geom name="fr_shin" fromto="0 0 0 0 0 -0.15" mass="0.2"/>
          <geom name="fr_foot" type="sphere" size="0.03" pos="0 0 -0.15" mass="0.05" condim="3" priority="1" friction="0.6 0.005 0.0001"/>
        </body>
      </body>
      <body name="rear" pos="-0.15 0 -0.05">
        <joint name="rear_slide" type="slide" axis="0 0 1" range="-0.05 0.05" stiffness="200" damping="2"/>
        <geom name="rear_leg" fromto="0 0 0 0 0 -0.25" mass="0.4"/>
      </body>
    </body>
  </worldbody>
  <actuator>
    <position name="fl_hip" joint="fl_hip" kp="20" kv="1" inheritrange="1" forcerange="-10 10"/>
    <position name="fl_knee" joint="fl_knee" kp="20" kv="1" inheritrange="1" forcerange="-10 10"/>
    <position name="fr_hip" joint="fr_hip" kp="20" kv="1" inheritrange="1" forcerange="-10 10"/>
    <motor name="fr_knee" joint="fr_knee" ctrlrange="-5 5"/>
  </actuator>
  <sensor>
    <gyro name="gyro" site="imu"/>
    <accelerometer name="acc" site="imu"/>
    <jointpos name="jp" joint="fl_knee"/>
    <jointvel name="jv" joint="fl_knee"/>
    <actuatorfrc name="af" actuator="fl_hip"/>
    <subtreecom name="com" body="trunk"/>
    <framelinvel name="flv" objtype="body" objname="fl2"/>
    <frameangvel name="fav" objtype="xbody" objname="fl2"/>
  </sensor>
</mujoco>
"""


BOX_ON_BOX = """
<mujoco>
  <compiler angle="radian"/>
  <option timestep="0.002"/>
  <worldbody>
    <body name="terrain"><geom name="slab" type="box" size="1 1 0.25" pos="0 0 -0.25"/></body>
    <body name="crate" pos="0.1 -0.05 0.1">
      <freejoint name="crate_joint"/>
      <geom name="crate_geom" type="box" size="0.15 0.1 0.1" mass="3.0"/>
    </body>
    <body name="rod" pos="-0.5 0.4 0.0499">
      <freejoint name="rod_joint"/>
      <geom name="rod_geom" type="capsule" fromto="-0.2 0 0 0.2 0 0" size="0.05" mass="1.5"/>
    </body>
  </worldbody>
</mujoco>
"""


TENDON_SLIDERS = """
<mujoco>
  <compiler angle="radian"/>
  <option timestep="0.005" integrator="implicitfast" gravity="0 0 0"/>
  <worldbody>
    <body name="c1" pos="0 0 1">
      <joint name="s1" type="slide" axis="1 0 0" range="-10 10"/>
      <geom type="box" size="0.1 0.1 0.1" mass="3.0" contype="0" conaffinity="0"/>
    </body>
    <body name="c2" pos="0 1 1">
      <joint name="s2" type="slide" axis="1 0 0" range="-10 10"/>
      <geom type="box" size="0.1 0.1 0.1" mass="0.5" contype="0" conaffinity="0"/>
    </body>
  </worldbody>
  <tendon>
    <fixed name="t" stiffness="20" damping="1.5" springlength="0.05"><joint joint="s1" coef="0.7"/><joint joint="s2" coef="-0.4"/></fixed>
  </tendon>
  <actuator>
    <position name="pd" tendon="t" kp="50" kv="4"/>
  </actuator>
</mujoco>
"""

# TENDON_ANKLE-like arm with one of every scalar / frame sensor mjlab's BuiltinSensorCfg accepts beyond the robots' own
# (sensor/builtin_sensor.py:59-101): tendon, actuator, joint-actuator-force, frame axes, frame accelerations, clock.
SENSOR_ARM = """
<mujoco>
  <compiler angle="radian"/>
  <option timestep="0.005" integrator="implicitfast"/>
  <worldbody>
    <geom name="floor" type="plane" size="5 5 0.1"/>
    <body name="upper" pos="0 0 1.2">
      <joint name="sh" type="hinge" axis="0 1 0" damping="0.2"/>
      <geom type="capsule" fromto="0 0 0 0.4 0 0" size="0.04" mass="1.0"/>
      <site name="mid" pos="0.2 0 0.05" quat="0.9238795 0 0 0.3826834"/>
      <body name="fore" pos="0.4 0 0">
        <joint name="el" type="hinge" axis="0 0 1" damping="0.1" range="-2 2"/>
        <geom type="capsule" fromto="0 0 0 0.3 0 0" size="0.03" mass="0.5"/>
        <site name="tip" pos="0.3 0 0"/>
      </body>
    </body>
  </worldbody>
  <tendon>
    <fixed name="couple"><joint joint="sh" coef="0.5"/><joint joint="el" coef="-1.5"/></fixed>
  </tendon>
  <actuator>
    <position name="a_sh" joint="sh" kp="30" kv="2" gear="2"/>
    <position name="a_t" tendon="couple" kp="10" kv="1"/>
  </actuator>
  <sensor>
    <tendonpos name="tp" tendon="couple"/><tendonvel name="tv" tendon="couple"/>
    <actuatorpos name="ap" actuator="a_sh"/><actuatorvel name="av" actuator="a_sh"/><actuatorfrc name="af" actuator="a_t"/>
    <jointactuatorfrc name="jf" joint="el"/>
    <framexaxis name="fx" objtype="site" objname="mid"/><frameyaxis name="fy" objtype="site" objname="tip" reftype="site" refname="mid"/>
    <framezaxis name="fz" objtype="xbody" objname="fore"/>
    <framelinacc name="la" objtype="site" objname="tip"/><frameangacc name="aa" objtype="site" objname="tip"/>
    <accelerometer name="acc" site="tip"/>
    <clock name="clk"/>
  </sensor>
</mujoco>
"""

# the remaining scalar sensors of mjlab's BuiltinSensorCfg list that need no ray casting / reaction forces: magnetometer, joint-limit
# position / velocity / force, tendon actuator force, potential and kinetic energy (sensor/builtin_sensor.py:59-101)
SENSOR_MISC = """
<mujoco>
  <compiler angle="radian"/>
  <option timestep="0.004" integrator="implicitfast" magnetic="0.2 -0.5 0.1"/>
  <worldbody>
    <body name="cart" pos="0 0 1">
      <joint name="sl" type="slide" axis="0 0 1" stiffness="40" springref="0.1" range="-0.3 0.3" damping="0.5"/>
      <geom type="box" size="0.1 0.1 0.1" mass="2.0" contype="0" conaffinity="0"/>
      <body name="arm" pos="0 0 0.1">
        <joint name="hg" type="hinge" axis="0 1 0" range="-0.5 0.6" damping="0.05" stiffness="3"/>
        <geom type="capsule" fromto="0 0 0 0.3 0 0" size="0.03" mass="0.7" contype="0" conaffinity="0"/>
        <site name="tip" pos="0.3 0 0" quat="0.9238795 0 0.3826834 0"/>
      </body>
    </body>
  </worldbody>
  <tendon>
    <fixed name="tn" stiffness="15" springlength="-0.05 0.05"><joint joint="sl" coef="1.0"/><joint joint="hg" coef="0.3"/></fixed>
  </tendon>
  <actuator>
    <position name="a1" tendon="tn" kp="20" kv="1"/>
    <motor name="a2" tendon="tn" gear="2"/>
    <motor name="a3" joint="hg"/>
  </actuator>
  <sensor>
    <magnetometer name="mag" site="tip"/>
    <jointlimitpos name="lp" joint="hg"/><jointlimitvel name="lv" joint="hg"/><jointlimitfrc name="lf" joint="hg"/>
    <jointlimitpos name="lps" joint="sl"/><jointlimitfrc name="lfs" joint="sl"/>
    <tendonactuatorfrc name="taf" tendon="tn"/>
    <e_potential name="ep"/><e_kinetic name="ek"/>
  </sensor>
</mujoco>
"""

# force / torque sensors (sensor/builtin_sensor.py:59-101; MuJoCo's mj_rnePostConstraint cfrc_int): a mast on a horizontal slider
# carrying a hinged boom (reaction forces through joints, a two-body subtree), and a free puck with a hinged lid resting on the
# plane (contact forces enter cfrc_ext; the root of a free body transmits nothing)
SENSOR_FORCE = """
<mujoco>
  <compiler angle="radian"/>
  <option timestep="0.004" integrator="implicitfast"/>
  <worldbody>
    <geom name="floor" type="plane" size="5 5 0.1"/>
    <body name="mast" pos="0 0 1">
      <joint name="sx" type="slide" axis="1 0 0" damping="0.2"/>
      <geom type="box" size="0.1 0.1 0.1" mass="2.0" contype="0" conaffinity="0"/>
      <site name="s_mast" pos="0.05 0.02 -0.03" quat="0.9238795 0 0.3826834 0"/>
      <body name="boom" pos="0 0 0.1">
        <joint name="hb" type="hinge" axis="0 1 0" damping="0.05"/>
        <geom type="capsule" fromto="0 0 0 0.4 0 0" size="0.03" mass="0.7" contype="0" conaffinity="0"/>
        <site name="s_boom" pos="0 0 0"/>
      </body>
    </body>
    <body name="puck" pos="1 0 0.06">
      <freejoint/>
      <geom name="puck_g" type="box" size="0.1 0.1 0.05" mass="1.5"/>
      <site name="s_puck" pos="0.02 0 0.01"/>
      <body name="lid" pos="0 0 0.05">
        <joint name="hl" type="hinge" axis="1 0 0" range="-0.2 1.2" damping="0.02"/>
        <geom name="lid_g" type="capsule" fromto="0 0 0.03 0 0.25 0.03" size="0.03" mass="0.4"/>
        <site name="s_lid" pos="0 0 0" quat="0.9659258 0.2588190 0 0"/>
      </body>
    </body>
  </worldbody>
  <actuator>
    <motor name="mb" joint="hb"/>
    <position name="pl" joint="hl" kp="5" kv="0.2"/>
  </actuator>
  <sensor>
    <force name="f_mast" site="s_mast"/><torque name="t_mast" site="s_mast"/>
    <force name="f_boom" site="s_boom"/><torque name="t_boom" site="s_boom"/>
    <force name="f_puck" site="s_puck"/><torque name="t_puck" site="s_puck"/>
    <force name="f_lid" site="s_lid"/><torque name="t_lid" site="s_lid" cutoff="0.5"/>
  </sensor>
</mujoco>
"""

# rangefinder (sensor/builtin_sensor.py:59-101; MuJoCo mj_ray along the site's z axis): a free "drone" with six sites looking along
# -z, +x, -x, +y, -y, +z at one primitive of each kind, and a free capsule that falls through the +x ray
SENSOR_RAY = """
<mujoco>
  <compiler angle="radian"/>
  <option timestep="0.004" integrator="implicitfast"/>
  <worldbody>
    <geom name="floor" type="plane" size="5 5 0.1"/>
    <geom name="ball" type="sphere" pos="3 0 2" size="0.5"/>
    <geom name="crate" type="box" pos="-2 0 2" size="0.3 0.4 0.5"/>
    <geom name="post" type="capsule" pos="0 3 2" size="0.2 0.5"/>
    <geom name="can" type="cylinder" pos="0 -2.5 2" quat="0.7071068 0.7071068 0 0" size="0.3 0.4"/>
    <geom name="egg" type="ellipsoid" pos="0 0 4" size="0.5 0.4 0.3"/>
    <body name="drone" pos="0 0 2">
      <freejoint/>
      <geom type="sphere" size="0.1" mass="1.0"/>
      <site name="s_down" quat="0 1 0 0"/>
      <site name="s_px" quat="0.7071068 0 0.7071068 0"/>
      <site name="s_nx" quat="0.7071068 0 -0.7071068 0"/>
      <site name="s_py" quat="0.7071068 -0.7071068 0 0"/>
      <site name="s_ny" quat="0.7071068 0.7071068 0 0"/>
      <site name="s_up"/>
    </body>
    <body name="rod" pos="1.5 0 2.6">
      <freejoint/>
      <geom type="capsule" fromto="0 -0.3 0 0 0.3 0" size="0.05" mass="0.2"/>
    </body>
  </worldbody>
  <sensor>
    <rangefinder name="r_down" site="s_down"/><rangefinder name="r_px" site="s_px"/><rangefinder name="r_nx" site="s_nx"/>
    <rangefinder name="r_py" site="s_py"/><rangefinder name="r_ny" site="s_ny"/><rangefinder name="r_up" site="s_up"/>
    <rangefinder name="r_down_cut" site="s_down" cutoff="1.0"/>
  </sensor>
</mujoco>
"""

# Asimov-toe's ankle in miniature: a pitch/roll hinge pair under gravity driven by two fixed tendons with position
# actuators (one clamped by ctrlrange / forcerange, one with a velocity gain), a sprung + damped third tendon with
# a dead band, and a joint motor on the same dof (asimov_toe.xml:175-204).
TENDON_ANKLE = """
<mujoco>
  <compiler angle="radian"/>
  <option timestep="0.005" integrator="implicitfast"/>
  <worldbody>
    <body name="shank" pos="0 0 1">
      <joint name="pitch" axis="0 1 0" range="-0.6 0.6" armature="0.01" damping="0.05"/>
      <geom type="capsule" fromto="0 0 0 0 0 -0.1" size="0.02" mass="0.3" contype="0" conaffinity="0"/>
      <body name="foot" pos="0 0 -0.1">
        <joint name="roll" axis="1 0 0" range="-0.4 0.4" armature="0.01"/>
        <geom type="box" size="0.1 0.04 0.015" pos="0.03 0 -0.02" mass="0.6" contype="0" conaffinity="0"/>
        <body name="toe" pos="0.13 0 -0.02">
          <joint name="toe" axis="0 1 0" range="-0.6 0.1" armature="0.001" stiffness="2" damping="0.02"/>
          <geom type="box" size="0.03 0.04 0.01" pos="0.03 0 0" mass="0.1" contype="0" conaffinity="0"/>
        </body>
      </body>
    </body>
  </worldbody>
  <tendon>
    <fixed name="A"><joint joint="pitch" coef="-0.09"/><joint joint="roll" coef="-0.02"/></fixed>
    <fixed name="B"><joint joint="pitch" coef="-0.09"/><joint joint="roll" coef="0.02"/></fixed>
    <fixed name="C" stiffness="40" damping="0.8" springlength="-0.01 0.02"><joint joint="roll" coef="0.05"/><joint joint="toe" coef="0.1"/></fixed>
  </tendon>
  <actuator>
    <position name="A" tendon="A" kp="300" ctrlrange="-0.08 0.08" forcerange="-12 12"/>
    <position name="B" tendon="B" kp="300" kv="10"/>
    <motor name="m" joint="pitch" gear="2"/>
  </actuator>
</mujoco>
"""


def compile_xml(xml, contact_sensors=()):
  spec = Spec.from_string(xml)
  for kw in contact_sensors:
    spec.add_sensor(**kw)
  return spec.compile()


def legged_with_contact_sensors():
  """LEGGED plus contact sensors in every reduce mode / object type (reference contact_sensor.py:16-47)."""
  S = []
  allbits = 1 | 2 | 4 | 8 | 16 | 32 | 64
  S.append(dict(name="c_feet_net", type="contact", objtype="xbody", objname="fl", reftype="body", refname="ground", intprm=[1 | 2, 3, 1]))
  S.append(dict(name="c_trunk_none", type="contact", objtype="body", objname="trunk", intprm=[allbits, 0, 2]))
  S.append(dict(name="c_any_mindist", type="contact", objtype="xbody", objname="trunk", intprm=[1 | 8 | 16, 1, 3]))
  S.append(dict(name="c_any_maxforce", type="contact", objtype="xbody", objname="trunk", reftype="geom", refname="floor", intprm=[1 | 2 | 32, 2, 2]))
  S.append(dict(name="c_geom", type="contact", objtype="geom", objname="fr_foot", intprm=[1 | 2 | 4, 3, 1]))
  return compile_xml(LEGGED, S)


def box_mesh_vertices(hx, hy, hz):
  return " ".join(f"{sx*hx} {sy*hy} {sz*hz}" for sx in (-1, 1) for sy in (-1, 1) for sz in (-1, 1))


# convex-mesh fixtures: "cube" is a mesh whose hull is a box, so that GJK/EPA and the plane-mesh manifold can be checked
# against the analytic box colliders on identical geometry
def convex_scene(ground="plane", movers=("mesh",), hfield_elev=None, hfield_size=(2.0, 2.0, 0.4, 0.2), nrow=9, ncol=9, margin=0.0):
  elev = hfield_elev if hfield_elev is not None else [0.0]*(nrow*ncol)
  g = {"plane": '<geom name="floor" type="plane" size="5 5 0.1"/>',
       "hfield": '<geom name="floor" type="hfield" hfield="hf"/>',
       "box": '<geom name="floor" type="box" size="1 1 0.1" pos="0 0 -0.1"/>',
       "meshbox": '<geom name="floor" type="mesh" mesh="slab" pos="0 0 -0.1"/>'}[ground]
  kinds = {"mesh": '<geom name="m{i}" type="mesh" mesh="cube" margin="%g"/>' % margin,
           "box": '<geom name="m{i}" type="box" size="0.1 0.08 0.06" margin="%g"/>' % margin,
           "sphere": '<geom name="m{i}" type="sphere" size="0.1" margin="%g"/>' % margin,
           "capsule": '<geom name="m{i}" type="capsule" size="0.05 0.1" margin="%g"/>' % margin}
  bodies = "".join(
    f'<body name="b{i}" pos="{0.5*i} 0 0.5"><freejoint/><inertial pos="0 0 0" mass="1" diaginertia="0.01 0.01 0.01"/>'
    + kinds[k].format(i=i) + "</body>" for i, k in enumerate(movers))
  return f"""
<mujoco>
  <compiler angle="radian"/>
  <option timestep="0.002"/>
  <asset>
    <mesh name="cube" vertex="{box_mesh_vertices(0.1, 0.08, 0.06)}"/>
    <mesh name="slab" vertex="{box_mesh_vertices(1.0, 1.0, 0.1)}"/>
    <hfield name="hf" nrow="{nrow}" ncol="{ncol}" size="{' '.join(map(str, hfield_size))}" elevation="{' '.join(map(str, elev))}"/>
  </asset>
  <worldbody>
    <body name="ground">{g}</body>
    {bodies}
  </worldbody>
</mujoco>
"""
