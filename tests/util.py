"""Shared helpers for the parity tests (seeded states, oracle/engine comparison)."""
import numpy as np


def g1_states(m, nworld, seed=0, spread=0.3, vel=0.5, zlo=0.70, zhi=0.80):
  """Seeded G1 states around the knees-bent keyframe: joint noise, base height/tilt noise, random velocities."""
  rng = np.random.default_rng(seed)
  key = m.keys["robot/init_state"]
  q0 = np.array(key["qpos"]); c0 = np.array(key["ctrl"])
  qpos = np.tile(q0, (nworld, 1)); qvel = np.zeros((nworld, m.nv)); ctrl = np.tile(c0, (nworld, 1))
  lo, hi = m.jnt_range[1:, 0], m.jnt_range[1:, 1]
  for w in range(nworld):
    s = spread * (w / max(nworld - 1, 1))
    qpos[w, 7:] = np.clip(q0[7:] + rng.normal(0, s, m.nq - 7), lo - 0.05, hi + 0.05)
    qpos[w, 2] = rng.uniform(zlo, zhi)
    qpos[w, 0:2] = rng.normal(0, 1.0, 2)
    ax = rng.normal(size=3); ax /= np.linalg.norm(ax)
    ang = rng.normal(0, 0.15)
    qpos[w, 3:7] = [np.cos(ang / 2), *(ax * np.sin(ang / 2))]
    qvel[w] = rng.normal(0, vel, m.nv) * (w > 0)
    ctrl[w] = c0 + rng.normal(0, 0.2, m.nu) * (w > 0)
  return qpos, qvel, ctrl


def contact_pairs(geom, n):
  return sorted(map(tuple, np.asarray(geom).reshape(-1, 2)[:n].tolist()))


def legged_states(m, nworld, seed=0, spread=0.3, vel=0.5, zlo=None, zhi=None):
  """Same for any floating-base robot with an init_state keyframe (Go1, ...)."""
  key = m.keys["robot/init_state"]
  z0 = key["qpos"][2]
  return g1_states(m, nworld, seed, spread, vel, z0 - 0.03 if zlo is None else zlo, z0 + 0.07 if zhi is None else zhi)


def asimov_states(m, nworld, seed=0, spread=0.2, vel=0.3, zlo=0.70, zhi=0.76):
  """Asimov-toe: like legged_states; the ctrl noise of the four tendon position actuators (length targets in metres,
  ctrlrange +-0.08) is scaled down so that some worlds sit inside the ctrl and force clamps and some outside."""
  qpos, qvel, ctrl = legged_states(m, nworld, seed, spread, vel, zlo, zhi)
  c0 = np.array(m.keys["robot/init_state"]["ctrl"])
  tend = np.asarray(m.actuator_trntype) == 3
  ctrl[:, tend] = c0[tend] + 0.25 * (ctrl[:, tend] - c0[tend])
  return qpos, qvel, ctrl


def rough_states(m, nworld, seed=0, spread=0.2, vel=0.3, drop=(0.0, 0.08), tumble=0.0):
  """Go1 on the box terrain: worlds spawn at the generator's terrain origins (all difficulty rows, flat / stairs /
  inverted-stairs columns) with a random xy offset (so feet land on step edges), a small drop height and tilt."""
  rng = np.random.default_rng(seed)
  org = np.asarray(m.meta["terrain_origins"])
  qpos, qvel, ctrl = legged_states(m, nworld, seed, spread, vel)
  key = m.keys["robot/init_state"]
  for w in range(nworld):
    o = org[rng.integers(org.shape[0]), rng.integers(org.shape[1])]
    off = rng.uniform(-2.2, 2.2, 2)
    qpos[w, 0:2] = o[:2] + off
    # ground height under the spawn point is not known in closed form on stairs: start from the platform height and
    # let the robot drop (stairs descend away from the platform on pyramids, ascend on inverted ones)
    qpos[w, 2] = max(o[2], 0.0) + key["qpos"][2] + rng.uniform(*drop)
    if tumble > 0.0 and w % 2 == 1:   # every other world: large random tilt, lower start -> leg capsules / trunk box touch
      ax = rng.normal(size=3); ax /= np.linalg.norm(ax)
      ang = rng.normal(0, tumble)
      qpos[w, 3:7] = [np.cos(ang / 2), *(ax * np.sin(ang / 2))]
      qpos[w, 2] -= 0.12
  return qpos, qvel, ctrl
