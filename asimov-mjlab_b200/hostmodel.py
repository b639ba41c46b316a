"""Host-side model container and compile-time derived quantities.

`HostModel` carries numpy arrays under MuJoCo `MjModel` field names (the
names the reference's managers and `expand_model_fields` use; reference
src/mjlab/envs/mdp/events.py:228-253, src/mjlab/sim/sim.py:177-184).
`finalize_model` derives what MuJoCo's compiler would (tree tables,
`body_subtreemass`, `dof_invweight0`, `body_invweight0`, `stat.meaninertia`)
with a small float64 numpy kinematics + CRBA evaluated at `qpos0`
(SURVEY.md §7.3 H1).  This runs once per model; it is not the hot path.
"""
from __future__ import annotations

import numpy as np


class _Opt:
  def __init__(self):
    self.timestep = 0.002
    self.gravity = np.array([0.0, 0.0, -9.81])
    self.integrator = 0
    self.cone = 0
    self.solver = 2
    self.iterations = 100
    self.ls_iterations = 50
    self.tolerance = 1e-8
    self.ls_tolerance = 0.01
    self.impratio = 1.0
    self.disableflags = 0
    self.jacobian = 2  # auto


class _Stat:
  meaninertia = 1.0
  extent = 2.0


class HostModel:
  """Attribute bag with MjModel field names (numpy arrays, float64 / int32)."""
  ntendon = 0     # defaults for model blobs exported before tendons / meshes / height fields existed
  nwrap = 0
  nmesh = 0
  nmeshvert = 0
  nhfield = 0
  nhfielddata = 0

  def __init__(self):
    self.opt = _Opt()
    self.stat = _Stat()
    self.names: dict[str, list[str]] = {}
    self.keys: dict = {}

  # name lookups in the style of mj_model.sensor(name).adr etc.
  def name2id(self, kind: str, name: str) -> int:
    return self.names[kind].index(name)

  def sensor_slice(self, name: str) -> slice:
    i = self.name2id("sensor", name)
    return slice(int(self.sensor_adr[i]), int(self.sensor_adr[i] + self.sensor_dim[i]))


class HostData:
  """Initial state container (the MjData fields `put_data` reads)."""

  def __init__(self, m: HostModel):
    self.qpos = m.qpos0.copy()
    self.qvel = np.zeros(m.nv)
    self.ctrl = np.zeros(m.nu)
    self.mocap_pos = np.array([m.body_pos[i] for i in range(m.nbody) if m.body_mocapid[i] >= 0]).reshape(-1, 3)
    self.mocap_quat = np.array([m.body_quat[i] for i in range(m.nbody) if m.body_mocapid[i] >= 0]).reshape(-1, 4)
    self.time = 0.0


def _qmul(a, b):
  aw, ax, ay, az = a
  bw, bx, by, bz = b
  return np.array([aw*bw - ax*bx - ay*by - az*bz, aw*bx + ax*bw + ay*bz - az*by,
                   aw*by - ax*bz + ay*bw + az*bx, aw*bz + ax*by - ay*bx + az*bw])


def _q2m(q):
  w, x, y, z = q
  return np.array([[w*w+x*x-y*y-z*z, 2*(x*y-w*z), 2*(x*z+w*y)],
                   [2*(x*y+w*z), w*w-x*x+y*y-z*z, 2*(y*z-w*x)],
                   [2*(x*z-w*y), 2*(y*z+w*x), w*w-x*x-y*y+z*z]])


def host_kinematics(m: HostModel, qpos):
  """float64 forward kinematics + com_pos (used for compile-time constants and
  as an independent cross-check in tests)."""
  nb = m.nbody
  xpos = np.zeros((nb, 3)); xquat = np.tile([1.0, 0, 0, 0], (nb, 1)); xmat = np.tile(np.eye(3), (nb, 1, 1))
  xanchor = np.zeros((m.njnt, 3)); xaxis = np.zeros((m.njnt, 3))
  for i in range(1, nb):
    p = m.body_parentid[i]
    pos = xpos[p] + xmat[p] @ m.body_pos[i]
    quat = _qmul(xquat[p], m.body_quat[i])
    for j in range(m.body_jntadr[i], m.body_jntadr[i] + m.body_jntnum[i]) if m.body_jntnum[i] else ():
      t, qa = m.jnt_type[j], m.jnt_qposadr[j]
      if t == 0:
        pos = qpos[qa:qa+3].copy(); quat = qpos[qa+3:qa+7] / np.linalg.norm(qpos[qa+3:qa+7])
        xanchor[j] = pos; xaxis[j] = _q2m(quat) @ m.jnt_axis[j]
        continue
      R = _q2m(quat)
      xanchor[j] = pos + R @ m.jnt_pos[j]; xaxis[j] = R @ m.jnt_axis[j]
      if t == 2:
        pos = pos + xaxis[j]*(qpos[qa] - m.qpos0[qa])
      elif t == 3:
        a = qpos[qa] - m.qpos0[qa]
        quat = _qmul(quat, np.array([np.cos(a/2), *(m.jnt_axis[j]*np.sin(a/2))]))
        pos = xanchor[j] - _q2m(quat) @ m.jnt_pos[j]
      elif t == 1:
        ql = qpos[qa:qa+4] / np.linalg.norm(qpos[qa:qa+4])
        quat = _qmul(quat, ql)
        pos = xanchor[j] - _q2m(quat) @ m.jnt_pos[j]
    quat = quat / np.linalg.norm(quat)
    xpos[i], xquat[i], xmat[i] = pos, quat, _q2m(quat)
  xipos = np.array([xpos[i] + xmat[i] @ m.body_ipos[i] for i in range(nb)])
  ximat = np.array([_q2m(_qmul(xquat[i], m.body_iquat[i])) for i in range(nb)])
  sub = m.body_mass[:, None] * xipos
  for i in range(nb - 1, 0, -1):
    sub[m.body_parentid[i]] += sub[i]
  com = np.where(m.body_subtreemass[:, None] > 1e-15, sub / np.maximum(m.body_subtreemass[:, None], 1e-15), xipos)
  return dict(xpos=xpos, xquat=xquat, xmat=xmat, xipos=xipos, ximat=ximat, xanchor=xanchor,
              xaxis=xaxis, subtree_com=com)


def host_cdof_cinert(m: HostModel, k):
  nb = m.nbody
  com = k["subtree_com"]
  cinert = np.zeros((nb, 10))
  for i in range(nb):
    R = k["ximat"][i]
    I = R @ np.diag(m.body_inertia[i]) @ R.T
    d = k["xipos"][i] - com[m.body_rootid[i]]
    ms = m.body_mass[i]
    I = I + ms*(np.dot(d, d)*np.eye(3) - np.outer(d, d))
    cinert[i] = [I[0, 0], I[1, 1], I[2, 2], I[0, 1], I[0, 2], I[1, 2], ms*d[0], ms*d[1], ms*d[2], ms]
  cdof = np.zeros((m.nv, 6))
  for j in range(m.njnt):
    b = m.jnt_bodyid[j]; da = m.jnt_dofadr[j]; t = m.jnt_type[j]
    off = com[m.body_rootid[b]] - k["xanchor"][j]
    if t == 0:
      for a in range(3):
        cdof[da+a, 3+a] = 1.0
      for a in range(3):
        ax = k["xmat"][b][:, a]
        cdof[da+3+a, :3] = ax; cdof[da+3+a, 3:] = np.cross(ax, off)
    elif t == 1:
      for a in range(3):
        ax = k["xmat"][b][:, a]
        cdof[da+a, :3] = ax; cdof[da+a, 3:] = np.cross(ax, off)
    elif t == 2:
      cdof[da, 3:] = k["xaxis"][j]
    else:
      cdof[da, :3] = k["xaxis"][j]; cdof[da, 3:] = np.cross(k["xaxis"][j], off)
  return cinert, cdof


def _mul_inert(I, v):
  r = np.zeros(6)
  r[0] = I[0]*v[0] + I[3]*v[1] + I[4]*v[2] - I[8]*v[4] + I[7]*v[5]
  r[1] = I[3]*v[0] + I[1]*v[1] + I[5]*v[2] + I[8]*v[3] - I[6]*v[5]
  r[2] = I[4]*v[0] + I[5]*v[1] + I[2]*v[2] - I[7]*v[3] + I[6]*v[4]
  r[3] = I[8]*v[1] - I[7]*v[2] + I[9]*v[3]
  r[4] = I[6]*v[2] - I[8]*v[0] + I[9]*v[4]
  r[5] = I[7]*v[0] - I[6]*v[1] + I[9]*v[5]
  return r


def host_mass_matrix(m: HostModel, qpos):
  k = host_kinematics(m, qpos)
  cinert, cdof = host_cdof_cinert(m, k)
  crb = cinert.copy()
  for i in range(m.nbody - 1, 0, -1):
    crb[m.body_parentid[i]] += crb[i]
  M = np.zeros((m.nv, m.nv))
  for i in range(m.nv):
    buf = _mul_inert(crb[m.dof_bodyid[i]], cdof[i])
    j = i
    while j >= 0:
      M[i, j] = M[j, i] = np.dot(cdof[j], buf)
      j = m.dof_parentid[j]
    M[i, i] += m.dof_armature[i]
  return M, k, cdof


def host_jac(m: HostModel, k, cdof, body, point):
  """3xnv translational and rotational Jacobians of `point` on `body`."""
  jp = np.zeros((3, m.nv)); jr = np.zeros((3, m.nv))
  off = point - k["subtree_com"][m.body_rootid[body]]
  b = body
  while b > 0 and m.body_dofnum[b] == 0:
    b = m.body_parentid[b]
  if b == 0:
    return jp, jr
  i = m.body_dofadr[b] + m.body_dofnum[b] - 1
  while i >= 0:
    jr[:, i] = cdof[i, :3]
    jp[:, i] = cdof[i, 3:] + np.cross(cdof[i, :3], off)
    i = m.dof_parentid[i]
  return jp, jr


def finalize_model(m: HostModel) -> None:
  nb = m.nbody
  m.body_rootid = np.zeros(nb, np.int32); m.body_weldid = np.zeros(nb, np.int32)
  for i in range(1, nb):
    p = m.body_parentid[i]
    m.body_rootid[i] = i if p == 0 else m.body_rootid[p]
    m.body_weldid[i] = i if m.body_jntnum[i] > 0 else m.body_weldid[p]
  m.dof_parentid = -np.ones(m.nv, np.int32)
  for i in range(1, nb):
    if m.body_dofnum[i] == 0:
      continue
    a = m.body_dofadr[i]
    p = m.body_parentid[i]
    while p > 0 and m.body_dofnum[p] == 0:
      p = m.body_parentid[p]
    m.dof_parentid[a] = -1 if p == 0 else m.body_dofadr[p] + m.body_dofnum[p] - 1
    for d in range(a + 1, a + m.body_dofnum[i]):
      m.dof_parentid[d] = d - 1
  m.body_subtreemass = m.body_mass.copy()
  for i in range(nb - 1, 0, -1):
    m.body_subtreemass[m.body_parentid[i]] += m.body_subtreemass[i]
  m.body_invweight0 = np.zeros((nb, 2)); m.dof_invweight0 = np.zeros(m.nv)
  if m.nv:
    M, k, cdof = host_mass_matrix(m, m.qpos0)
    m.stat.meaninertia = float(np.trace(M) / m.nv)
    Minv = np.linalg.inv(M)
    for i in range(1, nb):
      if m.body_weldid[i] == 0:
        continue
      jp, jr = host_jac(m, k, cdof, i, k["xipos"][i])
      m.body_invweight0[i, 0] = np.trace(jp @ Minv @ jp.T) / 3
      m.body_invweight0[i, 1] = np.trace(jr @ Minv @ jr.T) / 3
    d = np.diag(Minv)
    for j in range(m.njnt):
      a, t = m.jnt_dofadr[j], m.jnt_type[j]
      if t == 0:
        m.dof_invweight0[a:a+3] = d[a:a+3].mean(); m.dof_invweight0[a+3:a+6] = d[a+3:a+6].mean()
      elif t == 1:
        m.dof_invweight0[a:a+3] = d[a:a+3].mean()
      else:
        m.dof_invweight0[a] = d[a]
  else:
    m.stat.meaninertia = 1.0
