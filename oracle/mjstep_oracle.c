/* mjstep_oracle.c -- CPU restatement of the physics step the reference delegates
 * to mujoco_warp (reference call sites src/mjlab/sim/sim.py:143,146).
 *
 * TEST INFRASTRUCTURE ONLY (see oracle.h).  PARITY UNPINNED (see oracle.h).
 *
 * The algorithm is the one MuJoCo documents and implements in its engine
 * (third-party, absent from /root/reference; pins: mujoco 3.4.0.dev818719010,
 * mujoco-warp @ e605c406, uv.lock:1124-1126,1151-1153).  Each function names the
 * upstream routine whose published behaviour it restates:
 *   orc_kinematics      engine_core_smooth.c  mj_kinematics
 *   orc_com_pos         engine_core_smooth.c  mj_comPos
 *   orc_crb/factor_m    engine_core_smooth.c  mj_crb / mj_factorM (dense Cholesky here, as mjwarp's dense path)
 *   orc_collision       engine_collision_driver.c + engine_collision_primitive.c
 *   orc_make_constraint engine_core_constraint.c mj_instantiateLimit/Contact, mj_makeImpedance
 *   orc_fwd_velocity    mj_comVel, mj_passive, mj_rne
 *   orc_fwd_actuation   engine_forward.c mj_fwdActuation
 *   orc_fwd_acceleration engine_forward.c mj_fwdAcceleration
 *   orc_solve           engine_solver.c mj_solNewton / mj_solCG (mj_solPrimal) (+ mjwarp's parallel line search)
 *   orc_sensors         engine_sensor.c mj_sensorPos/Vel/Acc, mj_subtreeVel, mj_rnePostConstraint (cacc only)
 *   orc_integrate       engine_forward.c mj_Euler / mj_implicit (implicitfast)
 * and the reference files that fix WHAT must come out: sim.py:41-76 (options),
 * entity/data.py:20-31 (cvel frame), sensor/contact_sensor.py:16-41,472-533
 * (contact-sensor encoding), utils/spec_config.py:361-414 (actuator model).
 */
#include "oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define MINVAL ((real)1e-15)
#define MINMU ((real)1e-5)
#define NV_MAX 128

int orc_real_size(void) { return (int)sizeof(real); }

/* ------------------------------------------------------------------ math */
static real dot3(const real* a, const real* b) { return a[0]*b[0] + a[1]*b[1] + a[2]*b[2]; }
static void cross3(real* r, const real* a, const real* b) {
  real x = a[1]*b[2] - a[2]*b[1], y = a[2]*b[0] - a[0]*b[2], z = a[0]*b[1] - a[1]*b[0];
  r[0] = x; r[1] = y; r[2] = z;
}
static real norm3(const real* a) { return (real)sqrt((double)dot3(a, a)); }
static real normalize3(real* a) {
  real n = norm3(a);
  if (n < MINVAL) { a[0] = 1; a[1] = 0; a[2] = 0; }
  else { a[0] /= n; a[1] /= n; a[2] /= n; }
  return n;
}
static void normalize4(real* q) {
  real n = (real)sqrt((double)(q[0]*q[0] + q[1]*q[1] + q[2]*q[2] + q[3]*q[3]));
  if (n < MINVAL) { q[0] = 1; q[1] = q[2] = q[3] = 0; }
  else { q[0] /= n; q[1] /= n; q[2] /= n; q[3] /= n; }
}
static void quat_mul(real* r, const real* a, const real* b) {
  real w = a[0]*b[0] - a[1]*b[1] - a[2]*b[2] - a[3]*b[3];
  real x = a[0]*b[1] + a[1]*b[0] + a[2]*b[3] - a[3]*b[2];
  real y = a[0]*b[2] - a[1]*b[3] + a[2]*b[0] + a[3]*b[1];
  real z = a[0]*b[3] + a[1]*b[2] - a[2]*b[1] + a[3]*b[0];
  r[0] = w; r[1] = x; r[2] = y; r[3] = z;
}
static void quat2mat(real* m, const real* q) {
  real q00 = q[0]*q[0], q11 = q[1]*q[1], q22 = q[2]*q[2], q33 = q[3]*q[3];
  real q01 = q[0]*q[1], q02 = q[0]*q[2], q03 = q[0]*q[3], q12 = q[1]*q[2], q13 = q[1]*q[3], q23 = q[2]*q[3];
  m[0] = q00 + q11 - q22 - q33; m[4] = q00 - q11 + q22 - q33; m[8] = q00 - q11 - q22 + q33;
  m[1] = 2*(q12 - q03); m[2] = 2*(q13 + q02);
  m[3] = 2*(q12 + q03); m[5] = 2*(q23 - q01);
  m[6] = 2*(q13 - q02); m[7] = 2*(q23 + q01);
}
static void mat_vec(real* r, const real* m, const real* v) {
  real x = m[0]*v[0] + m[1]*v[1] + m[2]*v[2];
  real y = m[3]*v[0] + m[4]*v[1] + m[5]*v[2];
  real z = m[6]*v[0] + m[7]*v[1] + m[8]*v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
static void matT_vec(real* r, const real* m, const real* v) {
  real x = m[0]*v[0] + m[3]*v[1] + m[6]*v[2];
  real y = m[1]*v[0] + m[4]*v[1] + m[7]*v[2];
  real z = m[2]*v[0] + m[5]*v[1] + m[8]*v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
static void axis_angle_quat(real* q, const real* axis, real angle) {
  real s = (real)sin((double)angle * 0.5);
  q[0] = (real)cos((double)angle * 0.5); q[1] = axis[0]*s; q[2] = axis[1]*s; q[3] = axis[2]*s;
}
/* spatial helpers: vectors are [angular(3), linear(3)] */
static void mul_inert_vec(real* r, const real* I, const real* v) {
  r[0] = I[0]*v[0] + I[3]*v[1] + I[4]*v[2] - I[8]*v[4] + I[7]*v[5];
  r[1] = I[3]*v[0] + I[1]*v[1] + I[5]*v[2] + I[8]*v[3] - I[6]*v[5];
  r[2] = I[4]*v[0] + I[5]*v[1] + I[2]*v[2] - I[7]*v[3] + I[6]*v[4];
  r[3] = I[8]*v[1] - I[7]*v[2] + I[9]*v[3];
  r[4] = I[6]*v[2] - I[8]*v[0] + I[9]*v[4];
  r[5] = I[7]*v[0] - I[6]*v[1] + I[9]*v[5];
}
static void cross_motion(real* r, const real* vel, const real* v) {
  r[0] = -vel[2]*v[1] + vel[1]*v[2];
  r[1] =  vel[2]*v[0] - vel[0]*v[2];
  r[2] = -vel[1]*v[0] + vel[0]*v[1];
  r[3] = -vel[2]*v[4] + vel[1]*v[5] - vel[5]*v[1] + vel[4]*v[2];
  r[4] =  vel[2]*v[3] - vel[0]*v[5] + vel[5]*v[0] - vel[3]*v[2];
  r[5] = -vel[1]*v[3] + vel[0]*v[4] - vel[4]*v[0] + vel[3]*v[1];
}
static void cross_force(real* r, const real* vel, const real* f) {
  r[0] = -vel[2]*f[1] + vel[1]*f[2] - vel[5]*f[4] + vel[4]*f[5];
  r[1] =  vel[2]*f[0] - vel[0]*f[2] + vel[5]*f[3] - vel[3]*f[5];
  r[2] = -vel[1]*f[0] + vel[0]*f[1] - vel[4]*f[3] + vel[3]*f[4];
  r[3] = -vel[2]*f[4] + vel[1]*f[5];
  r[4] =  vel[2]*f[3] - vel[0]*f[5];
  r[5] = -vel[1]*f[3] + vel[0]*f[4];
}

/* ------------------------------------------------------------ kinematics */
void orc_kinematics(const b2s_model_desc* m, orc_data* d) {
  real* xpos = d->xpos; real* xquat = d->xquat; real* xmat = d->xmat;
  xpos[0] = xpos[1] = xpos[2] = 0; xquat[0] = 1; xquat[1] = xquat[2] = xquat[3] = 0;
  quat2mat(xmat, xquat);
  for (int i = 1; i < m->nbody; i++) {
    int p = m->body_parentid[i];
    real pos[3], quat[4], bp[3], bq[4];
    int mid = m->body_mocapid[i];
    if (mid >= 0) {
      for (int k = 0; k < 3; k++) bp[k] = d->mocap_pos[3*mid + k];
      for (int k = 0; k < 4; k++) bq[k] = d->mocap_quat[4*mid + k];
      normalize4(bq);
    } else {
      for (int k = 0; k < 3; k++) bp[k] = (real)m->body_pos[3*i + k];
      for (int k = 0; k < 4; k++) bq[k] = (real)m->body_quat[4*i + k];
    }
    mat_vec(pos, xmat + 9*p, bp);
    for (int k = 0; k < 3; k++) pos[k] += xpos[3*p + k];
    quat_mul(quat, xquat + 4*p, bq);
    for (int jj = 0; jj < m->body_jntnum[i]; jj++) {
      int j = m->body_jntadr[i] + jj, qa = m->jnt_qposadr[j], t = m->jnt_type[j];
      real jpos[3] = {(real)m->jnt_pos[3*j], (real)m->jnt_pos[3*j+1], (real)m->jnt_pos[3*j+2]};
      real jaxis[3] = {(real)m->jnt_axis[3*j], (real)m->jnt_axis[3*j+1], (real)m->jnt_axis[3*j+2]};
      real R[9];
      if (t == B2S_JNT_FREE) {
        for (int k = 0; k < 3; k++) pos[k] = d->qpos[qa + k];
        for (int k = 0; k < 4; k++) quat[k] = d->qpos[qa + 3 + k];
        normalize4(quat);
        quat2mat(R, quat);
        for (int k = 0; k < 3; k++) d->xanchor[3*j + k] = pos[k];
        mat_vec(d->xaxis + 3*j, R, jaxis);
        continue;
      }
      quat2mat(R, quat);
      real anchor[3];
      mat_vec(anchor, R, jpos);
      for (int k = 0; k < 3; k++) anchor[k] += pos[k];
      mat_vec(d->xaxis + 3*j, R, jaxis);
      for (int k = 0; k < 3; k++) d->xanchor[3*j + k] = anchor[k];
      if (t == B2S_JNT_SLIDE) {
        real q = d->qpos[qa] - (real)m->qpos0[qa];
        for (int k = 0; k < 3; k++) pos[k] += d->xaxis[3*j + k]*q;
      } else {
        real ql[4], qn[4], off[3];
        if (t == B2S_JNT_HINGE) axis_angle_quat(ql, jaxis, d->qpos[qa] - (real)m->qpos0[qa]);
        else { for (int k = 0; k < 4; k++) ql[k] = d->qpos[qa + k]; normalize4(ql); }
        quat_mul(qn, quat, ql);
        for (int k = 0; k < 4; k++) quat[k] = qn[k];
        quat2mat(R, quat);
        mat_vec(off, R, jpos);
        for (int k = 0; k < 3; k++) pos[k] = anchor[k] - off[k];
      }
    }
    normalize4(quat);
    for (int k = 0; k < 3; k++) xpos[3*i + k] = pos[k];
    for (int k = 0; k < 4; k++) xquat[4*i + k] = quat[k];
    quat2mat(xmat + 9*i, quat);
  }
  for (int i = 0; i < m->nbody; i++) {
    real ip[3] = {(real)m->body_ipos[3*i], (real)m->body_ipos[3*i+1], (real)m->body_ipos[3*i+2]};
    real iq[4] = {(real)m->body_iquat[4*i], (real)m->body_iquat[4*i+1], (real)m->body_iquat[4*i+2], (real)m->body_iquat[4*i+3]};
    real q[4];
    mat_vec(d->xipos + 3*i, xmat + 9*i, ip);
    for (int k = 0; k < 3; k++) d->xipos[3*i + k] += xpos[3*i + k];
    quat_mul(q, xquat + 4*i, iq);
    quat2mat(d->ximat + 9*i, q);
  }
  for (int g = 0; g < m->ngeom; g++) {
    int b = m->geom_bodyid[g];
    real gp[3] = {(real)m->geom_pos[3*g], (real)m->geom_pos[3*g+1], (real)m->geom_pos[3*g+2]};
    real gq[4] = {(real)m->geom_quat[4*g], (real)m->geom_quat[4*g+1], (real)m->geom_quat[4*g+2], (real)m->geom_quat[4*g+3]};
    real q[4];
    mat_vec(d->geom_xpos + 3*g, xmat + 9*b, gp);
    for (int k = 0; k < 3; k++) d->geom_xpos[3*g + k] += xpos[3*b + k];
    quat_mul(q, xquat + 4*b, gq);
    quat2mat(d->geom_xmat + 9*g, q);
  }
  for (int s = 0; s < m->nsite; s++) {
    int b = m->site_bodyid[s];
    real sp[3] = {(real)m->site_pos[3*s], (real)m->site_pos[3*s+1], (real)m->site_pos[3*s+2]};
    real sq[4] = {(real)m->site_quat[4*s], (real)m->site_quat[4*s+1], (real)m->site_quat[4*s+2], (real)m->site_quat[4*s+3]};
    real q[4];
    mat_vec(d->site_xpos + 3*s, xmat + 9*b, sp);
    for (int k = 0; k < 3; k++) d->site_xpos[3*s + k] += xpos[3*b + k];
    quat_mul(q, xquat + 4*b, sq);
    quat2mat(d->site_xmat + 9*s, q);
  }
}

/* --------------------------------------------------------------- com_pos */
void orc_com_pos(const b2s_model_desc* m, orc_data* d) {
  int nb = m->nbody;
  real* com = d->subtree_com;
  for (int i = 0; i < nb; i++)
    for (int k = 0; k < 3; k++) com[3*i + k] = (real)m->body_mass[i]*d->xipos[3*i + k];
  for (int i = nb - 1; i > 0; i--) {
    int p = m->body_parentid[i];
    for (int k = 0; k < 3; k++) com[3*p + k] += com[3*i + k];
  }
  for (int i = 0; i < nb; i++) {
    real sm = (real)m->body_subtreemass[i];
    if (sm < MINVAL) for (int k = 0; k < 3; k++) com[3*i + k] = d->xipos[3*i + k];
    else for (int k = 0; k < 3; k++) com[3*i + k] /= sm;
  }
  for (int i = 0; i < nb; i++) {
    const real* R = d->ximat + 9*i;
    real in[3] = {(real)m->body_inertia[3*i], (real)m->body_inertia[3*i+1], (real)m->body_inertia[3*i+2]};
    real mass = (real)m->body_mass[i];
    real off[3];
    for (int k = 0; k < 3; k++) off[k] = d->xipos[3*i + k] - com[3*m->body_rootid[i] + k];
    real* c = d->cinert + 10*i;
    /* R diag(in) R^T, upper triangle */
    real Ixx = 0, Iyy = 0, Izz = 0, Ixy = 0, Ixz = 0, Iyz = 0;
    for (int k = 0; k < 3; k++) {
      Ixx += R[0+k]*in[k]*R[0+k]; Iyy += R[3+k]*in[k]*R[3+k]; Izz += R[6+k]*in[k]*R[6+k];
      Ixy += R[0+k]*in[k]*R[3+k]; Ixz += R[0+k]*in[k]*R[6+k]; Iyz += R[3+k]*in[k]*R[6+k];
    }
    real o2 = dot3(off, off);
    c[0] = Ixx + mass*(o2 - off[0]*off[0]);
    c[1] = Iyy + mass*(o2 - off[1]*off[1]);
    c[2] = Izz + mass*(o2 - off[2]*off[2]);
    c[3] = Ixy - mass*off[0]*off[1];
    c[4] = Ixz - mass*off[0]*off[2];
    c[5] = Iyz - mass*off[1]*off[2];
    c[6] = mass*off[0]; c[7] = mass*off[1]; c[8] = mass*off[2]; c[9] = mass;
  }
  for (int j = 0; j < m->njnt; j++) {
    int b = m->jnt_bodyid[j], da = m->jnt_dofadr[j], t = m->jnt_type[j];
    real off[3];
    for (int k = 0; k < 3; k++) off[k] = com[3*m->body_rootid[b] + k] - d->xanchor[3*j + k];
    real* cd = d->cdof + 6*da;
    if (t == B2S_JNT_FREE) {
      for (int a = 0; a < 3; a++) { for (int k = 0; k < 6; k++) cd[6*a + k] = 0; cd[6*a + 3 + a] = 1; }
      cd += 18;
    }
    if (t == B2S_JNT_FREE || t == B2S_JNT_BALL) {
      for (int a = 0; a < 3; a++) {
        real ax[3] = {d->xmat[9*b + a], d->xmat[9*b + 3 + a], d->xmat[9*b + 6 + a]};
        for (int k = 0; k < 3; k++) cd[6*a + k] = ax[k];
        cross3(cd + 6*a + 3, ax, off);
      }
    } else if (t == B2S_JNT_SLIDE) {
      for (int k = 0; k < 3; k++) { cd[k] = 0; cd[3 + k] = d->xaxis[3*j + k]; }
    } else {
      for (int k = 0; k < 3; k++) cd[k] = d->xaxis[3*j + k];
      cross3(cd + 3, d->xaxis + 3*j, off);
    }
  }
}

/* ------------------------------------------------------- crb + factor_m */
void orc_crb(const b2s_model_desc* m, orc_data* d) {
  int nb = m->nbody, nv = m->nv;
  real* crb = (real*)malloc(sizeof(real)*10*nb);
  memcpy(crb, d->cinert, sizeof(real)*10*nb);
  for (int i = nb - 1; i > 0; i--) {
    int p = m->body_parentid[i];
    if (p > 0) for (int k = 0; k < 10; k++) crb[10*p + k] += crb[10*i + k];
  }
  for (int i = 0; i < nv*nv; i++) d->qM[i] = 0;
  for (int i = 0; i < nv; i++) {
    real buf[6];
    mul_inert_vec(buf, crb + 10*m->dof_bodyid[i], d->cdof + 6*i);
    int j = i;
    while (j >= 0) {
      real v = 0;
      for (int k = 0; k < 6; k++) v += d->cdof[6*j + k]*buf[k];
      d->qM[i*nv + j] = v; d->qM[j*nv + i] = v;
      j = m->dof_parentid[j];
    }
    d->qM[i*nv + i] += (real)m->dof_armature[i];
  }
  free(crb);
}

/* dense Cholesky A = L L^T (lower, row-major n x n); returns rank deficiency count */
static int chol_factor(real* L, const real* A, int n) {
  int bad = 0;
  for (int i = 0; i < n*n; i++) L[i] = 0;
  for (int j = 0; j < n; j++) {
    real s = A[j*n + j];
    for (int k = 0; k < j; k++) s -= L[j*n + k]*L[j*n + k];
    if (s < MINVAL) { s = MINVAL; bad++; }
    real dj = (real)sqrt((double)s);
    L[j*n + j] = dj;
    for (int i = j + 1; i < n; i++) {
      real t = A[i*n + j];
      for (int k = 0; k < j; k++) t -= L[i*n + k]*L[j*n + k];
      L[i*n + j] = t/dj;
    }
  }
  return bad;
}
static void chol_solve(real* x, const real* L, const real* b, int n) {
  for (int i = 0; i < n; i++) {
    real s = b[i];
    for (int k = 0; k < i; k++) s -= L[i*n + k]*x[k];
    x[i] = s/L[i*n + i];
  }
  for (int i = n - 1; i >= 0; i--) {
    real s = x[i];
    for (int k = i + 1; k < n; k++) s -= L[k*n + i]*x[k];
    x[i] = s/L[i*n + i];
  }
}
int orc_factor_m(const b2s_model_desc* m, orc_data* d) { return chol_factor(d->qL, d->qM, m->nv); }

/* -------------------------------------------------------------- collision */
typedef struct { real dist; real pos[3]; real frame[9]; } rawcon;

/* engine_util_misc / mju_makeFrame: complete a frame whose x axis (and maybe y axis) is given */
static void make_frame(real* f) {
  normalize3(f);
  if (norm3(f + 3) < (real)0.5) {
    f[3] = f[4] = f[5] = 0;
    if (f[1] < (real)0.5 && f[1] > (real)-0.5) f[4] = 1; else f[5] = 1;
  }
  real t = dot3(f, f + 3);
  for (int k = 0; k < 3; k++) f[3 + k] -= t*f[k];
  normalize3(f + 3);
  cross3(f + 6, f, f + 3);
}

static int raw_plane_sphere(rawcon* c, real margin, const real* ppos, const real* n, const real* spos, real r) {
  real tmp[3] = {spos[0] - ppos[0], spos[1] - ppos[1], spos[2] - ppos[2]};
  real cdist = dot3(tmp, n);
  if (cdist > margin + r) return 0;
  c->dist = cdist - r;
  for (int k = 0; k < 3; k++) { c->pos[k] = spos[k] - n[k]*(c->dist*(real)0.5 + r); c->frame[k] = n[k]; c->frame[3 + k] = 0; }
  return 1;
}
static int raw_sphere_sphere(rawcon* c, real margin, const real* p1, real r1, const real* p2, real r2) {
  real dif[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
  real cd = norm3(dif);
  if (cd > margin + r1 + r2) return 0;
  c->dist = cd - r1 - r2;
  if (cd < MINVAL) { c->frame[0] = 1; c->frame[1] = 0; c->frame[2] = 0; }
  else for (int k = 0; k < 3; k++) c->frame[k] = dif[k]/cd;
  for (int k = 0; k < 3; k++) { c->pos[k] = p1[k] + c->frame[k]*(r1 + c->dist*(real)0.5); c->frame[3 + k] = 0; }
  return 1;
}
static real clipr(real x, real lo, real hi) { return x < lo ? lo : (x > hi ? hi : x); }

static int col_plane_capsule(rawcon* c, real margin, const real* p1, const real* m1, const real* p2, const real* m2, const real* s2) {
  real n[3] = {m1[2], m1[5], m1[8]}, axis[3] = {m2[2], m2[5], m2[8]};
  real seg[3] = {axis[0]*s2[1], axis[1]*s2[1], axis[2]*s2[1]};
  real a[3] = {p2[0] + seg[0], p2[1] + seg[1], p2[2] + seg[2]};
  real b[3] = {p2[0] - seg[0], p2[1] - seg[1], p2[2] - seg[2]};
  int cnt = raw_plane_sphere(c, margin, p1, n, a, s2[0]);
  cnt += raw_plane_sphere(c + cnt, margin, p1, n, b, s2[0]);
  /* align the tangent with the capsule axis projected on the plane */
  real t = dot3(axis, n);
  real y[3] = {axis[0] - n[0]*t, axis[1] - n[1]*t, axis[2] - n[2]*t};
  if (norm3(y) >= (real)0.5) normalize3(y); else y[0] = y[1] = y[2] = 0;
  for (int i = 0; i < cnt; i++) for (int k = 0; k < 3; k++) c[i].frame[3 + k] = y[k];
  return cnt;
}
static int col_plane_box(rawcon* c, real margin, const real* p1, const real* m1, const real* p2, const real* m2, const real* s2) {
  real n[3] = {m1[2], m1[5], m1[8]};
  real dif[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
  real dist = dot3(dif, n);
  int cnt = 0;
  for (int i = 0; i < 8; i++) {
    real v[3] = {(i & 1 ? s2[0] : -s2[0]), (i & 2 ? s2[1] : -s2[1]), (i & 4 ? s2[2] : -s2[2])};
    real corner[3];
    mat_vec(corner, m2, v);
    real ld = dist + dot3(n, corner);
    if (ld > margin) continue;
    c[cnt].dist = ld;
    for (int k = 0; k < 3; k++) { c[cnt].pos[k] = p2[k] + corner[k] - n[k]*ld*(real)0.5; c[cnt].frame[k] = n[k]; c[cnt].frame[3 + k] = 0; }
    if (++cnt >= 4) break;
  }
  return cnt;
}
static int col_sphere_capsule(rawcon* c, real margin, const real* p1, const real* s1, const real* p2, const real* m2, const real* s2) {
  real axis[3] = {m2[2], m2[5], m2[8]};
  real vec[3] = {p1[0] - p2[0], p1[1] - p2[1], p1[2] - p2[2]};
  real x = clipr(dot3(axis, vec), -s2[1], s2[1]);
  real q[3] = {p2[0] + axis[0]*x, p2[1] + axis[1]*x, p2[2] + axis[2]*x};
  return raw_sphere_sphere(c, margin, p1, s1[0], q, s2[0]);
}
static int col_capsule_capsule(rawcon* c, real margin, const real* p1, const real* m1, const real* s1, const real* p2, const real* m2, const real* s2) {
  real a1[3] = {m1[2], m1[5], m1[8]}, a2[3] = {m2[2], m2[5], m2[8]};
  real dif[3] = {p1[0] - p2[0], p1[1] - p2[1], p1[2] - p2[2]};
  real ma = dot3(a1, a1), mb = -dot3(a1, a2), mc = dot3(a2, a2);
  real u = -dot3(a1, dif), v = dot3(a2, dif);
  real det = ma*mc - mb*mb;
  real v1[3], v2[3];
  if (fabs((double)det) >= (double)MINVAL) {
    real x1 = (mc*u - mb*v)/det, x2 = (ma*v - mb*u)/det;
    if (x1 > s1[1]) { x1 = s1[1]; x2 = (v - mb*s1[1])/mc; }
    else if (x1 < -s1[1]) { x1 = -s1[1]; x2 = (v + mb*s1[1])/mc; }
    if (x2 > s2[1]) { x2 = s2[1]; x1 = clipr((u - mb*s2[1])/ma, -s1[1], s1[1]); }
    else if (x2 < -s2[1]) { x2 = -s2[1]; x1 = clipr((u + mb*s2[1])/ma, -s1[1], s1[1]); }
    for (int k = 0; k < 3; k++) { v1[k] = p1[k] + a1[k]*x1; v2[k] = p2[k] + a2[k]*x2; }
    return raw_sphere_sphere(c, margin, v1, s1[0], v2, s2[0]);
  }
  /* parallel axes: up to two contacts at segment ends */
  int n = 0;
  real x2, x1;
  for (int k = 0; k < 3; k++) v1[k] = p1[k] + a1[k]*s1[1];
  x2 = clipr((v - mb*s1[1])/mc, -s2[1], s2[1]);
  for (int k = 0; k < 3; k++) v2[k] = p2[k] + a2[k]*x2;
  n += raw_sphere_sphere(c + n, margin, v1, s1[0], v2, s2[0]);
  for (int k = 0; k < 3; k++) v1[k] = p1[k] - a1[k]*s1[1];
  x2 = clipr((v + mb*s1[1])/mc, -s2[1], s2[1]);
  for (int k = 0; k < 3; k++) v2[k] = p2[k] + a2[k]*x2;
  n += raw_sphere_sphere(c + n, margin, v1, s1[0], v2, s2[0]);
  if (n >= 2) return n;
  for (int k = 0; k < 3; k++) v2[k] = p2[k] + a2[k]*s2[1];
  x1 = clipr((u - mb*s2[1])/ma, -s1[1], s1[1]);
  for (int k = 0; k < 3; k++) v1[k] = p1[k] + a1[k]*x1;
  n += raw_sphere_sphere(c + n, margin, v1, s1[0], v2, s2[0]);
  if (n >= 2) return n;
  for (int k = 0; k < 3; k++) v2[k] = p2[k] - a2[k]*s2[1];
  x1 = clipr((u + mb*s2[1])/ma, -s1[1], s1[1]);
  for (int k = 0; k < 3; k++) v1[k] = p1[k] + a1[k]*x1;
  n += raw_sphere_sphere(c + n, margin, v1, s1[0], v2, s2[0]);
  return n;
}
static int col_sphere_box(rawcon* c, real margin, const real* p1, const real* s1, const real* p2, const real* m2, const real* s2) {
  real dif[3] = {p1[0] - p2[0], p1[1] - p2[1], p1[2] - p2[2]}, loc[3], cl[3];
  matT_vec(loc, m2, dif);
  int inside = 1;
  for (int k = 0; k < 3; k++) { cl[k] = clipr(loc[k], -s2[k], s2[k]); if (cl[k] != loc[k]) inside = 0; }
  real nl[3], dist, r = s1[0];
  if (!inside) {
    real dv[3] = {loc[0] - cl[0], loc[1] - cl[1], loc[2] - cl[2]};
    real dn = norm3(dv);
    if (dn - r > margin) return 0;
    dist = dn - r;
    for (int k = 0; k < 3; k++) nl[k] = -dv[k]/dn;   /* sphere -> box */
  } else {
    int kb = 0; real best = s2[0] - (real)fabs((double)loc[0]);
    for (int k = 1; k < 3; k++) { real t = s2[k] - (real)fabs((double)loc[k]); if (t < best) { best = t; kb = k; } }
    nl[0] = nl[1] = nl[2] = 0;
    nl[kb] = loc[kb] >= 0 ? (real)-1 : (real)1;
    dist = -best - r;
  }
  mat_vec(c->frame, m2, nl);
  c->dist = dist;
  for (int k = 0; k < 3; k++) { c->pos[k] = p1[k] + c->frame[k]*(r + dist*(real)0.5); c->frame[3 + k] = 0; }
  return 1;
}

/* capsule-box: the capsule is a swept sphere, so each contact is the sphere-box contact of one point of the
 * segment.  Point 1 is the segment point closest to the box (or, if the segment enters the box, the sampled point
 * that is deepest inside); point 2 is a segment end whose own sphere-box contact exists and faces the same way
 * (capsule lying along a face or an edge).  Own formulation (the reference's dependency ports MuJoCo's
 * mjc_CapsuleBox, whose source is not available here): same contact model, possibly different tie-breaking. */
static real seg_box_closest(const real* a, const real* dd, const real* h, int* inside_out) {
  real ts[8]; int nt = 0;
  ts[nt++] = 0; ts[nt++] = 1;
  for (int i = 0; i < 3; i++) if (dd[i] != 0) for (int sgn = -1; sgn <= 1; sgn += 2) {
    real t = (sgn*h[i] - a[i])/dd[i];
    if (t > 0 && t < 1) ts[nt++] = t;
  }
  for (int i = 1; i < nt; i++) { real v = ts[i]; int j = i - 1; while (j >= 0 && ts[j] > v) { ts[j + 1] = ts[j]; j--; } ts[j + 1] = v; }
  real best = (real)1e30, bt = 0, tin = 0; int inside = 0;
  for (int k = 0; k + 1 < nt; k++) {
    real t0 = ts[k], t1 = ts[k + 1], tm = (real)0.5*(t0 + t1), A = 0, B = 0;
    int nout = 0;
    for (int i = 0; i < 3; i++) {
      real pm = a[i] + tm*dd[i];
      real r = pm > h[i] ? h[i] : (pm < -h[i] ? -h[i] : 0);
      if (pm > h[i] || pm < -h[i]) { A += dd[i]*dd[i]; B += dd[i]*(a[i] - r); nout++; }
    }
    real t = t0;
    if (nout == 0) { inside = 1; tin = tm; }      /* the part of the segment inside the box is ONE interval: take its middle */
    else if (A > MINVAL) t = clipr(-B/A, t0, t1);
    real d2 = 0;
    for (int i = 0; i < 3; i++) { real pv = a[i] + t*dd[i]; real c = clipr(pv, -h[i], h[i]); d2 += (pv - c)*(pv - c); }
    if (d2 < best) { best = d2; bt = t; }
  }
  *inside_out = inside;
  return inside ? tin : bt;
}
static int col_capsule_box(rawcon* c, real margin, const real* p1, const real* m1, const real* s1, const real* p2, const real* m2, const real* s2) {
  real axis[3] = {m1[2], m1[5], m1[8]}, e0[3], e1[3], a[3], b[3], dd[3], tmp[3];
  for (int k = 0; k < 3; k++) { e0[k] = p1[k] - axis[k]*s1[1]; e1[k] = p1[k] + axis[k]*s1[1]; }
  for (int k = 0; k < 3; k++) tmp[k] = e0[k] - p2[k];
  matT_vec(a, m2, tmp);
  for (int k = 0; k < 3; k++) tmp[k] = e1[k] - p2[k];
  matT_vec(b, m2, tmp);
  for (int k = 0; k < 3; k++) dd[k] = b[k] - a[k];
  int inside;
  real tb = seg_box_closest(a, dd, s2, &inside);
  real pt[3], rad[3] = {s1[0], 0, 0};
  for (int k = 0; k < 3; k++) pt[k] = e0[k] + tb*(e1[k] - e0[k]);
  int n;
  if (!inside) n = col_sphere_box(c, margin, pt, rad, p2, m2, s2);
  else {
    /* the segment passes through the box: contact at the middle of the inside part, pushed out along the box axis of
     * least depth, towards the side the capsule CENTRE is on (robust for thin boxes, where both faces are equally near) */
    real ps[3], ctr[3], nl[3] = {0, 0, 0};
    for (int k = 0; k < 3; k++) { ps[k] = a[k] + tb*dd[k]; ctr[k] = a[k] + (real)0.5*dd[k]; }
    int kb = 0; real bestd = s2[0] - (real)fabs((double)ps[0]);
    for (int k = 1; k < 3; k++) { real e = s2[k] - (real)fabs((double)ps[k]); if (e < bestd) { bestd = e; kb = k; } }
    real sg = ctr[kb] >= 0 ? (real)1 : (real)-1;
    nl[kb] = -sg;
    real depth = s2[kb] - sg*ps[kb];
    mat_vec(c->frame, m2, nl);
    c->dist = -depth - s1[0];
    for (int k = 0; k < 3; k++) { c->pos[k] = pt[k] + c->frame[k]*(s1[0] + c->dist*(real)0.5); c->frame[3 + k] = 0; }
    n = 1;
  }
  if (n == 0) return 0;
  for (int e = 0; e < 2 && n < 2; e++) {
    real te = (real)e;
    if ((real)fabs((double)(te - tb))*2*s1[1] < (real)1e-3*s1[0] + (real)1e-6) continue;   /* same point */
    const real* pe = e == 0 ? e0 : e1;
    if (col_sphere_box(c + n, margin, pe, rad, p2, m2, s2) && dot3(c[n].frame, c[0].frame) > (real)0.9) n++;
  }
  return n;
}

/* box-box: separating-axis test over the 15 candidate axes, then either a face contact (the incident face is
 * clipped against the side planes of the reference face; every clipped vertex below the reference face within
 * `margin` is a contact, up to 8) or one edge-edge contact.  Own formulation of the standard method (the
 * reference's dependency ports MuJoCo's mjc_BoxBox); face axes are preferred over edge axes by a small bias. */
static int col_box_box(rawcon* c, real margin, const real* p1, const real* m1, const real* s1, const real* p2, const real* m2, const real* s2) {
  real R[9], AR[9], t[3], dp[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
  matT_vec(t, m1, dp);
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) {
    real v = 0;
    for (int k = 0; k < 3; k++) v += m1[3*k + i]*m2[3*k + j];
    R[3*i + j] = v; AR[3*i + j] = (real)fabs((double)v) + (real)1e-6;
  }
  real best = (real)-1e30; int bestax = -1; real bestsign = 1;
  /* face axes of box 1 */
  for (int i = 0; i < 3; i++) {
    real rb = s2[0]*AR[3*i] + s2[1]*AR[3*i + 1] + s2[2]*AR[3*i + 2];
    real sep = (real)fabs((double)t[i]) - (s1[i] + rb);
    if (sep > margin) return 0;
    if (sep > best) { best = sep; bestax = i; bestsign = t[i] >= 0 ? (real)1 : (real)-1; }
  }
  /* face axes of box 2 */
  for (int j = 0; j < 3; j++) {
    real tj = t[0]*R[j] + t[1]*R[3 + j] + t[2]*R[6 + j];
    real ra = s1[0]*AR[j] + s1[1]*AR[3 + j] + s1[2]*AR[6 + j];
    real sep = (real)fabs((double)tj) - (ra + s2[j]);
    if (sep > margin) return 0;
    if (sep > best) { best = sep; bestax = 3 + j; bestsign = tj >= 0 ? (real)1 : (real)-1; }
  }
  /* edge-edge axes */
  real ebest = (real)-1e30; int ei = -1, ej = -1; real esign = 1, eaxis[3] = {0, 0, 0};
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) {
    int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
    /* L = A_i x B_j in box-1 coordinates */
    real L[3] = {0, 0, 0};
    L[i1] = -R[3*i2 + j]; L[i2] = R[3*i1 + j];
    real ln = norm3(L);
    if (ln < (real)1e-6) continue;   /* parallel edges: covered by the face axes */
    real ra = s1[i1]*(real)fabs((double)R[3*i2 + j]) + s1[i2]*(real)fabs((double)R[3*i1 + j]);
    real rb = s2[j1]*(real)fabs((double)R[3*i + j2]) + s2[j2]*(real)fabs((double)R[3*i + j1]);
    real tl = t[i2]*R[3*i1 + j] - t[i1]*R[3*i2 + j];
    real sep = ((real)fabs((double)tl) - (ra + rb))/ln;
    if (sep > margin) return 0;
    if (sep > ebest) { ebest = sep; ei = i; ej = j; esign = tl >= 0 ? (real)1 : (real)-1; for (int k = 0; k < 3; k++) eaxis[k] = L[k]/ln; }
  }
  if (ei >= 0 && ebest > best + (real)1e-3*(real)fabs((double)best) + (real)1e-5) {
    /* edge-edge contact: closest points of the two supporting edges */
    real nl[3] = {eaxis[0]*esign, eaxis[1]*esign, eaxis[2]*esign}, nw[3];   /* box1 -> box2, box-1 frame */
    mat_vec(nw, m1, nl);
    real pa[3], pb[3];
    for (int k = 0; k < 3; k++) { pa[k] = p1[k]; pb[k] = p2[k]; }
    for (int k = 0; k < 3; k++) if (k != ei) {
      real sg = nl[k] > 0 ? (real)1 : (real)-1;
      for (int q = 0; q < 3; q++) pa[q] += sg*s1[k]*m1[3*q + k];
    }
    for (int k = 0; k < 3; k++) if (k != ej) {
      real nb = nl[0]*R[k] + nl[1]*R[3 + k] + nl[2]*R[6 + k];
      real sg = nb > 0 ? (real)-1 : (real)1;
      for (int q = 0; q < 3; q++) pb[q] += sg*s2[k]*m2[3*q + k];
    }
    real ua[3] = {m1[ei], m1[3 + ei], m1[6 + ei]}, ub[3] = {m2[ej], m2[3 + ej], m2[6 + ej]};
    real w[3] = {pb[0] - pa[0], pb[1] - pa[1], pb[2] - pa[2]};
    real uaub = dot3(ua, ub), q1 = dot3(ua, w), q2 = -dot3(ub, w), den = 1 - uaub*uaub;
    real al = 0, be = 0;
    if (den > (real)1e-8) { al = (q1 + uaub*q2)/den; be = (uaub*q1 + q2)/den; }
    al = clipr(al, -s1[ei], s1[ei]); be = clipr(be, -s2[ej], s2[ej]);
    for (int k = 0; k < 3; k++) { pa[k] += al*ua[k]; pb[k] += be*ub[k]; }
    c->dist = ebest;
    for (int k = 0; k < 3; k++) { c->pos[k] = (real)0.5*(pa[k] + pb[k]); c->frame[k] = nw[k]; c->frame[3 + k] = 0; }
    return 1;
  }
  /* face contact */
  const real *pr, *mr, *sr, *pi, *mi, *si; int ax; real nref[3], flip;
  if (bestax < 3) { pr = p1; mr = m1; sr = s1; pi = p2; mi = m2; si = s2; ax = bestax; flip = 1; }
  else { pr = p2; mr = m2; sr = s2; pi = p1; mi = m1; si = s1; ax = bestax - 3; flip = -1; }
  /* outward normal of the reference face, pointing at the incident box */
  real sgn = bestsign*flip;
  for (int k = 0; k < 3; k++) nref[k] = sgn*mr[3*k + ax];
  /* incident face: the face of the incident box most anti-parallel to nref */
  int iax = 0; real imin = 2;
  real isg = 1;
  for (int k = 0; k < 3; k++) {
    real dcol = nref[0]*mi[k] + nref[1]*mi[3 + k] + nref[2]*mi[6 + k];
    if (-(real)fabs((double)dcol) < imin) { imin = -(real)fabs((double)dcol); iax = k; isg = dcol > 0 ? (real)-1 : (real)1; }
  }
  int u = (iax + 1) % 3, v = (iax + 2) % 3;
  real poly[16][3], tmpoly[16][3]; int np = 4;
  static const real su[4] = {1, -1, -1, 1}, sv[4] = {1, 1, -1, -1};
  for (int q = 0; q < 4; q++) for (int k = 0; k < 3; k++)
    poly[q][k] = pi[k] + isg*si[iax]*mi[3*k + iax] + su[q]*si[u]*mi[3*k + u] + sv[q]*si[v]*mi[3*k + v];
  /* clip against the four side planes of the reference face */
  int ru = (ax + 1) % 3, rv = (ax + 2) % 3;
  for (int side = 0; side < 4 && np > 0; side++) {
    int ka = side < 2 ? ru : rv; real sg = (side & 1) ? (real)-1 : (real)1;
    real pn[3] = {sg*mr[ka], sg*mr[3 + ka], sg*mr[6 + ka]}, lim = sr[ka];
    int nn = 0;
    for (int q = 0; q < np; q++) {
      const real* A = poly[q]; const real* B = poly[(q + 1) % np];
      real da = (A[0] - pr[0])*pn[0] + (A[1] - pr[1])*pn[1] + (A[2] - pr[2])*pn[2] - lim;
      real db = (B[0] - pr[0])*pn[0] + (B[1] - pr[1])*pn[1] + (B[2] - pr[2])*pn[2] - lim;
      if (da <= 0) { for (int k = 0; k < 3; k++) tmpoly[nn][k] = A[k]; nn++; }
      if ((da < 0 && db > 0) || (da > 0 && db < 0)) { real f = da/(da - db); for (int k = 0; k < 3; k++) tmpoly[nn][k] = A[k] + f*(B[k] - A[k]); nn++; }
    }
    np = nn > 16 ? 16 : nn;
    for (int q = 0; q < np; q++) for (int k = 0; k < 3; k++) poly[q][k] = tmpoly[q][k];
  }
  int n = 0;
  for (int q = 0; q < np && n < 8; q++) {
    real dist = (poly[q][0] - pr[0])*nref[0] + (poly[q][1] - pr[1])*nref[1] + (poly[q][2] - pr[2])*nref[2] - sr[ax];
    if (dist > margin) continue;
    c[n].dist = dist;
    for (int k = 0; k < 3; k++) { c[n].pos[k] = poly[q][k] - (real)0.5*dist*nref[k]; c[n].frame[k] = flip*nref[k]; c[n].frame[3 + k] = 0; }
    n++;
  }
  return n;
}

/* mj_contactParam */
#include "convex_oracle.h"

static void contact_param(const b2s_model_desc* m, int g1, int g2, int* condim, real* fri5, real* solref, real* solimp, real* margin, real* gap) {
  int p1 = m->geom_priority[g1], p2 = m->geom_priority[g2];
  real f[3];
  *margin = (real)fmax(m->geom_margin[g1], m->geom_margin[g2]);
  *gap = (real)fmax(m->geom_gap[g1], m->geom_gap[g2]);
  if (p1 != p2) {
    int g = p1 > p2 ? g1 : g2;
    *condim = m->geom_condim[g];
    for (int k = 0; k < 3; k++) f[k] = (real)m->geom_friction[3*g + k];
    for (int k = 0; k < 2; k++) solref[k] = (real)m->geom_solref[2*g + k];
    for (int k = 0; k < 5; k++) solimp[k] = (real)m->geom_solimp[5*g + k];
  } else {
    *condim = m->geom_condim[g1] > m->geom_condim[g2] ? m->geom_condim[g1] : m->geom_condim[g2];
    for (int k = 0; k < 3; k++) f[k] = (real)fmax(m->geom_friction[3*g1 + k], m->geom_friction[3*g2 + k]);
    real s1 = (real)m->geom_solmix[g1], s2 = (real)m->geom_solmix[g2], mix;
    if (s1 >= MINVAL && s2 >= MINVAL) mix = s1/(s1 + s2);
    else if (s1 < MINVAL && s2 < MINVAL) mix = (real)0.5;
    else if (s1 < MINVAL) mix = 0; else mix = 1;
    if (m->geom_solref[2*g1] > 0 && m->geom_solref[2*g2] > 0)
      for (int k = 0; k < 2; k++) solref[k] = mix*(real)m->geom_solref[2*g1 + k] + (1 - mix)*(real)m->geom_solref[2*g2 + k];
    else
      for (int k = 0; k < 2; k++) solref[k] = (real)fmin(m->geom_solref[2*g1 + k], m->geom_solref[2*g2 + k]);
    for (int k = 0; k < 5; k++) solimp[k] = mix*(real)m->geom_solimp[5*g1 + k] + (1 - mix)*(real)m->geom_solimp[5*g2 + k];
  }
  fri5[0] = fri5[1] = f[0] > MINMU ? f[0] : MINMU;
  fri5[2] = f[1] > MINMU ? f[1] : MINMU;
  fri5[3] = fri5[4] = f[2] > MINMU ? f[2] : MINMU;
}

/* body-pair and geom-pair filters of mj_collision (static pair list in mjwarp) */
static int pair_filtered(const b2s_model_desc* m, int g1, int g2) {
  int b1 = m->geom_bodyid[g1], b2 = m->geom_bodyid[g2];
  if (!((m->geom_contype[g1] & m->geom_conaffinity[g2]) || (m->geom_contype[g2] & m->geom_conaffinity[g1]))) return 1;
  int w1 = m->body_weldid[b1], w2 = m->body_weldid[b2];
  if (w1 == w2) return 1;
  int wp1 = m->body_weldid[m->body_parentid[w1]], wp2 = m->body_weldid[m->body_parentid[w2]];
  if (!(m->opt_disableflags & B2S_DSBL_FILTERPARENT) && w1 != 0 && w2 != 0 && (w1 == wp2 || w2 == wp1)) return 1;
  for (int e = 0; e < m->nexclude; e++) {
    int e1 = m->exclude_body1[e], e2 = m->exclude_body2[e];
    if ((e1 == b1 && e2 == b2) || (e1 == b2 && e2 == b1)) return 1;
  }
  return 0;
}

void orc_collision(const b2s_model_desc* m, orc_data* d) {
  int ncon = 0;
  d->ncon[0] = 0;
  if (m->opt_disableflags & (B2S_DSBL_CONSTRAINT | B2S_DSBL_CONTACT)) return;
  /* pairs are visited in (ga, gb) order, ga < gb; two geoms on static (world-welded) bodies never collide, so for a
   * static ga only the dynamic gb are visited (same order, far fewer iterations with thousands of terrain geoms) */
  int ndyn = 0;
  int* dyn = (int*)malloc(sizeof(int)*(size_t)(m->ngeom > 0 ? m->ngeom : 1));
  for (int g = 0; g < m->ngeom; g++) if (m->body_weldid[m->geom_bodyid[g]] != 0) dyn[ndyn++] = g;
  for (int ga = 0; ga < m->ngeom; ga++) {
    const int ga_static = m->body_weldid[m->geom_bodyid[ga]] == 0;
    int k0 = 0;
    if (ga_static) { while (k0 < ndyn && dyn[k0] <= ga) k0++; }
    const int cnt = ga_static ? ndyn - k0 : m->ngeom - ga - 1;
    for (int kk = 0; kk < cnt; kk++) {
    const int gb = ga_static ? dyn[k0 + kk] : ga + 1 + kk;
    if (pair_filtered(m, ga, gb)) continue;
    int g1 = ga, g2 = gb;
    if (m->geom_type[g1] > m->geom_type[g2]) { g1 = gb; g2 = ga; }
    int t1 = m->geom_type[g1], t2 = m->geom_type[g2];
    int condim; real fri[5], solref[2], solimp[5], margin, gap;
    const real* p1 = d->geom_xpos + 3*g1; const real* p2 = d->geom_xpos + 3*g2;
    const real* m1 = d->geom_xmat + 9*g1; const real* m2 = d->geom_xmat + 9*g2;
    real s1[3] = {(real)m->geom_size[3*g1], (real)m->geom_size[3*g1+1], (real)m->geom_size[3*g1+2]};
    real s2[3] = {(real)m->geom_size[3*g2], (real)m->geom_size[3*g2+1], (real)m->geom_size[3*g2+2]};
    margin = (real)fmax(m->geom_margin[g1], m->geom_margin[g2]);
    /* bounding rejection (exact narrowphase follows, so this never changes the result) */
    if (t1 == B2S_GEOM_PLANE) {
      real n[3] = {m1[2], m1[5], m1[8]}, dif[3] = {p2[0]-p1[0], p2[1]-p1[1], p2[2]-p1[2]};
      if (dot3(dif, n) > (real)m->geom_rbound[g2] + margin) continue;
    } else if (t1 == B2S_GEOM_HFIELD) {
      /* the sub-grid test inside col_hfield is the rejection */
    } else {
      real dif[3] = {p2[0]-p1[0], p2[1]-p1[1], p2[2]-p1[2]};
      real bound = (real)m->geom_rbound[g1] + (real)m->geom_rbound[g2] + margin;
      if (dot3(dif, dif) > bound*bound) continue;
      if (t2 == B2S_GEOM_BOX) {   /* bounding sphere of geom 1 against the oriented box 2 */
        real loc[3], ndif[3] = {-dif[0], -dif[1], -dif[2]}, d2 = 0, rb = (real)m->geom_rbound[g1] + margin;
        matT_vec(loc, m2, ndif);
        for (int k = 0; k < 3; k++) { real e = (real)fabs((double)loc[k]) - s2[k]; if (e > 0) d2 += e*e; }
        if (d2 > rb*rb) continue;
      }
    }
    contact_param(m, g1, g2, &condim, fri, solref, solimp, &margin, &gap);
    rawcon rc[8];
    int n = 0;
    if (t1 == B2S_GEOM_PLANE && t2 == B2S_GEOM_SPHERE) { real nn[3] = {m1[2], m1[5], m1[8]}; n = raw_plane_sphere(rc, margin, p1, nn, p2, s2[0]); }
    else if (t1 == B2S_GEOM_PLANE && t2 == B2S_GEOM_CAPSULE) n = col_plane_capsule(rc, margin, p1, m1, p2, m2, s2);
    else if (t1 == B2S_GEOM_PLANE && t2 == B2S_GEOM_BOX) n = col_plane_box(rc, margin, p1, m1, p2, m2, s2);
    else if (t1 == B2S_GEOM_SPHERE && t2 == B2S_GEOM_SPHERE) n = raw_sphere_sphere(rc, margin, p1, s1[0], p2, s2[0]);
    else if (t1 == B2S_GEOM_SPHERE && t2 == B2S_GEOM_CAPSULE) n = col_sphere_capsule(rc, margin, p1, s1, p2, m2, s2);
    else if (t1 == B2S_GEOM_CAPSULE && t2 == B2S_GEOM_CAPSULE) n = col_capsule_capsule(rc, margin, p1, m1, s1, p2, m2, s2);
    else if (t1 == B2S_GEOM_SPHERE && t2 == B2S_GEOM_BOX) n = col_sphere_box(rc, margin, p1, s1, p2, m2, s2);
    else if (t1 == B2S_GEOM_CAPSULE && t2 == B2S_GEOM_BOX) n = col_capsule_box(rc, margin, p1, m1, s1, p2, m2, s2);
    else if (t1 == B2S_GEOM_BOX && t2 == B2S_GEOM_BOX) n = col_box_box(rc, margin, p1, m1, s1, p2, m2, s2);
    else if (t1 == B2S_GEOM_PLANE && t2 == B2S_GEOM_MESH) {
      int id = m->geom_dataid[g2];
      n = col_plane_mesh(rc, margin, p1, m1, p2, m2, m->mesh_vert + 3*m->mesh_vertadr[id], m->mesh_vertnum[id]);
    } else if (t1 == B2S_GEOM_PLANE && t2 == B2S_GEOM_CYLINDER) n = col_plane_cylinder(rc, margin, p1, m1, p2, m2, s2);
    else if (t1 == B2S_GEOM_PLANE && t2 == B2S_GEOM_ELLIPSOID) { cvx B; cvx_from_geom(m, d, g2, &B); n = col_plane_ellipsoid(rc, margin, p1, m1, &B); }
    else if (t1 == B2S_GEOM_HFIELD && t2 >= B2S_GEOM_SPHERE) {
      cvx B; cvx_from_geom(m, d, g2, &B);
      n = col_hfield(rc, 8, margin, m, m->geom_dataid[g1], p1, m1, &B, (real)m->geom_rbound[g2]);
    } else if (t1 >= B2S_GEOM_SPHERE && (t2 == B2S_GEOM_MESH || t1 == B2S_GEOM_ELLIPSOID || t1 == B2S_GEOM_CYLINDER ||
                                         t2 == B2S_GEOM_ELLIPSOID || t2 == B2S_GEOM_CYLINDER)) {
      /* every pair with a mesh, a cylinder or an ellipsoid goes through GJK / EPA (one contact) */
      cvx A, B; cvx_from_geom(m, d, g1, &A); cvx_from_geom(m, d, g2, &B);
      n = col_convex(rc, margin, &A, &B);
    }
    else n = 0; /* unsupported pair types are rejected at model creation by the product */
    for (int i = 0; i < n && ncon < d->nconmax; i++) {
      int c = ncon++;
      make_frame(rc[i].frame);
      d->con_dist[c] = rc[i].dist;
      for (int k = 0; k < 3; k++) d->con_pos[3*c + k] = rc[i].pos[k];
      for (int k = 0; k < 9; k++) d->con_frame[9*c + k] = rc[i].frame[k];
      d->con_includemargin[c] = margin - gap;
      for (int k = 0; k < 5; k++) { d->con_friction[5*c + k] = fri[k]; d->con_solimp[5*c + k] = solimp[k]; }
      d->con_solref[2*c] = solref[0]; d->con_solref[2*c + 1] = solref[1];
      d->con_dim[c] = condim; d->con_geom[2*c] = g1; d->con_geom[2*c + 1] = g2; d->con_efc_address[c] = -1;
    }
    }
  }
  free(dyn);
  d->ncon[0] = ncon;
}

/* ------------------------------------------------------------- Jacobians */
/* mj_jac: translational (jp) and rotational (jr) 3 x nv Jacobians of `point` attached to `body` */
static void jac_point(const b2s_model_desc* m, const orc_data* d, int body, const real* point, real* jp, real* jr) {
  int nv = m->nv;
  for (int i = 0; i < 3*nv; i++) { jp[i] = 0; if (jr) jr[i] = 0; }
  while (body > 0 && m->body_dofnum[body] == 0) body = m->body_parentid[body];
  if (body == 0) return;
  real off[3];
  for (int k = 0; k < 3; k++) off[k] = point[k] - d->subtree_com[3*m->body_rootid[body] + k];
  int i = m->body_dofadr[body] + m->body_dofnum[body] - 1;
  while (i >= 0) {
    const real* c = d->cdof + 6*i;
    real t[3];
    cross3(t, c, off);
    for (int k = 0; k < 3; k++) { jp[k*nv + i] = c[3 + k] + t[k]; if (jr) jr[k*nv + i] = c[k]; }
    i = m->dof_parentid[i];
  }
}

/* ------------------------------------------------------- make_constraint */
/* engine_core_constraint.c getimpedance */
static real impedance(const real* solimp, real pos, real margin) {
  real d0 = solimp[0], dw = solimp[1], width = solimp[2], mid = solimp[3], power = solimp[4];
  if (d0 == dw || width <= MINVAL) return (real)0.5*(d0 + dw);
  real x = (pos - margin)/width;
  if (x < 0) x = -x;
  if (x >= 1) return dw;
  if (x <= 0) return d0;
  real y;
  if (power == 1) y = x;
  else if (x <= mid) y = (real)(pow((double)x, (double)power)/pow((double)mid, (double)power - 1));
  else y = (real)(1 - pow(1 - (double)x, (double)power)/pow(1 - (double)mid, (double)power - 1));
  return d0 + y*(dw - d0);
}
/* mj_makeImpedance row finish: D from impedance and diagApprox, aref from solref (K, B) */
static void finish_row(const b2s_model_desc* m, orc_data* d, int r, const real* solref, const real* solimp, real diagApprox, real Rscale) {
  real pos = d->efc_pos[r], margin = d->efc_margin[r];
  real imp = impedance(solimp, pos, margin);
  real R = (1 - imp)*diagApprox/imp*Rscale;
  if (R < MINVAL) R = MINVAL;
  d->efc_D[r] = 1/R;
  real K, B, dmax = solimp[1];
  if (solref[0] > 0) {
    real tc = solref[0], dr = solref[1];
    if (!(m->opt_disableflags & B2S_DSBL_REFSAFE) && tc < 2*(real)m->opt_timestep) tc = 2*(real)m->opt_timestep;
    real kd = dmax*dmax*tc*tc*dr*dr; if (kd < MINVAL) kd = MINVAL;
    real bd = dmax*tc; if (bd < MINVAL) bd = MINVAL;
    K = 1/kd; B = 2/bd;
  } else {
    real kd = dmax*dmax; if (kd < MINVAL) kd = MINVAL;
    real bd = dmax; if (bd < MINVAL) bd = MINVAL;
    K = -solref[0]/kd; B = -solref[1]/bd;
  }
  real vel = 0;
  for (int i = 0; i < m->nv; i++) vel += d->efc_J[r*m->nv + i]*d->qvel[i];
  d->efc_vel[r] = vel;
  d->efc_aref[r] = -B*vel - K*imp*(pos - margin);
  d->efc_force[r] = 0;
}

/* aref of a row without position term: -B vel with B from solref as in finish_row (efc_vel[r] already holds J qvel) */
static real friction_row_aref(const b2s_model_desc* m, const orc_data* d, int r, const real* solref, const real* solimp) {
  real B, dmax = solimp[1];
  if (solref[0] > 0) {
    real tc = solref[0];
    if (!(m->opt_disableflags & B2S_DSBL_REFSAFE) && tc < 2*(real)m->opt_timestep) tc = 2*(real)m->opt_timestep;
    real bd = dmax*tc; if (bd < MINVAL) bd = MINVAL;
    B = 2/bd;
  } else { real bd = dmax; if (bd < MINVAL) bd = MINVAL; B = -solref[1]/bd; }
  return -B*d->efc_vel[r];
}

void orc_make_constraint(const b2s_model_desc* m, orc_data* d) {
  int nv = m->nv, nefc = 0;
  d->nefc[0] = 0; d->nl[0] = 0; d->nf[0] = 0;
  if (m->opt_disableflags & B2S_DSBL_CONSTRAINT) return;
  /* dof friction loss (mj_instantiateFriction): one row per dof with frictionloss > 0, J = e_dof, pos = margin = 0, default
   * solreffriction / solimpfriction; the rows come before the limits (MuJoCo order: equality, friction, limit, contact) */
  if (!(m->opt_disableflags & B2S_DSBL_FRICTIONLOSS))
    for (int i = 0; i < nv; i++) {
      if (!(m->dof_frictionloss[i] > 0) || nefc >= d->njmax) continue;
      int r = nefc++;
      for (int k = 0; k < nv; k++) d->efc_J[r*nv + k] = 0;
      d->efc_J[r*nv + i] = 1;
      d->efc_pos[r] = 0; d->efc_margin[r] = 0; d->efc_frictionloss[r] = (real)m->dof_frictionloss[i];
      d->efc_type[r] = ORC_EFC_FRICTION_DOF; d->efc_id[r] = i;
      real sr[2] = {(real)0.02, 1}, si[5] = {(real)0.9, (real)0.95, (real)0.001, (real)0.5, 2};
      finish_row(m, d, r, sr, si, (real)m->dof_invweight0[i], 1);
    }
  d->nf[0] = nefc;
  /* joint limits: lower side then upper side (mj_instantiateLimit) */
  if (!(m->opt_disableflags & B2S_DSBL_LIMIT))
    for (int j = 0; j < m->njnt; j++) {
      if (!m->jnt_limited[j]) continue;
      int t = m->jnt_type[j];
      if (t != B2S_JNT_HINGE && t != B2S_JNT_SLIDE) continue;
      real q = d->qpos[m->jnt_qposadr[j]], margin = (real)m->jnt_margin[j];
      for (int side = -1; side <= 1; side += 2) {
        real dist = side*((real)m->jnt_range[2*j + (side + 1)/2] - q);
        if (dist >= margin || nefc >= d->njmax) continue;
        int r = nefc++;
        for (int i = 0; i < nv; i++) d->efc_J[r*nv + i] = 0;
        d->efc_J[r*nv + m->jnt_dofadr[j]] = (real)(-side);
        d->efc_pos[r] = dist; d->efc_margin[r] = margin;
        d->efc_type[r] = ORC_EFC_LIMIT; d->efc_id[r] = j;
        real sr[2] = {(real)m->jnt_solref[2*j], (real)m->jnt_solref[2*j + 1]}, si[5];
        for (int k = 0; k < 5; k++) si[k] = (real)m->jnt_solimp[5*j + k];
        finish_row(m, d, r, sr, si, (real)m->dof_invweight0[m->jnt_dofadr[j]], 1);
      }
    }
  d->nl[0] = nefc;
  /* contacts (mj_instantiateContact, pyramidal cone) */
  real* jp1 = (real*)malloc(sizeof(real)*nv*12); real* jr1 = jp1 + 3*nv; real* jp2 = jr1 + 3*nv; real* jr2 = jp2 + 3*nv;
  int full = 0;  /* once the row budget overflows, this and all later contacts are dropped */
  for (int c = 0; c < d->ncon[0]; c++) {
    d->con_efc_address[c] = -1;
    if (m->opt_disableflags & B2S_DSBL_CONTACT) continue;
    if (d->con_dist[c] >= d->con_includemargin[c]) continue;
    int dim = d->con_dim[c];
    const int elliptic = m->opt_cone == B2S_CONE_ELLIPTIC && dim > 1;
    int nrow = dim == 1 ? 1 : (elliptic ? dim : 2*(dim - 1));
    if (full || nefc + nrow > d->njmax) { full = 1; continue; }
    int g1 = d->con_geom[2*c], g2 = d->con_geom[2*c + 1];
    int b1 = m->geom_bodyid[g1], b2 = m->geom_bodyid[g2];
    const real* fr = d->con_frame + 9*c;
    jac_point(m, d, b1, d->con_pos + 3*c, jp1, jr1);
    jac_point(m, d, b2, d->con_pos + 3*c, jp2, jr2);
    /* rows of the difference Jacobian expressed in the contact frame: 3 translational, 3 rotational */
    real Jc[6*NV_MAX];
    for (int a = 0; a < 3; a++) for (int i = 0; i < nv; i++) {
      real st = 0, sr = 0;
      for (int k = 0; k < 3; k++) { st += fr[3*a + k]*(jp2[k*nv + i] - jp1[k*nv + i]); sr += fr[3*a + k]*(jr2[k*nv + i] - jr1[k*nv + i]); }
      Jc[a*nv + i] = st; Jc[(3 + a)*nv + i] = sr;
    }
    real tran = (real)m->body_invweight0[2*b1] + (real)m->body_invweight0[2*b2];
    real sr[2] = {d->con_solref[2*c], d->con_solref[2*c + 1]};
    d->con_efc_address[c] = nefc;
    if (dim == 1) {
      int r = nefc++;
      for (int i = 0; i < nv; i++) d->efc_J[r*nv + i] = Jc[i];
      d->efc_pos[r] = d->con_dist[c]; d->efc_margin[r] = d->con_includemargin[c];
      d->efc_type[r] = ORC_EFC_CONTACT_FRICTIONLESS; d->efc_id[r] = c;
      finish_row(m, d, r, sr, d->con_solimp + 5*c, tran, 1);
    } else if (elliptic) {
      /* elliptic cone (mj_instantiateContact + mj_makeImpedance): rows = the contact-frame axes (normal, tangents, then
       * torsional / rolling); the friction rows have pos = 0, share the normal row's impedance, and are regularised by
       * R_1 = R_n / impratio, R_j = R_1 mu_1^2 / mu_j^2; their aref is -B vel */
      real rot = (real)m->body_invweight0[2*b1 + 1] + (real)m->body_invweight0[2*b2 + 1];
      int r0 = nefc;
      for (int k = 0; k < dim; k++) {
        int r = nefc++;
        for (int i = 0; i < nv; i++) d->efc_J[r*nv + i] = Jc[k*nv + i];
        d->efc_pos[r] = d->con_dist[c]; d->efc_margin[r] = d->con_includemargin[c];
        d->efc_type[r] = ORC_EFC_CONTACT_ELLIPTIC; d->efc_id[r] = c;
        finish_row(m, d, r, sr, d->con_solimp + 5*c, k < 3 ? tran : rot, 1);
        if (k > 0) {
          real mu1 = d->con_friction[5*c], muk = d->con_friction[5*c + k - 1];
          real Rn = 1/d->efc_D[r0], R1 = Rn/(real)m->opt_impratio, Rk = R1*mu1*mu1/(muk*muk);
          if (Rk < MINVAL) Rk = MINVAL;
          d->efc_D[r] = 1/Rk;
          /* friction rows: no position term (K = 0) */
          d->efc_aref[r] = friction_row_aref(m, d, r, sr, d->con_solimp + 5*c);
          d->efc_pos[r] = 0; d->efc_margin[r] = 0;
        }
      }
    } else {
      real mu0 = d->con_friction[5*c];
      /* diagApprox of the first pyramid edge and the common pyramid regulariser Rpy = 2 mu^2 R / impratio */
      real diag = tran + mu0*mu0*tran;
      real scale = 2*mu0*mu0/(real)m->opt_impratio;
      for (int k = 0; k < nrow; k++) {
        int r = nefc++;
        int ax = 1 + k/2;
        real mu = d->con_friction[5*c + k/2], sgn = (k & 1) ? (real)-1 : (real)1;
        for (int i = 0; i < nv; i++) d->efc_J[r*nv + i] = Jc[i] + sgn*mu*Jc[ax*nv + i];
        d->efc_pos[r] = d->con_dist[c]; d->efc_margin[r] = d->con_includemargin[c];
        d->efc_type[r] = ORC_EFC_CONTACT_PYRAMIDAL; d->efc_id[r] = c;
        finish_row(m, d, r, sr, d->con_solimp + 5*c, diag, scale);
      }
    }
  }
  free(jp1);
  d->nefc[0] = nefc;
}

/* ----------------------------------------------------------- fwd_velocity */
void orc_fwd_velocity(const b2s_model_desc* m, orc_data* d) {
  int nb = m->nbody, nv = m->nv;
  /* mj_comVel */
  for (int k = 0; k < 6; k++) d->cvel[k] = 0;
  for (int i = 1; i < nb; i++) {
    real v[6];
    for (int k = 0; k < 6; k++) v[k] = d->cvel[6*m->body_parentid[i] + k];
    for (int jj = 0; jj < m->body_jntnum[i]; jj++) {
      int j = m->body_jntadr[i] + jj, da = m->jnt_dofadr[j], t = m->jnt_type[j];
      if (t == B2S_JNT_FREE) {
        for (int a = 0; a < 3; a++) {
          for (int k = 0; k < 6; k++) d->cdof_dot[6*(da + a) + k] = 0;
          for (int k = 0; k < 6; k++) v[k] += d->cdof[6*(da + a) + k]*d->qvel[da + a];
        }
        da += 3;
      }
      if (t == B2S_JNT_FREE || t == B2S_JNT_BALL) {
        for (int a = 0; a < 3; a++) cross_motion(d->cdof_dot + 6*(da + a), v, d->cdof + 6*(da + a));
        for (int a = 0; a < 3; a++) for (int k = 0; k < 6; k++) v[k] += d->cdof[6*(da + a) + k]*d->qvel[da + a];
      } else {
        cross_motion(d->cdof_dot + 6*da, v, d->cdof + 6*da);
        for (int k = 0; k < 6; k++) v[k] += d->cdof[6*da + k]*d->qvel[da];
      }
    }
    for (int k = 0; k < 6; k++) d->cvel[6*i + k] = v[k];
  }
  /* fixed tendons (mj_tendon, mjWRAP_JOINT only): length = sum coef*qpos, velocity = ten_J . qvel */
  for (int t = 0; t < m->ntendon; t++) {
    real len = 0, vel = 0;
    for (int k = m->tendon_adr[t]; k < m->tendon_adr[t] + m->tendon_num[t]; k++) {
      int j = m->wrap_objid[k];
      len += (real)m->wrap_prm[k]*d->qpos[m->jnt_qposadr[j]];
      vel += (real)m->wrap_prm[k]*d->qvel[m->jnt_dofadr[j]];
    }
    d->ten_length[t] = len; d->ten_velocity[t] = vel;
  }
  /* actuator length / velocity (joint or tendon transmission; mj_transmission + mj_fwdVelocity) */
  for (int u = 0; u < m->nu; u++) {
    int j = m->actuator_trnid[u];
    if (m->actuator_trntype[u] == B2S_TRN_TENDON) {
      d->actuator_length[u] = (real)m->actuator_gear[u]*d->ten_length[j];
      d->actuator_velocity[u] = (real)m->actuator_gear[u]*d->ten_velocity[j];
    } else {
      d->actuator_length[u] = (real)m->actuator_gear[u]*d->qpos[m->jnt_qposadr[j]];
      d->actuator_velocity[u] = (real)m->actuator_gear[u]*d->qvel[m->jnt_dofadr[j]];
    }
  }
  /* mj_passive: joint springs and dampers, tendon springs (deadband lengthspring) and dampers */
  for (int i = 0; i < nv; i++) d->qfrc_passive[i] = 0;
  if (!(m->opt_disableflags & B2S_DSBL_PASSIVE)) {
    for (int j = 0; j < m->njnt; j++) {
      int t = m->jnt_type[j];
      real k = (real)m->jnt_stiffness[j];
      if (k == 0 || (t != B2S_JNT_HINGE && t != B2S_JNT_SLIDE)) continue;
      int qa = m->jnt_qposadr[j];
      d->qfrc_passive[m->jnt_dofadr[j]] -= k*(d->qpos[qa] - (real)m->qpos_spring[qa]);
    }
    for (int i = 0; i < nv; i++) d->qfrc_passive[i] -= (real)m->dof_damping[i]*d->qvel[i];
    for (int t = 0; t < m->ntendon; t++) {
      real k = (real)m->tendon_stiffness[t], b = (real)m->tendon_damping[t], f = 0;
      if (k == 0 && b == 0) continue;
      real lo = (real)m->tendon_lengthspring[2*t], hi = (real)m->tendon_lengthspring[2*t + 1];
      if (d->ten_length[t] > hi) f = k*(hi - d->ten_length[t]);
      else if (d->ten_length[t] < lo) f = k*(lo - d->ten_length[t]);
      f -= b*d->ten_velocity[t];
      for (int w = m->tendon_adr[t]; w < m->tendon_adr[t] + m->tendon_num[t]; w++)
        d->qfrc_passive[m->jnt_dofadr[m->wrap_objid[w]]] += (real)m->wrap_prm[w]*f;
    }
  }
  /* mj_rne (flg_acc = 0): Coriolis, centrifugal and gravity */
  real* cfrc = (real*)malloc(sizeof(real)*6*nb);
  real* cacc = (real*)malloc(sizeof(real)*6*nb);
  int grav = !(m->opt_disableflags & B2S_DSBL_GRAVITY);
  for (int k = 0; k < 3; k++) { cacc[k] = 0; cacc[3 + k] = grav ? -(real)m->opt_gravity[k] : 0; }
  for (int k = 0; k < 6; k++) cfrc[k] = 0;
  for (int i = 1; i < nb; i++) {
    real* a = cacc + 6*i;
    for (int k = 0; k < 6; k++) a[k] = cacc[6*m->body_parentid[i] + k];
    for (int dd = 0; dd < m->body_dofnum[i]; dd++) {
      int da = m->body_dofadr[i] + dd;
      for (int k = 0; k < 6; k++) a[k] += d->cdof_dot[6*da + k]*d->qvel[da];
    }
    real t0[6], t1[6];
    mul_inert_vec(t0, d->cinert + 10*i, d->cvel + 6*i);
    cross_force(t1, d->cvel + 6*i, t0);
    mul_inert_vec(t0, d->cinert + 10*i, a);
    for (int k = 0; k < 6; k++) cfrc[6*i + k] = t0[k] + t1[k];
  }
  for (int i = nb - 1; i > 0; i--) {
    int p = m->body_parentid[i];
    for (int k = 0; k < 6; k++) cfrc[6*p + k] += cfrc[6*i + k];
  }
  for (int i = 0; i < nv; i++) {
    real s = 0;
    for (int k = 0; k < 6; k++) s += d->cdof[6*i + k]*cfrc[6*m->dof_bodyid[i] + k];
    d->qfrc_bias[i] = s;
  }
  free(cfrc); free(cacc);
}

/* ---------------------------------------------------------- fwd_actuation */
void orc_fwd_actuation(const b2s_model_desc* m, orc_data* d) {
  for (int i = 0; i < m->nv; i++) d->qfrc_actuator[i] = 0;
  for (int u = 0; u < m->nu; u++) {
    d->actuator_force[u] = 0;
    if (m->opt_disableflags & B2S_DSBL_ACTUATION) continue;
    real c = d->ctrl[u];
    if (m->actuator_ctrllimited[u] && !(m->opt_disableflags & B2S_DSBL_CLAMPCTRL))
      c = clipr(c, (real)m->actuator_ctrlrange[2*u], (real)m->actuator_ctrlrange[2*u + 1]);
    real gain = (real)m->actuator_gainprm[3*u];
    if (m->actuator_gaintype[u] == B2S_GAIN_AFFINE)
      gain += (real)m->actuator_gainprm[3*u + 1]*d->actuator_length[u] + (real)m->actuator_gainprm[3*u + 2]*d->actuator_velocity[u];
    real bias = 0;
    if (m->actuator_biastype[u] == B2S_BIAS_AFFINE)
      bias = (real)m->actuator_biasprm[3*u] + (real)m->actuator_biasprm[3*u + 1]*d->actuator_length[u]
           + (real)m->actuator_biasprm[3*u + 2]*d->actuator_velocity[u];
    real f = gain*c + bias;
    if (m->actuator_forcelimited[u]) f = clipr(f, (real)m->actuator_forcerange[2*u], (real)m->actuator_forcerange[2*u + 1]);
    d->actuator_force[u] = f;
    if (m->actuator_trntype[u] == B2S_TRN_TENDON) {   /* moment = gear * ten_J row */
      int t = m->actuator_trnid[u];
      for (int w = m->tendon_adr[t]; w < m->tendon_adr[t] + m->tendon_num[t]; w++)
        d->qfrc_actuator[m->jnt_dofadr[m->wrap_objid[w]]] += (real)m->actuator_gear[u]*(real)m->wrap_prm[w]*f;
    } else d->qfrc_actuator[m->jnt_dofadr[m->actuator_trnid[u]]] += (real)m->actuator_gear[u]*f;
  }
}

/* ------------------------------------------------------- fwd_acceleration */
void orc_fwd_acceleration(const b2s_model_desc* m, orc_data* d) {
  int nv = m->nv;
  real* jp = (real*)malloc(sizeof(real)*6*nv); real* jr = jp + 3*nv;
  for (int i = 0; i < nv; i++)
    d->qfrc_smooth[i] = d->qfrc_passive[i] - d->qfrc_bias[i] + d->qfrc_applied[i] + d->qfrc_actuator[i];
  /* mj_xfrcAccumulate: Cartesian wrench applied at the body's inertial frame origin */
  for (int b = 1; b < m->nbody; b++) {
    const real* x = d->xfrc_applied + 6*b;
    if (x[0] == 0 && x[1] == 0 && x[2] == 0 && x[3] == 0 && x[4] == 0 && x[5] == 0) continue;
    jac_point(m, d, b, d->xipos + 3*b, jp, jr);
    for (int i = 0; i < nv; i++)
      for (int k = 0; k < 3; k++) d->qfrc_smooth[i] += jp[k*nv + i]*x[k] + jr[k*nv + i]*x[3 + k];
  }
  free(jp);
  chol_solve(d->qacc_smooth, d->qL, d->qfrc_smooth, nv);
}

/* ----------------------------------------------------------------- solve */
typedef struct {
  const b2s_model_desc* m; orc_data* d;
  int nv, nefc;
  real *Ma, *Jaref, *grad, *Mgrad, *search, *Mv, *Jv, *H, *HL, *quad;
  real *gradold, *Mgradold;   /* CG: previous gradient and M^-1 gradient (Polak-Ribiere) */
  int cg, have_ml;
  int nf;                     /* rows [0, nf) are friction-loss rows: quadratic inside |Jaref| < R f, linear outside */
  int* active;
  real cost, gauss, quadGauss[3];
} solctx;

/* One elliptic contact (rows r .. r+dim-1) at residual `jar` (mj_constraintUpdate, elliptic branch): U = jar scaled by
 * (mu, friction), N = U_0, T = |U_1..|; zone 0 top (N >= mu T: no force), 1 bottom (mu N + T <= 0: every row quadratic),
 * 2 middle (cost 1/2 Dm (N - mu T)^2).  mu is the regularised cone slope friction_1 sqrt(R_1 / R_n). */
typedef struct { int dim, zone; real mu, Dm, N, T, U[6], fri[6]; } conezone;
static void cone_eval(const orc_data* d, int r, const real* jar, conezone* z) {
  int con = d->efc_id[r], dim = d->con_dim[con];
  z->dim = dim;
  z->mu = d->con_friction[5*con]*(real)sqrt((double)(d->efc_D[r]/d->efc_D[r + 1]));
  z->fri[0] = z->mu;
  for (int j = 1; j < dim; j++) z->fri[j] = d->con_friction[5*con + j - 1];
  real tt = 0;
  for (int k = 0; k < dim; k++) { z->U[k] = jar[r + k]*z->fri[k]; if (k) tt += z->U[k]*z->U[k]; }
  z->N = z->U[0]; z->T = (real)sqrt((double)tt);
  real den = z->mu*z->mu*(1 + z->mu*z->mu);
  z->Dm = d->efc_D[r]/(den > MINVAL ? den : MINVAL);
  if (z->N >= z->mu*z->T || (z->T <= 0 && z->N >= 0)) z->zone = 0;
  else if (z->mu*z->N + z->T <= 0 || (z->T <= 0 && z->N < 0)) z->zone = 1;
  else z->zone = 2;
}
/* mj_constraintUpdate: forces, active set and cost s(Jaref) for limit / contact rows, plus the Gauss term */
static void sol_update_constraint(solctx* c) {
  orc_data* d = c->d; int nv = c->nv;
  real cost = 0;
  for (int r = 0; r < c->nefc; r++) {
    real jar = c->Jaref[r], D = d->efc_D[r];
    if (r < c->nf) {
      /* friction loss (mj_constraintUpdate): force saturates at +-f outside the quadratic zone |jar| < R f */
      real f = d->efc_frictionloss[r], rf = f/D;
      if (jar <= -rf) { d->efc_force[r] = f; c->active[r] = 0; cost += -f*((real)0.5*rf + jar); }
      else if (jar >= rf) { d->efc_force[r] = -f; c->active[r] = 0; cost += -f*((real)0.5*rf - jar); }
      else { d->efc_force[r] = -D*jar; c->active[r] = 1; cost += (real)0.5*D*jar*jar; }
      continue;
    }
    if (d->efc_type[r] == ORC_EFC_CONTACT_ELLIPTIC) {
      /* active: 1 = row in the quadratic Hessian, 2 = first row of a middle-zone contact (cone Hessian block) */
      conezone z; cone_eval(d, r, c->Jaref, &z);
      for (int k = 0; k < z.dim; k++) { d->efc_force[r + k] = 0; c->active[r + k] = 0; }
      if (z.zone == 1) for (int k = 0; k < z.dim; k++) {
        real x = c->Jaref[r + k], Dk = d->efc_D[r + k];
        d->efc_force[r + k] = -Dk*x; c->active[r + k] = 1; cost += (real)0.5*Dk*x*x;
      } else if (z.zone == 2) {
        real nmt = z.N - z.mu*z.T;
        cost += (real)0.5*z.Dm*nmt*nmt;
        d->efc_force[r] = -z.Dm*nmt*z.mu;
        for (int j = 1; j < z.dim; j++) d->efc_force[r + j] = -d->efc_force[r]/z.T*z.U[j]*z.fri[j];
        c->active[r] = 2;
      }
      r += z.dim - 1;
      continue;
    }
    if (jar < 0) { d->efc_force[r] = -D*jar; c->active[r] = 1; cost += (real)0.5*D*jar*jar; }
    else { d->efc_force[r] = 0; c->active[r] = 0; }
  }
  for (int i = 0; i < nv; i++) {
    real s = 0;
    for (int r = 0; r < c->nefc; r++) s += d->efc_J[r*nv + i]*d->efc_force[r];
    d->qfrc_constraint[i] = s;
  }
  real g = 0;
  for (int i = 0; i < nv; i++) g += (c->Ma[i] - d->qfrc_smooth[i])*(d->qacc[i] - d->qacc_smooth[i]);
  c->gauss = (real)0.5*g;
  c->cost = cost + c->gauss;
}
/* gradient; Newton: Hessian H = M + J^T diag(D active) J, search = -H^-1 grad; CG (engine_solver.c mj_solPrimal with
 * flg_Newton = 0): Mgrad = M^-1 grad, search = -Mgrad + beta search with the Polak-Ribiere beta, clamped at 0 */
static void sol_update_gradient(solctx* c) {
  orc_data* d = c->d; int nv = c->nv;
  for (int i = 0; i < nv; i++) c->grad[i] = c->Ma[i] - d->qfrc_smooth[i] - d->qfrc_constraint[i];
  if (c->cg) {
    if (!c->have_ml) { chol_factor(c->HL, d->qM, nv); c->have_ml = 1; }
    chol_solve(c->Mgrad, c->HL, c->grad, nv);
    return;
  }
  for (int i = 0; i < nv*nv; i++) c->H[i] = d->qM[i];
  for (int r = 0; r < c->nefc; r++) {
    if (!c->active[r]) continue;
    if (c->active[r] == 2) {
      /* cone Hessian of a middle-zone contact: Dm diag(fri) [1, -mu U'/T; ., mu N/T^3 U U' + (mu^2 - mu N/T) I] diag(fri) */
      conezone z; cone_eval(d, r, c->Jaref, &z);
      real Hc[36];
      for (int a = 0; a < z.dim; a++) for (int b = 0; b < z.dim; b++) {
        real v;
        if (a == 0 && b == 0) v = 1;
        else if (a == 0 || b == 0) v = -z.mu*z.U[a + b]/z.T;
        else v = z.mu*z.N/(z.T*z.T*z.T)*z.U[a]*z.U[b] + (a == b ? z.mu*z.mu - z.mu*z.N/z.T : 0);
        Hc[a*z.dim + b] = v*z.Dm*z.fri[a]*z.fri[b];
      }
      for (int a = 0; a < z.dim; a++) for (int b = 0; b < z.dim; b++) {
        const real* Ja = d->efc_J + (r + a)*nv; const real* Jb = d->efc_J + (r + b)*nv; real h = Hc[a*z.dim + b];
        for (int i = 0; i < nv; i++) { real t = h*Ja[i]; if (t == 0) continue; for (int j = 0; j <= i; j++) c->H[i*nv + j] += t*Jb[j]; }
      }
      continue;
    }
    const real* J = d->efc_J + r*nv; real D = d->efc_D[r];
    for (int i = 0; i < nv; i++) { real t = D*J[i]; if (t == 0) continue; for (int j = 0; j <= i; j++) c->H[i*nv + j] += t*J[j]; }
  }
  for (int i = 0; i < nv; i++) for (int j = 0; j < i; j++) c->H[j*nv + i] = c->H[i*nv + j];
  chol_factor(c->HL, c->H, nv);
  chol_solve(c->Mgrad, c->HL, c->grad, nv);
  for (int i = 0; i < nv; i++) c->search[i] = -c->Mgrad[i];
}
typedef struct { real alpha, cost, d0, d1; } lspoint;
static lspoint ls_eval(const solctx* c, real alpha) {
  real q0 = c->quadGauss[0], q1 = c->quadGauss[1], q2 = c->quadGauss[2];
  real xc = 0, xd0 = 0, xd1 = 0;      /* non-quadratic part: middle zones of elliptic contacts */
  for (int r = 0; r < c->nefc; r++) {
    real x = c->Jaref[r] + alpha*c->Jv[r];
    if (r < c->nf) {
      real f = c->d->efc_frictionloss[r], rf = f/c->d->efc_D[r];
      if (x <= -rf) { q0 += f*(-(real)0.5*rf - c->Jaref[r]); q1 += -f*c->Jv[r]; }
      else if (x >= rf) { q0 += f*(-(real)0.5*rf + c->Jaref[r]); q1 += f*c->Jv[r]; }
      else { q0 += c->quad[3*r]; q1 += c->quad[3*r + 1]; q2 += c->quad[3*r + 2]; }
      continue;
    }
    if (c->d->efc_type[r] == ORC_EFC_CONTACT_ELLIPTIC) {
      /* PrimalEval, elliptic contact: N and T along the search direction, zones as in cone_eval */
      const orc_data* d = c->d;
      conezone z; cone_eval(d, r, c->Jaref, &z);       /* gives mu, fri, Dm and U = U(0) */
      real V[6], UU = 0, UV = 0, VV = 0;
      for (int k = 0; k < z.dim; k++) { V[k] = c->Jv[r + k]*z.fri[k]; if (k) { UU += z.U[k]*z.U[k]; UV += z.U[k]*V[k]; VV += V[k]*V[k]; } }
      real N = z.U[0] + alpha*V[0], Tsqr = UU + alpha*(2*UV + alpha*VV);
      real T = Tsqr > 0 ? (real)sqrt((double)Tsqr) : 0;
      if (N >= z.mu*T || (Tsqr <= 0 && N >= 0)) { /* top zone */ }
      else if (z.mu*N + T <= 0 || (Tsqr <= 0 && N < 0)) {
        for (int k = 0; k < z.dim; k++) { q0 += c->quad[3*(r + k)]; q1 += c->quad[3*(r + k) + 1]; q2 += c->quad[3*(r + k) + 2]; }
      } else {
        real N1 = V[0], T1 = (UV + alpha*VV)/T, T2 = VV/T - (UV + alpha*VV)*(UV + alpha*VV)/(T*T*T);
        real nmt = N - z.mu*T;
        xc += (real)0.5*z.Dm*nmt*nmt;
        xd0 += z.Dm*nmt*(N1 - z.mu*T1);
        xd1 += z.Dm*((N1 - z.mu*T1)*(N1 - z.mu*T1) + nmt*(-z.mu*T2));
      }
      r += z.dim - 1;
      continue;
    }
    if (x < 0) { q0 += c->quad[3*r]; q1 += c->quad[3*r + 1]; q2 += c->quad[3*r + 2]; }
  }
  lspoint p;
  p.alpha = alpha; p.cost = alpha*alpha*q2 + alpha*q1 + q0 + xc; p.d0 = 2*alpha*q2 + q1 + xd0; p.d1 = 2*q2 + xd1;
  if (p.d1 < MINVAL) p.d1 = MINVAL;
  return p;
}
static void ls_prepare(solctx* c) {
  orc_data* d = c->d; int nv = c->nv;
  for (int i = 0; i < nv; i++) { real s = 0; for (int j = 0; j < nv; j++) s += d->qM[i*nv + j]*c->search[j]; c->Mv[i] = s; }
  for (int r = 0; r < c->nefc; r++) { real s = 0; for (int j = 0; j < nv; j++) s += d->efc_J[r*nv + j]*c->search[j]; c->Jv[r] = s; }
  real g1 = 0, g2 = 0;
  for (int i = 0; i < nv; i++) { g1 += c->search[i]*(c->Ma[i] - d->qfrc_smooth[i]); g2 += c->search[i]*c->Mv[i]; }
  c->quadGauss[0] = c->gauss; c->quadGauss[1] = g1; c->quadGauss[2] = (real)0.5*g2;
  for (int r = 0; r < c->nefc; r++) {
    real D = d->efc_D[r], ja = c->Jaref[r], jv = c->Jv[r];
    c->quad[3*r] = (real)0.5*D*ja*ja; c->quad[3*r + 1] = D*ja*jv; c->quad[3*r + 2] = (real)0.5*D*jv*jv;
  }
}
/* exact 1-D Newton line search with bracketing (engine_solver.c PrimalLinesearch, restated) */
static real ls_iterative(solctx* c) {
  const b2s_model_desc* m = c->m;
  real snorm = 0;
  for (int i = 0; i < c->nv; i++) snorm += c->search[i]*c->search[i];
  snorm = (real)sqrt((double)snorm);
  if (snorm < MINVAL) return 0;
  real gtol = (real)(m->opt_tolerance*m->opt_ls_tolerance)*snorm*(real)m->stat_meaninertia*(real)(c->nv > 1 ? c->nv : 1);
  int maxit = m->opt_ls_iterations, it = 0;
  lspoint p0 = ls_eval(c, 0);
  lspoint p1 = ls_eval(c, p0.alpha - p0.d0/p0.d1); it++;
  if (p0.cost < p1.cost) p1 = p0;
  if (fabs((double)p1.d0) < gtol) return p1.alpha;
  real dir = p1.d0 < 0 ? (real)1 : (real)-1;
  lspoint p2 = p0; int p2upd = 0;
  while (p1.d0*dir <= -gtol && it < maxit) {
    p2 = p1; p2upd = 1;
    p1 = ls_eval(c, p1.alpha - p1.d0/p1.d1); it++;
    if (fabs((double)p1.d0) < gtol) return p1.alpha;
  }
  if (it >= maxit || !p2upd) return p1.alpha;
  /* bracket [p2, p1]: derivative changes sign between them */
  lspoint p2next = p1, p1next = ls_eval(c, p1.alpha - p1.d0/p1.d1); it++;
  while (it < maxit) {
    lspoint pmid = ls_eval(c, (real)0.5*(p1.alpha + p2.alpha)); it++;
    lspoint cand[3] = {p1next, p2next, pmid};
    int b1 = 0, b2 = 0;
    for (int k = 0; k < 3; k++) {
      if (fabs((double)cand[k].d0) < gtol) return cand[k].alpha;
      /* tighten whichever end has the same derivative sign and lies inside the bracket */
      if (cand[k].d0*dir < 0 && (cand[k].alpha - p2.alpha)*dir > 0 && (p1.alpha - cand[k].alpha)*dir > 0 && !b2) { p2 = cand[k]; b2 = 1; }
      else if (cand[k].d0*dir > 0 && (cand[k].alpha - p2.alpha)*dir > 0 && (p1.alpha - cand[k].alpha)*dir > 0 && !b1) { p1 = cand[k]; b1 = 1; }
    }
    if (!b1 && !b2) return pmid.cost < p0.cost ? pmid.alpha : 0;
    if (b1) p1next = ls_eval(c, p1.alpha - p1.d0/p1.d1), it++;
    if (b2) p2next = ls_eval(c, p2.alpha - p2.d0/p2.d1), it++;
  }
  if (p1.cost <= p2.cost && p1.cost < p0.cost) return p1.alpha;
  if (p2.cost <= p1.cost && p2.cost < p0.cost) return p2.alpha;
  return 0;
}
/* mjwarp's parallel line search: ls_iterations log-spaced step sizes in [min_step, 1], best cost wins
 * (reference option: sim.py:91 ls_parallel=True).  A step is taken only if it improves on alpha = 0. */
static real ls_parallel(solctx* c) {
  const b2s_model_desc* m = c->m;
  int n = m->opt_ls_iterations;
  real best = ls_eval(c, 0).cost, balpha = 0;
  double lg = log(m->opt_ls_parallel_min_step);
  for (int k = 0; k < n; k++) {
    real alpha = n > 1 ? (real)exp((double)k/(double)(n - 1)*lg) : (real)1;
    real cost = ls_eval(c, alpha).cost;
    if (cost < best) { best = cost; balpha = alpha; }
  }
  return balpha;
}

void orc_solve(const b2s_model_desc* m, orc_data* d) {
  int nv = m->nv, nefc = d->nefc[0];
  d->solver_niter[0] = 0; d->solver_cost[0] = 0;
  if (nefc == 0) {
    for (int i = 0; i < nv; i++) { d->qacc[i] = d->qacc_smooth[i]; d->qfrc_constraint[i] = 0; }
    return;
  }
  solctx c; c.m = m; c.d = d; c.nv = nv; c.nefc = nefc; c.nf = d->nf[0];
  real* buf = (real*)calloc((size_t)(9*nv + 2*nv*nv + 5*nefc), sizeof(real));
  c.cg = m->opt_solver == B2S_SOL_CG; c.have_ml = 0;
  c.Ma = buf; c.grad = c.Ma + nv; c.Mgrad = c.grad + nv; c.search = c.Mgrad + nv; c.Mv = c.search + nv;
  c.H = c.Mv + nv; c.HL = c.H + nv*nv; c.Jaref = c.HL + nv*nv; c.Jv = c.Jaref + nefc; c.quad = c.Jv + nefc;
  c.gradold = c.quad + 3*nefc; c.Mgradold = c.gradold + nv;
  c.active = (int*)calloc((size_t)nefc, sizeof(int));
  /* warm start: keep qacc_warmstart only if its cost beats qacc_smooth (mj_fwdConstraint) */
  for (int pass = 0; pass < 2; pass++) {
    const real* src = (pass == 0 && !(m->opt_disableflags & B2S_DSBL_WARMSTART)) ? d->qacc_warmstart : d->qacc_smooth;
    for (int i = 0; i < nv; i++) d->qacc[i] = src[i];
    for (int i = 0; i < nv; i++) { real s = 0; for (int j = 0; j < nv; j++) s += d->qM[i*nv + j]*d->qacc[j]; c.Ma[i] = s; }
    for (int r = 0; r < nefc; r++) { real s = 0; for (int j = 0; j < nv; j++) s += d->efc_J[r*nv + j]*d->qacc[j]; c.Jaref[r] = s - d->efc_aref[r]; }
    sol_update_constraint(&c);
    if (pass == 0) {
      if (src == d->qacc_smooth) break;
      real cost_warm = c.cost, cost_smooth = 0;
      for (int r = 0; r < nefc; r++) {
        real s = 0; for (int j = 0; j < nv; j++) s += d->efc_J[r*nv + j]*d->qacc_smooth[j];
        s -= d->efc_aref[r];
        if (r < c.nf) {
          real f = d->efc_frictionloss[r], rf = f/d->efc_D[r];
          if (s <= -rf) cost_smooth += -f*((real)0.5*rf + s);
          else if (s >= rf) cost_smooth += -f*((real)0.5*rf - s);
          else cost_smooth += (real)0.5*d->efc_D[r]*s*s;
        } else if (d->efc_type[r] == ORC_EFC_CONTACT_ELLIPTIC) c.Jv[r] = s;      /* collected, evaluated per contact below */
        else if (s < 0) cost_smooth += (real)0.5*d->efc_D[r]*s*s;
      }
      for (int r = 0; r < nefc; r++) {
        if (d->efc_type[r] != ORC_EFC_CONTACT_ELLIPTIC) continue;
        conezone z; cone_eval(d, r, c.Jv, &z);
        if (z.zone == 1) for (int k = 0; k < z.dim; k++) cost_smooth += (real)0.5*d->efc_D[r + k]*c.Jv[r + k]*c.Jv[r + k];
        else if (z.zone == 2) { real nmt = z.N - z.mu*z.T; cost_smooth += (real)0.5*z.Dm*nmt*nmt; }
        r += z.dim - 1;
      }
      if (cost_warm <= cost_smooth) break;
    }
  }
  sol_update_gradient(&c);
  if (c.cg) for (int i = 0; i < nv; i++) c.search[i] = -c.Mgrad[i];
  real scale = 1/((real)m->stat_meaninertia*(real)(nv > 1 ? nv : 1));
  int it = 0;
  for (; it < m->opt_iterations; it++) {
    ls_prepare(&c);
    real alpha = m->opt_ls_parallel ? ls_parallel(&c) : ls_iterative(&c);
    if (alpha == 0) break;
    for (int i = 0; i < nv; i++) { d->qacc[i] += alpha*c.search[i]; c.Ma[i] += alpha*c.Mv[i]; }
    for (int r = 0; r < nefc; r++) c.Jaref[r] += alpha*c.Jv[r];
    real old = c.cost;
    if (c.cg) for (int i = 0; i < nv; i++) { c.gradold[i] = c.grad[i]; c.Mgradold[i] = c.Mgrad[i]; }
    sol_update_constraint(&c);
    sol_update_gradient(&c);
    if (c.cg) {
      real num = 0, den = 0;
      for (int i = 0; i < nv; i++) { num += c.grad[i]*(c.Mgrad[i] - c.Mgradold[i]); den += c.gradold[i]*c.Mgradold[i]; }
      real beta = num/(den > MINVAL ? den : MINVAL);
      if (beta < 0) beta = 0;
      for (int i = 0; i < nv; i++) c.search[i] = -c.Mgrad[i] + beta*c.search[i];
    }
    real improvement = scale*(old - c.cost), gn = 0;
    for (int i = 0; i < nv; i++) gn += c.grad[i]*c.grad[i];
    real gradient = scale*(real)sqrt((double)gn);
    if (improvement < (real)m->opt_tolerance || gradient < (real)m->opt_tolerance) { it++; break; }
  }
  d->solver_niter[0] = it; d->solver_cost[0] = c.cost;
  free(buf); free(c.active);
}

/* --------------------------------------------------------------- sensors */
/* mj_objectVelocity: 6D velocity [ang, lin] of a frame at `pos` on `body`; rotated by rot^T if rot != NULL */
static void object_velocity(const b2s_model_desc* m, const orc_data* d, int body, const real* pos, const real* rot, real* out) {
  const real* cv = d->cvel + 6*body;
  real dif[3], lin[3], t[3];
  for (int k = 0; k < 3; k++) dif[k] = pos[k] - d->subtree_com[3*m->body_rootid[body] + k];
  cross3(t, cv, dif);
  for (int k = 0; k < 3; k++) lin[k] = cv[3 + k] + t[k];
  if (rot) { matT_vec(out, rot, cv); matT_vec(out + 3, rot, lin); }
  else for (int k = 0; k < 3; k++) { out[k] = cv[k]; out[3 + k] = lin[k]; }
}
static int in_subtree(const b2s_model_desc* m, int body, int root) {
  while (body > 0 && body != root) body = m->body_parentid[body];
  return body == root;
}
static int side_match(const b2s_model_desc* m, int objtype, int objid, int geom) {
  int body = m->geom_bodyid[geom];
  if (objtype == B2S_OBJ_GEOM) return geom == objid;
  if (objtype == B2S_OBJ_BODY) return body == objid;
  if (objtype == B2S_OBJ_XBODY) return in_subtree(m, body, objid);
  return 1;  /* unspecified side matches anything */
}
/* mj_contactForce for a pyramidal / frictionless contact: wrench in the contact frame */
static void contact_wrench(const orc_data* d, int c, real* f6) {
  for (int k = 0; k < 6; k++) f6[k] = 0;
  int adr = d->con_efc_address[c], dim = d->con_dim[c];
  if (adr < 0) return;
  if (dim == 1) { f6[0] = d->efc_force[adr]; return; }
  if (d->efc_type[adr] == ORC_EFC_CONTACT_ELLIPTIC) { for (int k = 0; k < dim; k++) f6[k] = d->efc_force[adr + k]; return; }
  for (int k = 0; k < dim - 1; k++) {
    real a = d->efc_force[adr + 2*k], b = d->efc_force[adr + 2*k + 1];
    f6[0] += a + b;
    f6[1 + k] = d->con_friction[5*c + k]*(a - b);
  }
}
static void subtree_vel(const b2s_model_desc* m, orc_data* d) {
  int nb = m->nbody;
  real* bvel = (real*)malloc(sizeof(real)*6*nb);
  for (int i = 0; i < nb; i++) {
    object_velocity(m, d, i, d->xipos + 3*i, NULL, bvel + 6*i);
    real mass = (real)m->body_mass[i], t[3];
    for (int k = 0; k < 3; k++) d->subtree_linvel[3*i + k] = mass*bvel[6*i + 3 + k];
    matT_vec(t, d->ximat + 9*i, bvel + 6*i);
    for (int k = 0; k < 3; k++) t[k] *= (real)m->body_inertia[3*i + k];
    mat_vec(d->subtree_angmom + 3*i, d->ximat + 9*i, t);
  }
  for (int i = nb - 1; i >= 0; i--) {
    if (i) for (int k = 0; k < 3; k++) d->subtree_linvel[3*m->body_parentid[i] + k] += d->subtree_linvel[3*i + k];
    real sm = (real)m->body_subtreemass[i]; if (sm < MINVAL) sm = MINVAL;
    for (int k = 0; k < 3; k++) d->subtree_linvel[3*i + k] /= sm;
  }
  for (int i = nb - 1; i > 0; i--) {
    int p = m->body_parentid[i];
    real dx[3], dv[3], dL[3], mass = (real)m->body_mass[i], sm = (real)m->body_subtreemass[i];
    for (int k = 0; k < 3; k++) { dx[k] = d->xipos[3*i + k] - d->subtree_com[3*i + k]; dv[k] = mass*(bvel[6*i + 3 + k] - d->subtree_linvel[3*i + k]); }
    cross3(dL, dx, dv);
    for (int k = 0; k < 3; k++) d->subtree_angmom[3*i + k] += dL[k];
    for (int k = 0; k < 3; k++) d->subtree_angmom[3*p + k] += d->subtree_angmom[3*i + k];
    for (int k = 0; k < 3; k++) { dx[k] = d->subtree_com[3*i + k] - d->subtree_com[3*p + k]; dv[k] = sm*(d->subtree_linvel[3*i + k] - d->subtree_linvel[3*p + k]); }
    cross3(dL, dx, dv);
    for (int k = 0; k < 3; k++) d->subtree_angmom[3*p + k] += dL[k];
  }
  free(bvel);
}
static void obj_frame(const b2s_model_desc* m, const orc_data* d, int type, int id, const real** pos, const real** mat, int* body) {
  if (type == B2S_OBJ_SITE) { *pos = d->site_xpos + 3*id; *mat = d->site_xmat + 9*id; *body = m->site_bodyid[id]; }
  else if (type == B2S_OBJ_GEOM) { *pos = d->geom_xpos + 3*id; *mat = d->geom_xmat + 9*id; *body = m->geom_bodyid[id]; }
  else if (type == B2S_OBJ_XBODY) { *pos = d->xpos + 3*id; *mat = d->xmat + 9*id; *body = id; }
  else { *pos = d->xipos + 3*id; *mat = d->ximat + 9*id; *body = id; }
}
static void mat2quat_(real* q, const real* R) {
  real tr = R[0] + R[4] + R[8];
  if (tr > 0) { real s = (real)sqrt((double)tr + 1)*2; q[0] = s/4; q[1] = (R[7] - R[5])/s; q[2] = (R[2] - R[6])/s; q[3] = (R[3] - R[1])/s; }
  else if (R[0] > R[4] && R[0] > R[8]) { real s = (real)sqrt(1.0 + (double)R[0] - R[4] - R[8])*2; q[0] = (R[7] - R[5])/s; q[1] = s/4; q[2] = (R[1] + R[3])/s; q[3] = (R[2] + R[6])/s; }
  else if (R[4] > R[8]) { real s = (real)sqrt(1.0 + (double)R[4] - R[0] - R[8])*2; q[0] = (R[2] - R[6])/s; q[1] = (R[1] + R[3])/s; q[2] = s/4; q[3] = (R[5] + R[7])/s; }
  else { real s = (real)sqrt(1.0 + (double)R[8] - R[0] - R[4])*2; q[0] = (R[3] - R[1])/s; q[1] = (R[2] + R[6])/s; q[2] = (R[5] + R[7])/s; q[3] = s/4; }
  normalize4(q);
}

/* engine_ray.c [ext]: distance along the ray (pnt, vec) to one primitive geom, -1 if none.  The ray is moved into the geom
   frame; ray_quad returns the smaller non-negative root of a x^2 + 2 b x + c = 0. */
static real ray_quad(real a, real b, real c) {
  real det = b*b - a*c;
  if (det < MINVAL) return -1;
  real sq = (real)sqrt((double)det);
  real x0 = (-b - sq)/a, x1 = (-b + sq)/a;
  if (x0 >= 0) return x0;
  if (x1 >= 0) return x1;
  return -1;
}
static real ray_geom(int type, const double* size, const real* gpos, const real* gmat, const real* pnt, const real* vec) {
  real dif[3], p[3], v[3], x = -1, t;
  for (int k = 0; k < 3; k++) dif[k] = pnt[k] - gpos[k];
  matT_vec(p, gmat, dif); matT_vec(v, gmat, vec);
  real r = (real)size[0], h = (real)size[1];
  switch (type) {
    case B2S_GEOM_PLANE:
      if (v[2] > -MINVAL) return -1;
      t = -p[2]/v[2];
      if (t < 0) return -1;
      if ((size[0] <= 0 || fabs((double)(p[0] + t*v[0])) <= size[0]) && (size[1] <= 0 || fabs((double)(p[1] + t*v[1])) <= size[1])) x = t;
      break;
    case B2S_GEOM_SPHERE:
      x = ray_quad(dot3(v, v), dot3(v, p), dot3(p, p) - r*r);
      break;
    case B2S_GEOM_CAPSULE:
      t = ray_quad(v[0]*v[0] + v[1]*v[1], v[0]*p[0] + v[1]*p[1], p[0]*p[0] + p[1]*p[1] - r*r);
      if (t >= 0 && fabs((double)(p[2] + t*v[2])) <= h) x = t;
      for (int side = -1; side <= 1; side += 2) {
        real q[3] = {p[0], p[1], p[2] - side*h};
        t = ray_quad(dot3(v, v), dot3(v, q), dot3(q, q) - r*r);
        if (t >= 0 && (x < 0 || t < x) && side*(p[2] + t*v[2]) >= h) x = t;
      }
      break;
    case B2S_GEOM_ELLIPSOID: {
      real s[3];
      for (int k = 0; k < 3; k++) s[k] = (real)(1.0/(size[k]*size[k]));
      x = ray_quad(s[0]*v[0]*v[0] + s[1]*v[1]*v[1] + s[2]*v[2]*v[2], s[0]*p[0]*v[0] + s[1]*p[1]*v[1] + s[2]*p[2]*v[2],
                   s[0]*p[0]*p[0] + s[1]*p[1]*p[1] + s[2]*p[2]*p[2] - 1);
    } break;
    case B2S_GEOM_CYLINDER:
      if (fabs((double)v[2]) > MINVAL)
        for (int side = -1; side <= 1; side += 2) {
          t = (side*h - p[2])/v[2];
          if (t >= 0 && (x < 0 || t < x)) {
            real px = p[0] + t*v[0], py = p[1] + t*v[1];
            if (px*px + py*py <= r*r) x = t;
          }
        }
      t = ray_quad(v[0]*v[0] + v[1]*v[1], v[0]*p[0] + v[1]*p[1], p[0]*p[0] + p[1]*p[1] - r*r);
      if (t >= 0 && (x < 0 || t < x) && fabs((double)(p[2] + t*v[2])) <= h) x = t;
      break;
    case B2S_GEOM_BOX:
      for (int i = 0; i < 3; i++) {
        if (fabs((double)v[i]) <= MINVAL) continue;
        for (int side = -1; side <= 1; side += 2) {
          t = (side*(real)size[i] - p[i])/v[i];
          if (t >= 0 && (x < 0 || t < x)) {
            int j = (i + 1)%3, k = (i + 2)%3;
            if (fabs((double)(p[j] + t*v[j])) <= size[j] && fabs((double)(p[k] + t*v[k])) <= size[k]) x = t;
          }
        }
      }
      break;
    default: break;
  }
  return x;
}

void orc_sensors(const b2s_model_desc* m, orc_data* d) {
  int nb = m->nbody;
  /* mj_rnePostConstraint, acceleration part only: cacc with the solved qacc */
  int grav = !(m->opt_disableflags & B2S_DSBL_GRAVITY);
  for (int k = 0; k < 3; k++) { d->cacc[k] = 0; d->cacc[3 + k] = grav ? -(real)m->opt_gravity[k] : 0; }
  for (int i = 1; i < nb; i++) {
    real* a = d->cacc + 6*i;
    for (int k = 0; k < 6; k++) a[k] = d->cacc[6*m->body_parentid[i] + k];
    for (int dd = 0; dd < m->body_dofnum[i]; dd++) {
      int da = m->body_dofadr[i] + dd;
      for (int k = 0; k < 6; k++) a[k] += d->cdof_dot[6*da + k]*d->qvel[da] + d->cdof[6*da + k]*d->qacc[da];
    }
  }
  int need_sub = 0;
  for (int s = 0; s < m->nsensor; s++)
    if (m->sensor_type[s] == B2S_SENS_SUBTREELINVEL || m->sensor_type[s] == B2S_SENS_SUBTREEANGMOM) need_sub = 1;
  if (need_sub) subtree_vel(m, d);
  /* mj_rnePostConstraint, force part (only when a force / torque sensor asks for it): cfrc_ext from xfrc_applied and the
     contact forces, cfrc_int = cinert cacc + cvel x* (cinert cvel) - cfrc_ext, summed from the leaves to the root.  Spatial
     vectors are [torque, force] about the subtree COM of the body's root, world orientation. */
  real* cfrc_int = NULL;
  for (int s = 0; s < m->nsensor; s++)
    if (m->sensor_type[s] == B2S_SENS_FORCE || m->sensor_type[s] == B2S_SENS_TORQUE) need_sub |= 2;
  if (need_sub & 2) {
    cfrc_int = (real*)malloc(sizeof(real)*12*nb);
    real* cfrc_ext = cfrc_int + 6*nb;
    for (int k = 0; k < 12*nb; k++) cfrc_int[k] = 0;
    for (int b = 1; b < nb; b++) {
      const real* x = d->xfrc_applied + 6*b;
      if (x[0] == 0 && x[1] == 0 && x[2] == 0 && x[3] == 0 && x[4] == 0 && x[5] == 0) continue;
      real dif[3], cr[3];
      for (int k = 0; k < 3; k++) dif[k] = d->subtree_com[3*m->body_rootid[b] + k] - d->xipos[3*b + k];
      cross3(cr, dif, x);
      for (int k = 0; k < 3; k++) { cfrc_ext[6*b + k] += x[3 + k] - cr[k]; cfrc_ext[6*b + 3 + k] += x[k]; }
    }
    for (int c = 0; c < d->ncon[0]; c++) {
      if (d->con_efc_address[c] < 0) continue;
      real f6[6], fw[3], tw[3];
      contact_wrench(d, c, f6);
      matT_vec(fw, d->con_frame + 9*c, f6);
      matT_vec(tw, d->con_frame + 9*c, f6 + 3);
      for (int side = 0; side < 2; side++) {
        int b = m->geom_bodyid[d->con_geom[2*c + side]];
        if (!b) continue;
        real dif[3], cr[3], sg = side ? (real)1 : (real)-1;
        for (int k = 0; k < 3; k++) dif[k] = d->subtree_com[3*m->body_rootid[b] + k] - d->con_pos[3*c + k];
        cross3(cr, dif, fw);
        for (int k = 0; k < 3; k++) { cfrc_ext[6*b + k] += sg*(tw[k] - cr[k]); cfrc_ext[6*b + 3 + k] += sg*fw[k]; }
      }
    }
    for (int i = 1; i < nb; i++) {
      real t0[6], t1[6], t2[6];
      mul_inert_vec(t0, d->cinert + 10*i, d->cacc + 6*i);
      mul_inert_vec(t1, d->cinert + 10*i, d->cvel + 6*i);
      cross_force(t2, d->cvel + 6*i, t1);
      for (int k = 0; k < 6; k++) cfrc_int[6*i + k] = t0[k] + t2[k] - cfrc_ext[6*i + k];
    }
    for (int i = nb - 1; i > 0; i--) {
      int p = m->body_parentid[i];
      if (p) for (int k = 0; k < 6; k++) cfrc_int[6*p + k] += cfrc_int[6*i + k];
    }
  }
  for (int s = 0; s < m->nsensor; s++) {
    int type = m->sensor_type[s], id = m->sensor_objid[s], ot = m->sensor_objtype[s], dim = m->sensor_dim[s];
    real* out = d->sensordata + m->sensor_adr[s];
    const real *pos, *mat; int body;
    real v6[6];
    switch (type) {
      case B2S_SENS_GYRO:
        obj_frame(m, d, B2S_OBJ_SITE, id, &pos, &mat, &body);
        object_velocity(m, d, body, pos, mat, v6);
        for (int k = 0; k < 3; k++) out[k] = v6[k];
        break;
      case B2S_SENS_VELOCIMETER:
        obj_frame(m, d, B2S_OBJ_SITE, id, &pos, &mat, &body);
        object_velocity(m, d, body, pos, mat, v6);
        for (int k = 0; k < 3; k++) out[k] = v6[3 + k];
        break;
      case B2S_SENS_RANGEFINDER: {
        /* mj_ray from the site along its z axis, geoms of the site's own body excluded; -1 without a hit */
        obj_frame(m, d, B2S_OBJ_SITE, id, &pos, &mat, &body);
        real dir[3] = {mat[2], mat[5], mat[8]}, best = -1;
        for (int g = 0; g < m->ngeom; g++) {
          if (m->geom_bodyid[g] == body) continue;
          real x = ray_geom(m->geom_type[g], m->geom_size + 3*g, d->geom_xpos + 3*g, d->geom_xmat + 9*g, pos, dir);
          if (x >= 0 && (best < 0 || x < best)) best = x;
        }
        out[0] = best;
      } break;
      case B2S_SENS_FORCE: case B2S_SENS_TORQUE: {
        /* interaction force between the site's body and its parent, moved to the site and expressed in its frame
           (mju_transformSpatial with flg_force: the torque changes by -(newpos - oldpos) x force) */
        obj_frame(m, d, B2S_OBJ_SITE, id, &pos, &mat, &body);
        const real* ci = cfrc_int + 6*body;
        real dif[3], cr[3], tq[3];
        for (int k = 0; k < 3; k++) dif[k] = pos[k] - d->subtree_com[3*m->body_rootid[body] + k];
        cross3(cr, dif, ci + 3);
        for (int k = 0; k < 3; k++) tq[k] = ci[k] - cr[k];
        matT_vec(out, mat, type == B2S_SENS_FORCE ? ci + 3 : tq);
      } break;
      case B2S_SENS_ACCELEROMETER: {
        /* mj_objectAcceleration (local frame): cacc moved to the site + centripetal correction */
        obj_frame(m, d, B2S_OBJ_SITE, id, &pos, &mat, &body);
        object_velocity(m, d, body, pos, mat, v6);
        const real* ca = d->cacc + 6*body;
        real dif[3], t[3], lin[3], loc[3], corr[3];
        for (int k = 0; k < 3; k++) dif[k] = pos[k] - d->subtree_com[3*m->body_rootid[body] + k];
        cross3(t, ca, dif);
        for (int k = 0; k < 3; k++) lin[k] = ca[3 + k] + t[k];
        matT_vec(loc, mat, lin);
        cross3(corr, v6, v6 + 3);
        for (int k = 0; k < 3; k++) out[k] = loc[k] + corr[k];
      } break;
      case B2S_SENS_JOINTPOS: out[0] = d->qpos[m->jnt_qposadr[id]]; break;
      case B2S_SENS_JOINTVEL: out[0] = d->qvel[m->jnt_dofadr[id]]; break;
      case B2S_SENS_ACTUATORFRC: out[0] = d->actuator_force[id]; break;
      /* engine_sensor.c mj_sensorPos / mj_sensorVel / mj_sensorAcc scalar read-outs */
      case B2S_SENS_TENDONPOS: out[0] = d->ten_length[id]; break;
      case B2S_SENS_TENDONVEL: out[0] = d->ten_velocity[id]; break;
      case B2S_SENS_ACTUATORPOS: out[0] = d->actuator_length[id]; break;
      case B2S_SENS_ACTUATORVEL: out[0] = d->actuator_velocity[id]; break;
      case B2S_SENS_JOINTACTFRC: out[0] = d->qfrc_actuator[m->jnt_dofadr[id]]; break;
      case B2S_SENS_CLOCK: out[0] = d->time[0]; break;
      case B2S_SENS_MAGNETOMETER: {      /* mj_sensorPos: the global field opt.magnetic in the site frame */
        obj_frame(m, d, B2S_OBJ_SITE, id, &pos, &mat, &body);
        for (int k = 0; k < 3; k++) out[k] = mat[k]*(real)m->opt_magnetic[0] + mat[3 + k]*(real)m->opt_magnetic[1] + mat[6 + k]*(real)m->opt_magnetic[2];
      } break;
      case B2S_SENS_JOINTLIMITPOS: case B2S_SENS_JOINTLIMITVEL: case B2S_SENS_JOINTLIMITFRC: {
        /* the joint's limit row, if one is active: pos - margin (mj_sensorPos), its velocity (mj_sensorVel), its force (mj_sensorAcc) */
        out[0] = 0;
        for (int r = 0; r < d->nefc[0]; r++)
          if (d->efc_type[r] == ORC_EFC_LIMIT && d->efc_id[r] == id) {
            out[0] = type == B2S_SENS_JOINTLIMITPOS ? d->efc_pos[r] - d->efc_margin[r] : (type == B2S_SENS_JOINTLIMITVEL ? d->efc_vel[r] : d->efc_force[r]);
            break;
          }
      } break;
      case B2S_SENS_TENDONACTFRC: {      /* sum of the forces of the actuators that pull on this tendon */
        real f = 0;
        for (int u = 0; u < m->nu; u++) if (m->actuator_trntype[u] == B2S_TRN_TENDON && m->actuator_trnid[u] == id) f += d->actuator_force[u];
        out[0] = f;
      } break;
      case B2S_SENS_E_POTENTIAL: {       /* mj_energyPos: gravity + joint springs (hinge / slide) + tendon springs outside their dead band */
        real e = 0;
        if (grav) for (int i = 1; i < nb; i++)
          e -= (real)m->body_mass[i]*((real)m->opt_gravity[0]*d->xipos[3*i] + (real)m->opt_gravity[1]*d->xipos[3*i + 1] + (real)m->opt_gravity[2]*d->xipos[3*i + 2]);
        if (!(m->opt_disableflags & B2S_DSBL_PASSIVE)) {
          for (int j = 0; j < m->njnt; j++) {
            real k = (real)m->jnt_stiffness[j];
            if (k == 0 || (m->jnt_type[j] != B2S_JNT_HINGE && m->jnt_type[j] != B2S_JNT_SLIDE)) continue;
            real dq = d->qpos[m->jnt_qposadr[j]] - (real)m->qpos_spring[m->jnt_qposadr[j]];
            e += (real)0.5*k*dq*dq;
          }
          for (int t = 0; t < m->ntendon; t++) {
            real k = (real)m->tendon_stiffness[t], len = d->ten_length[t], lo = (real)m->tendon_lengthspring[2*t], hi = (real)m->tendon_lengthspring[2*t + 1];
            real dl = len > hi ? len - hi : (len < lo ? len - lo : 0);
            e += (real)0.5*k*dl*dl;
          }
        }
        out[0] = e;
      } break;
      case B2S_SENS_E_KINETIC: {         /* mj_energyVel: 1/2 v' M v */
        real e = 0;
        for (int i = 0; i < m->nv; i++) { real r = 0; for (int j = 0; j < m->nv; j++) r += d->qM[i*m->nv + j]*d->qvel[j]; e += d->qvel[i]*r; }
        out[0] = (real)0.5*e;
      } break;
      case B2S_SENS_SUBTREECOM: for (int k = 0; k < 3; k++) out[k] = d->subtree_com[3*id + k]; break;
      case B2S_SENS_SUBTREELINVEL: for (int k = 0; k < 3; k++) out[k] = d->subtree_linvel[3*id + k]; break;
      case B2S_SENS_SUBTREEANGMOM: for (int k = 0; k < 3; k++) out[k] = d->subtree_angmom[3*id + k]; break;
      case B2S_SENS_FRAMELINACC: case B2S_SENS_FRAMEANGACC: {
        /* mj_objectAcceleration, world orientation: cacc moved to the frame origin + the centripetal term w x v */
        obj_frame(m, d, ot, id, &pos, &mat, &body);
        object_velocity(m, d, body, pos, NULL, v6);
        const real* ca = d->cacc + 6*body;
        if (type == B2S_SENS_FRAMEANGACC) { for (int k = 0; k < 3; k++) out[k] = ca[k]; }
        else {
          real dif[3], t[3], corr[3];
          for (int k = 0; k < 3; k++) dif[k] = pos[k] - d->subtree_com[3*m->body_rootid[body] + k];
          cross3(t, ca, dif); cross3(corr, v6, v6 + 3);
          for (int k = 0; k < 3; k++) out[k] = ca[3 + k] + t[k] + corr[k];
        }
      } break;
      case B2S_SENS_FRAMEPOS: case B2S_SENS_FRAMEQUAT: case B2S_SENS_FRAMEXAXIS: case B2S_SENS_FRAMEYAXIS: case B2S_SENS_FRAMEZAXIS: {
        obj_frame(m, d, ot, id, &pos, &mat, &body);
        const real *rp = NULL, *rm = NULL; int rb;
        if (m->sensor_refid[s] >= 0) obj_frame(m, d, m->sensor_reftype[s], m->sensor_refid[s], &rp, &rm, &rb);
        if (type == B2S_SENS_FRAMEPOS) {
          if (rp) { real df[3] = {pos[0] - rp[0], pos[1] - rp[1], pos[2] - rp[2]}; matT_vec(out, rm, df); }
          else for (int k = 0; k < 3; k++) out[k] = pos[k];
        } else if (type != B2S_SENS_FRAMEQUAT) {
          const int ax = type - B2S_SENS_FRAMEXAXIS;      /* column 0 / 1 / 2 of the frame */
          real z[3] = {mat[ax], mat[3 + ax], mat[6 + ax]};
          if (rm) matT_vec(out, rm, z); else for (int k = 0; k < 3; k++) out[k] = z[k];
        } else {
          real q[4], R[9];
          if (rm) { for (int a = 0; a < 3; a++) for (int b = 0; b < 3; b++) { real sacc = 0; for (int k = 0; k < 3; k++) sacc += rm[3*k + a]*mat[3*k + b]; R[3*a + b] = sacc; } mat2quat_(q, R); }
          else mat2quat_(q, mat);
          for (int k = 0; k < 4; k++) out[k] = q[k];
        }
      } break;
      case B2S_SENS_FRAMELINVEL: case B2S_SENS_FRAMEANGVEL:
        obj_frame(m, d, ot, id, &pos, &mat, &body);
        object_velocity(m, d, body, pos, NULL, v6);
        for (int k = 0; k < 3; k++) out[k] = type == B2S_SENS_FRAMELINVEL ? v6[3 + k] : v6[k];
        break;
      case B2S_SENS_CONTACT: {
        int bits = m->sensor_intprm[3*s], reduce = m->sensor_intprm[3*s + 1], nslot = m->sensor_intprm[3*s + 2];
        int rt = m->sensor_reftype[s], rid = m->sensor_refid[s];
        int maxm = m->opt_contact_sensor_maxmatch;
        int* mc = (int*)malloc(sizeof(int)*(size_t)(maxm > 0 ? maxm : 1)); real* mdir = (real*)malloc(sizeof(real)*(size_t)(maxm > 0 ? maxm : 1));
        int nmatch = 0;
        for (int c = 0; c < d->ncon[0] && nmatch < maxm; c++) {
          int g1 = d->con_geom[2*c], g2 = d->con_geom[2*c + 1];
          int fwd = side_match(m, ot, id, g1) && (rid < 0 || side_match(m, rt, rid, g2));
          int bwd = side_match(m, ot, id, g2) && (rid < 0 || side_match(m, rt, rid, g1));
          if (!fwd && !bwd) continue;
          mc[nmatch] = c; mdir[nmatch] = fwd ? (real)1 : (real)-1; nmatch++;
        }
        for (int k = 0; k < dim; k++) out[k] = 0;
        int slotdim = nslot > 0 ? dim/nslot : 0;
        if (reduce == B2S_REDUCE_NETFORCE) {
          /* one synthetic contact: force-weighted centroid, net wrench in the GLOBAL frame */
          real F[3] = {0, 0, 0}, T[3] = {0, 0, 0}, P[3] = {0, 0, 0}, wsum = 0;
          for (int i = 0; i < nmatch; i++) {
            real f6[6]; contact_wrench(d, mc[i], f6);
            real w = (real)sqrt((double)dot3(f6, f6));
            for (int k = 0; k < 3; k++) P[k] += w*d->con_pos[3*mc[i] + k];
            wsum += w;
          }
          if (wsum > MINVAL) for (int k = 0; k < 3; k++) P[k] /= wsum;
          else if (nmatch) { for (int k = 0; k < 3; k++) P[k] = 0; for (int i = 0; i < nmatch; i++) for (int k = 0; k < 3; k++) P[k] += d->con_pos[3*mc[i] + k]/(real)nmatch; }
          for (int i = 0; i < nmatch; i++) {
            int c = mc[i]; real f6[6], fg[3], tg[3], arm[3], tx[3];
            contact_wrench(d, c, f6);
            matT_vec(fg, d->con_frame + 9*c, f6); matT_vec(tg, d->con_frame + 9*c, f6 + 3);
            for (int k = 0; k < 3; k++) { fg[k] *= mdir[i]; tg[k] *= mdir[i]; arm[k] = d->con_pos[3*c + k] - P[k]; }
            cross3(tx, arm, fg);
            for (int k = 0; k < 3; k++) { F[k] += fg[k]; T[k] += tg[k] + tx[k]; }
          }
          if (nmatch && nslot > 0) {
            real* o = out;
            if (bits & B2S_CON_FOUND) *o++ = (real)nmatch;
            if (bits & B2S_CON_FORCE) { for (int k = 0; k < 3; k++) o[k] = F[k]; o += 3; }
            if (bits & B2S_CON_TORQUE) { for (int k = 0; k < 3; k++) o[k] = T[k]; o += 3; }
            if (bits & B2S_CON_DIST) *o++ = 0;
            if (bits & B2S_CON_POS) { for (int k = 0; k < 3; k++) o[k] = P[k]; o += 3; }
            if (bits & B2S_CON_NORMAL) { o[0] = 1; o[1] = 0; o[2] = 0; o += 3; }
            if (bits & B2S_CON_TANGENT) { o[0] = 0; o[1] = 1; o[2] = 0; o += 3; }
          }
        } else {
          /* order matches: none = detection order, mindist = ascending dist, maxforce = descending |force| (stable) */
          if (reduce != B2S_REDUCE_NONE) {
            real* key = (real*)malloc(sizeof(real)*(size_t)(nmatch > 0 ? nmatch : 1));
            for (int i = 0; i < nmatch; i++) {
              if (reduce == B2S_REDUCE_MINDIST) key[i] = d->con_dist[mc[i]];
              else { real f6[6]; contact_wrench(d, mc[i], f6); key[i] = -(real)sqrt((double)dot3(f6, f6)); }
            }
            for (int i = 1; i < nmatch; i++) {
              real kk = key[i], dd = mdir[i]; int cc = mc[i], j = i - 1;
              while (j >= 0 && key[j] > kk) { key[j + 1] = key[j]; mc[j + 1] = mc[j]; mdir[j + 1] = mdir[j]; j--; }
              key[j + 1] = kk; mc[j + 1] = cc; mdir[j + 1] = dd;
            }
            free(key);
          }
          for (int sl = 0; sl < nslot && sl < nmatch; sl++) {
            int c = mc[sl]; real dir = mdir[sl], f6[6];
            contact_wrench(d, c, f6);
            real* o = out + sl*slotdim;
            if (bits & B2S_CON_FOUND) *o++ = (real)nmatch;
            if (bits & B2S_CON_FORCE) { o[0] = f6[0]; o[1] = dir*f6[1]; o[2] = dir*f6[2]; o += 3; }
            if (bits & B2S_CON_TORQUE) { o[0] = f6[3]; o[1] = dir*f6[4]; o[2] = dir*f6[5]; o += 3; }
            if (bits & B2S_CON_DIST) *o++ = d->con_dist[c];
            if (bits & B2S_CON_POS) { for (int k = 0; k < 3; k++) o[k] = d->con_pos[3*c + k]; o += 3; }
            if (bits & B2S_CON_NORMAL) { for (int k = 0; k < 3; k++) o[k] = dir*d->con_frame[9*c + k]; o += 3; }
            if (bits & B2S_CON_TANGENT) { for (int k = 0; k < 3; k++) o[k] = dir*d->con_frame[9*c + 3 + k]; o += 3; }
          }
        }
        free(mc); free(mdir);
      } break;
      default: for (int k = 0; k < dim; k++) out[k] = 0;
    }
    real cut = (real)m->sensor_cutoff[s];
    if (cut > 0 && type == B2S_SENS_RANGEFINDER) { if (out[0] > cut) out[0] = cut; }      /* "positive" data type: caps hits only */
    else if (cut > 0 && type != B2S_SENS_CONTACT)
      for (int k = 0; k < dim; k++) out[k] = clipr(out[k], -cut, cut);
  }
  free(cfrc_int);
}

/* -------------------------------------------------------------- integrate */
void orc_integrate(const b2s_model_desc* m, orc_data* d) {
  int nv = m->nv;
  real dt = (real)m->opt_timestep;
  real* qa = (real*)malloc(sizeof(real)*(size_t)(3*nv + 2*nv*nv));
  real* rhs = qa + nv; real* dg = rhs + nv; real* A = dg + nv; real* L = A + nv*nv;
  for (int i = 0; i < nv; i++) { rhs[i] = d->qfrc_smooth[i] + d->qfrc_constraint[i]; dg[i] = 0; qa[i] = d->qacc[i]; }
  int use = 0, ntrm = 0;
  real* tact = (real*)calloc((size_t)(m->nu + 1), sizeof(real));   /* velocity gain of tendon-borne actuators */
  if (m->opt_integrator == B2S_INT_IMPLICITFAST) {
    /* qDeriv = d qfrc_smooth / d qvel: actuator velocity terms (mjd_actuator_vel) + joint damping (mjd_passive_vel) */
    if (!(m->opt_disableflags & B2S_DSBL_ACTUATION))
      for (int u = 0; u < m->nu; u++) {
        real bv = 0;
        if (m->actuator_biastype[u] == B2S_BIAS_AFFINE) bv += (real)m->actuator_biasprm[3*u + 2];
        if (m->actuator_gaintype[u] == B2S_GAIN_AFFINE) {
          real c = d->ctrl[u];
          if (m->actuator_ctrllimited[u] && !(m->opt_disableflags & B2S_DSBL_CLAMPCTRL))
            c = clipr(c, (real)m->actuator_ctrlrange[2*u], (real)m->actuator_ctrlrange[2*u + 1]);
          bv += (real)m->actuator_gainprm[3*u + 2]*c;
        }
        if (bv == 0) continue;
        if (m->actuator_forcelimited[u]) {  /* a force clamped by forcerange has zero velocity derivative */
          real f = d->actuator_force[u];
          if (f <= (real)m->actuator_forcerange[2*u] || f >= (real)m->actuator_forcerange[2*u + 1]) continue;
        }
        real g = (real)m->actuator_gear[u];
        if (m->actuator_trntype[u] == B2S_TRN_TENDON) { tact[u] = g*g*bv; ntrm++; }
        else dg[m->jnt_dofadr[m->actuator_trnid[u]]] += g*g*bv;
      }
    if (!(m->opt_disableflags & B2S_DSBL_PASSIVE)) for (int i = 0; i < nv; i++) dg[i] -= (real)m->dof_damping[i];
    use = 1;
  } else if (m->opt_integrator == B2S_INT_EULER) {
    if (!(m->opt_disableflags & (B2S_DSBL_EULERDAMP | B2S_DSBL_PASSIVE)))
      for (int i = 0; i < nv; i++) if (m->dof_damping[i] > 0) { use = 1; dg[i] = -(real)m->dof_damping[i]; }
  }
  if (use) {
    for (int i = 0; i < nv*nv; i++) A[i] = d->qM[i];
    for (int i = 0; i < nv; i++) A[i*nv + i] -= dt*dg[i];
    if (m->opt_integrator == B2S_INT_IMPLICITFAST && m->ntendon > 0) {
      /* rank-1 terms moment^T s moment through fixed tendons: actuators in order, then tendon dampers */
      for (int u = 0; u < m->nu && ntrm; u++) {
        if (tact[u] == 0) continue;
        int t = m->actuator_trnid[u];
        for (int a = m->tendon_adr[t]; a < m->tendon_adr[t] + m->tendon_num[t]; a++)
          for (int b = m->tendon_adr[t]; b < m->tendon_adr[t] + m->tendon_num[t]; b++)
            A[m->jnt_dofadr[m->wrap_objid[a]]*nv + m->jnt_dofadr[m->wrap_objid[b]]] -= dt*(tact[u]*(real)m->wrap_prm[a]*(real)m->wrap_prm[b]);
      }
      if (!(m->opt_disableflags & B2S_DSBL_PASSIVE))
        for (int t = 0; t < m->ntendon; t++) {
          real bd = (real)m->tendon_damping[t];
          if (bd == 0) continue;
          for (int a = m->tendon_adr[t]; a < m->tendon_adr[t] + m->tendon_num[t]; a++)
            for (int b = m->tendon_adr[t]; b < m->tendon_adr[t] + m->tendon_num[t]; b++)
              A[m->jnt_dofadr[m->wrap_objid[a]]*nv + m->jnt_dofadr[m->wrap_objid[b]]] += dt*(bd*(real)m->wrap_prm[a]*(real)m->wrap_prm[b]);
        }
    }
    chol_factor(L, A, nv);
    chol_solve(qa, L, rhs, nv);
  }
  /* mj_advance: semi-implicit velocity then position update */
  for (int i = 0; i < nv; i++) d->qvel[i] += dt*qa[i];
  for (int j = 0; j < m->njnt; j++) {
    int pa = m->jnt_qposadr[j], da = m->jnt_dofadr[j], t = m->jnt_type[j];
    if (t == B2S_JNT_FREE) { for (int k = 0; k < 3; k++) d->qpos[pa + k] += dt*d->qvel[da + k]; pa += 3; da += 3; }
    if (t == B2S_JNT_FREE || t == B2S_JNT_BALL) {
      real w[3] = {d->qvel[da], d->qvel[da + 1], d->qvel[da + 2]};
      real ang = norm3(w)*dt;
      real q[4] = {d->qpos[pa], d->qpos[pa + 1], d->qpos[pa + 2], d->qpos[pa + 3]}, dq[4], qn[4];
      normalize3(w);
      axis_angle_quat(dq, w, ang);
      normalize4(q);
      quat_mul(qn, q, dq);
      normalize4(qn);
      for (int k = 0; k < 4; k++) d->qpos[pa + k] = qn[k];
    } else d->qpos[pa] += dt*d->qvel[da];
  }
  d->time[0] += dt;
  for (int i = 0; i < nv; i++) d->qacc_warmstart[i] = d->qacc[i];
  free(qa); free(tact);
}

/* -------------------------------------------------------------- top level */
int orc_forward(const b2s_model_desc* m, orc_data* d) {
  if (m->nv > NV_MAX) return -1;
  orc_kinematics(m, d);
  orc_com_pos(m, d);
  orc_crb(m, d);
  orc_factor_m(m, d);
  orc_collision(m, d);
  orc_make_constraint(m, d);
  orc_fwd_velocity(m, d);
  orc_fwd_actuation(m, d);
  orc_fwd_acceleration(m, d);
  orc_solve(m, d);
  orc_sensors(m, d);
  return 0;
}
int orc_step(const b2s_model_desc* m, orc_data* d) {
  int rc = orc_forward(m, d);
  if (rc) return rc;
  orc_integrate(m, d);
  return 0;
}
int orc_rollout(const b2s_model_desc* m, orc_data* worlds, int nworld, int nstep, int nthread) {
  int rc = 0;
#ifdef _OPENMP
#pragma omp parallel for num_threads(nthread > 0 ? nthread : 1) schedule(static)
#endif
  for (int w = 0; w < nworld; w++)
    for (int s = 0; s < nstep; s++) if (orc_step(m, worlds + w)) rc = -1;
  (void)nthread;
  return rc;
}
