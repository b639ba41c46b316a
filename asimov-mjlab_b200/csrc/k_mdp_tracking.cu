// k_mdp_tracking.cu -- manager terms of the motion-tracking task as two fused kernels (BASELINE.json
// configs[3]).  One warp per env; one lane per tracked body (14 for the G1) for the body terms, one lane per joint for the
// joint terms.  Restates tasks/tracking/mdp/rewards.py, terminations.py, observations.py and the metrics / re-anchoring
// of commands.py:219-259,396-415; quaternion helpers follow third_party/isaaclab/isaaclab/utils/math.py
// (quat_error_magnitude, yaw_quat, subtract_frame_transforms, matrix_from_quat).  The torch restatement of the same terms
// (rl/tracking_env.py, use_fused=False) is the test oracle.
#include "b2s_math.cuh"
#include "b2s_api.h"

#define TRK_WARPS 4

struct qf { float w, x, y, z; };
DEV qf ldqf(const float* p) { qf q = {p[0], p[1], p[2], p[3]}; return q; }
DEV qf qfmul(qf a, qf b) {
  qf r = {a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z, a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y,
          a.w * b.y - a.x * b.z + a.y * b.w + a.z * b.x, a.w * b.z + a.x * b.y - a.y * b.x + a.z * b.w};
  return r;
}
DEV qf qfconj(qf q) { qf r = {q.w, -q.x, -q.y, -q.z}; return r; }
DEV qf qfinv(qf q) {
  const float s = 1.0f / fmaxf(q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z, 1e-9f);
  qf r = {q.w * s, -q.x * s, -q.y * s, -q.z * s};
  return r;
}
DEV v3 qfapply(qf q, v3 v) {
  const v3 u = mk3(q.x, q.y, q.z);
  const v3 t = cross(u, v) * 2.0f;
  return v + t * q.w + cross(u, t);
}
DEV v3 qfapply_inv(qf q, v3 v) {
  const v3 u = mk3(q.x, q.y, q.z);
  const v3 t = cross(u, v) * 2.0f;
  return v - t * q.w + cross(u, t);
}
// |axis-angle of a * conj(b)|
DEV float qferr(qf a, qf b) {
  qf q = qfmul(a, qfconj(b));
  if (q.w < 0.f) { q.w = -q.w; q.x = -q.x; q.y = -q.y; q.z = -q.z; }
  const float mag = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z);
  const float half = atan2f(mag, q.w), angle = 2.0f * half;
  const float s = fabsf(angle) > 1e-6f ? sinf(half) / angle : 0.5f - angle * angle / 48.0f;
  return mag / s;
}
DEV qf qfyaw(qf q) {
  const float yaw = atan2f(2.0f * (q.w * q.z + q.x * q.y), 1.0f - 2.0f * (q.y * q.y + q.z * q.z));
  const float c = cosf(0.5f * yaw), s = sinf(0.5f * yaw);
  const float n = 1.0f / fmaxf(sqrtf(c * c + s * s), 1e-9f);
  qf r = {c * n, 0.f, 0.f, s * n};
  return r;
}
// first two columns of R(q), row by row: r00 r01 r10 r11 r20 r21
DEV void qfmat6(qf q, float* o) {
  const float two_s = 2.0f / (q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z);
  o[0] = 1.0f - two_s * (q.y * q.y + q.z * q.z); o[1] = two_s * (q.x * q.y - q.z * q.w);
  o[2] = two_s * (q.x * q.y + q.z * q.w);        o[3] = 1.0f - two_s * (q.x * q.x + q.z * q.z);
  o[4] = two_s * (q.x * q.z - q.y * q.w);        o[5] = two_s * (q.y * q.z + q.x * q.w);
}
DEV float bcast(float v, int src) { return __shfl_sync(FULL, v, src); }

__global__ void __launch_bounds__(32 * TRK_WARPS) k_trk_terms(const b2s_tracking_mdp p) {
  const int lane = threadIdx.x & 31, w = blockIdx.x * TRK_WARPS + (threadIdx.x >> 5);
  if (w >= p.nworld) return;
  const int nb = p.nb, nu = p.nu, a = p.anchor_k;
  const long long ts = p.time_steps[w];
  const float* qpos = p.qpos + (size_t)w * p.nq;
  const float* qvel = p.qvel + (size_t)w * p.nv;
  // ---- body terms: lanes 0..nb-1
  float e_pos = 0.f, e_ori = 0.f, e_lin = 0.f, e_ang = 0.f, m_pos = 0.f, m_rot = 0.f;
  float a_pos2 = 0.f, a_rot = 0.f;
  bool bad_ee = false, bad_z = false, bad_ori = false;
  if (lane < nb) {
    const size_t mo = (size_t)ts * nb + lane, rb = (size_t)w * p.nbody + p.body_ids[lane];
    const v3 bpos = ld3(p.mo_bpos + 3 * mo) + ld3(p.env_origins + 3 * (size_t)w);
    const qf bquat = ldqf(p.mo_bquat + 4 * mo);
    const v3 rpos = ld3(p.xpos + 3 * rb);
    const qf rquat = ldqf(p.xquat + 4 * rb);
    const float* cv = p.cvel + 6 * rb;
    const v3 rang = ld3(cv);
    const v3 rlin = ld3(cv + 3) - cross(rang, ld3(p.subtree_com + 3 * ((size_t)w * p.nbody + p.body_root[lane])) - rpos);
    const v3 relp = ld3(p.body_pos_relative_w + 3 * ((size_t)w * nb + lane));
    const qf relq = ldqf(p.body_quat_relative_w + 4 * ((size_t)w * nb + lane));
    const v3 dp = relp - rpos, dl = ld3(p.mo_blin + 3 * mo) - rlin, da = ld3(p.mo_bang + 3 * mo) - rang;
    e_pos = dot(dp, dp); e_lin = dot(dl, dl); e_ang = dot(da, da);
    m_pos = sqrtf(e_pos);
    m_rot = qferr(relq, rquat); e_ori = m_rot * m_rot;
    bad_ee = p.ee_mask[lane] && fabsf(relp.z - rpos.z) > 0.25f;
    if (lane == a) {
      const v3 d0 = bpos - rpos;
      a_pos2 = dot(d0, d0);
      a_rot = qferr(bquat, rquat);
      bad_z = fabsf(bpos.z - rpos.z) > 0.25f;
      const v3 g = mk3(0.f, 0.f, -1.f);
      bad_ori = fabsf(qfapply_inv(bquat, g).z - qfapply_inv(rquat, g).z) > 0.8f;
    }
  }
  const float inb = 1.0f / (float)nb;
  e_pos = warp_sum(e_pos) * inb; e_ori = warp_sum(e_ori) * inb; e_lin = warp_sum(e_lin) * inb; e_ang = warp_sum(e_ang) * inb;
  m_pos = warp_sum(m_pos) * inb; m_rot = warp_sum(m_rot) * inb;
  a_pos2 = bcast(a_pos2, a); a_rot = bcast(a_rot, a);
  const bool terminated = __any_sync(FULL, bad_ee || bad_z || bad_ori);
  // ---- joint terms: one lane per joint
  float rate = 0.f, lim = 0.f, ejp = 0.f, ejv = 0.f;
  for (int j = lane; j < nu; j += 32) {
    const float q = qpos[7 + j];
    const float d = p.action[(size_t)w * nu + j] - p.prev_action[(size_t)w * nu + j];
    rate += d * d;
    lim += -fminf(q - p.soft_lo[j], 0.f) + fmaxf(q - p.soft_hi[j], 0.f);
    const float dq = p.mo_jp[(size_t)ts * nu + j] - q, dv = p.mo_jv[(size_t)ts * nu + j] - qvel[6 + j];
    ejp += dq * dq; ejv += dv * dv;
  }
  rate = warp_sum(rate); lim = warp_sum(lim); ejp = warp_sum(ejp); ejv = warp_sum(ejv);
  if (lane != 0) return;
  float term[B2S_TRK_NTERM];
  term[B2S_TRK_TERM_ROOT_POS] = expf(-a_pos2 * p.inv_std2[0]);
  term[B2S_TRK_TERM_ROOT_ORI] = expf(-a_rot * a_rot * p.inv_std2[1]);
  term[B2S_TRK_TERM_BODY_POS] = expf(-e_pos * p.inv_std2[2]);
  term[B2S_TRK_TERM_BODY_ORI] = expf(-e_ori * p.inv_std2[3]);
  term[B2S_TRK_TERM_BODY_LIN_VEL] = expf(-e_lin * p.inv_std2[4]);
  term[B2S_TRK_TERM_BODY_ANG_VEL] = expf(-e_ang * p.inv_std2[5]);
  term[B2S_TRK_TERM_ACTION_RATE] = rate;
  term[B2S_TRK_TERM_JOINT_LIMIT] = lim;
  term[B2S_TRK_TERM_SELF_COLLISIONS] = p.sensordata[(size_t)w * p.nsensordata + p.s_self];
  float rew = 0.f;
#pragma unroll
  for (int k = 0; k < B2S_TRK_NTERM; k++) {
    const float val = term[k] * p.weight[k];
    rew += val;
    p.term_sums[(size_t)k * p.nworld + w] += val;
  }
  p.rew[w] = rew;
  const long long ep = p.episode_length_buf[w] + 1;
  p.episode_length_buf[w] = ep;
  const bool time_out = ep >= (long long)p.max_episode_length;
  p.time_outs[w] = time_out; p.terminated[w] = terminated; p.dones[w] = terminated || time_out;
  float* met = p.metrics + w;
  met[(size_t)B2S_TRK_MET_ANCHOR_POS * p.nworld] = sqrtf(a_pos2);
  met[(size_t)B2S_TRK_MET_ANCHOR_ROT * p.nworld] = a_rot;
  met[(size_t)B2S_TRK_MET_BODY_POS * p.nworld] = m_pos;
  met[(size_t)B2S_TRK_MET_BODY_ROT * p.nworld] = m_rot;
  met[(size_t)B2S_TRK_MET_JOINT_POS * p.nworld] = sqrtf(ejp);
  met[(size_t)B2S_TRK_MET_JOINT_VEL * p.nworld] = sqrtf(ejv);
}

__global__ void __launch_bounds__(32 * TRK_WARPS) k_trk_observe(const b2s_tracking_mdp p) {
  const int lane = threadIdx.x & 31, w = blockIdx.x * TRK_WARPS + (threadIdx.x >> 5);
  if (w >= p.nworld) return;
  const int nb = p.nb, nu = p.nu, a = p.anchor_k;
  const long long ts = p.time_steps[w];
  const float* qpos = p.qpos + (size_t)w * p.nq;
  const float* qvel = p.qvel + (size_t)w * p.nv;
  const float* sd = p.sensordata + (size_t)w * p.nsensordata;
  const float* U = p.U + (size_t)w * p.nrand;
  const v3 org = ld3(p.env_origins + 3 * (size_t)w);
  // anchors (every lane: broadcast loads)
  const v3 m_apos = ld3(p.mo_bpos + 3 * ((size_t)ts * nb + a)) + org;
  const qf m_aq = ldqf(p.mo_bquat + 4 * ((size_t)ts * nb + a));
  const size_t ra = (size_t)w * p.nbody + p.body_ids[a];
  const v3 r_apos = ld3(p.xpos + 3 * ra);
  const qf r_aq = ldqf(p.xquat + 4 * ra);
  const qf q10 = qfinv(r_aq);
  float* o = p.obs + (size_t)w * (15 + 5 * nu);
  float* c = p.critic_obs + (size_t)w * (15 + 5 * nu + 9 * nb);
  float* cb = c + 2 * nu + 9;                 // body_pos (3 nb) then body_ori (6 nb)
  // ---- re-anchored motion bodies and the robot's bodies in its anchor frame: one lane per tracked body
  if (lane < nb) {
    const size_t mo = (size_t)ts * nb + lane, rb = (size_t)w * p.nbody + p.body_ids[lane];
    const qf dori = qfyaw(qfmul(r_aq, qfinv(m_aq)));
    const v3 dpos = mk3(r_apos.x, r_apos.y, m_apos.z);
    const qf relq = qfmul(dori, ldqf(p.mo_bquat + 4 * mo));
    const v3 relp = dpos + qfapply(dori, ld3(p.mo_bpos + 3 * mo) + org - m_apos);
    st3(p.body_pos_relative_w + 3 * ((size_t)w * nb + lane), relp);
    float* rq = p.body_quat_relative_w + 4 * ((size_t)w * nb + lane);
    rq[0] = relq.w; rq[1] = relq.x; rq[2] = relq.y; rq[3] = relq.z;
    st3(cb + 3 * lane, qfapply(q10, ld3(p.xpos + 3 * rb) - r_apos));
    qfmat6(qfmul(q10, ldqf(p.xquat + 4 * rb)), cb + 3 * nb + 6 * lane);
  }
  // ---- command, anchor error in the robot's anchor frame, proprioception
  for (int j = lane; j < nu; j += 32) {
    const float jm = p.mo_jp[(size_t)ts * nu + j], vm = p.mo_jv[(size_t)ts * nu + j];
    const float jp = qpos[7 + j] - p.default_joint_pos[j], jv = qvel[6 + j], ac = p.action[(size_t)w * nu + j];
    o[j] = jm; o[nu + j] = vm; c[j] = jm; c[nu + j] = vm;
    o[2 * nu + 15 + j] = jp + (2.0f * U[15 + j] - 1.0f) * 0.01f;
    o[3 * nu + 15 + j] = jv + (2.0f * U[15 + nu + j] - 1.0f) * 0.5f;
    o[4 * nu + 15 + j] = ac;
    float* cp = c + 2 * nu + 9 + 9 * nb + 6;
    cp[j] = jp; cp[nu + j] = jv; cp[2 * nu + j] = ac;
  }
  if (lane == 0) {
    const v3 pos_b = qfapply(q10, m_apos - r_apos);
    float m6[6];
    qfmat6(qfmul(q10, m_aq), m6);
    float* op = o + 2 * nu;
    float* cp = c + 2 * nu;
    const float pb[3] = {pos_b.x, pos_b.y, pos_b.z};
#pragma unroll
    for (int k = 0; k < 3; k++) { op[k] = pb[k] + (2.0f * U[k] - 1.0f) * 0.25f; cp[k] = pb[k]; }
#pragma unroll
    for (int k = 0; k < 6; k++) { op[3 + k] = m6[k] + (2.0f * U[3 + k] - 1.0f) * 0.05f; cp[3 + k] = m6[k]; }
    float* cs = c + 2 * nu + 9 + 9 * nb;
#pragma unroll
    for (int k = 0; k < 3; k++) {
      const float lin = sd[p.s_linvel + k], gy = sd[p.s_gyro + k];
      op[9 + k] = lin + (2.0f * U[9 + k] - 1.0f) * 0.5f; op[12 + k] = gy + (2.0f * U[12 + k] - 1.0f) * 0.2f;
      cs[k] = lin; cs[3 + k] = gy;
    }
  }
}

static int trk_check(const b2s_tracking_mdp* p) {
  if (!p || p->nworld < 1 || p->nu < 1 || p->nb < 1 || p->nb > 32 || p->anchor_k < 0 || p->anchor_k >= p->nb || p->nrand < 15 + 2 * p->nu) return -1;
  return 0;
}
extern "C" int b2s_trk_terms(const b2s_tracking_mdp* p, void* stream) {
  if (trk_check(p)) return -1;
  k_trk_terms<<<(p->nworld + TRK_WARPS - 1) / TRK_WARPS, 32 * TRK_WARPS, 0, (cudaStream_t)stream>>>(*p);
  b2s_count_launch();
  return 0;
}
extern "C" int b2s_trk_observe(const b2s_tracking_mdp* p, void* stream) {
  if (trk_check(p)) return -1;
  k_trk_observe<<<(p->nworld + TRK_WARPS - 1) / TRK_WARPS, 32 * TRK_WARPS, 0, (cudaStream_t)stream>>>(*p);
  b2s_count_launch();
  return 0;
}
