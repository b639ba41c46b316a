// k_mdp.cu -- the manager side of the velocity task's env step as four fused kernels.
//
// The reference evaluates the action / reward / termination / event / command / observation managers as ~500 small
// torch kernels per env step (envs/manager_based_rl_env.py:111-152).  Here one warp owns one env and reads the
// engine's state arrays directly:
//   k_vel_apply_action  action manager: process_action + apply_action          (envs/mdp/actions/joint_actions.py)
//   k_vel_air_time      ContactSensor._update_air_time_tracking, per sub-step  (sensor/contact_sensor.py:327-367)
//   k_vel_post          episode counter, terminations, the 14 reward terms of the G1 velocity task
//                       (tasks/velocity/mdp/rewards.py:23-359, envs/mdp/rewards.py:48-70, terminations.py:24-32),
//                       then reset_root_state_uniform / reset_joints_by_offset of finished envs and the command resample
//   k_vel_observe       command timers + heading control (velocity_command.py:70-125), push_by_setting_velocity,
//                       policy and critic observations with uniform noise (velocity_env_cfg.py:138-193)
// Randomness comes from the caller's uniform buffer U at fixed slots, so the torch restatement of the same terms
// (rl/velocity_env.py, use_fused=False) is the test oracle on identical draws.  State written back into the physics
// (ctrl, reset qpos / qvel) is computed with explicitly rounded multiplies and adds (no FMA contraction) in the
// torch restatement's operation order, so both paths feed bit-identical inputs to the step.
#include "b2s_math.cuh"
#include "b2s_api.h"

#define MDP_WARPS 4
#define MUL(a, b) __fmul_rn((a), (b))
#define ADD(a, b) __fadd_rn((a), (b))
#define SUB(a, b) __fsub_rn((a), (b))

DEV v3 quat_apply_inverse(const float* q, v3 v) {
  const float w = q[0];
  const v3 u = mk3(q[1], q[2], q[3]);
  const v3 t = cross(u, v) * 2.0f;
  return v - t * w + cross(u, t);
}
DEV float uni(float lo, float span, float r) { return ADD(lo, MUL(span, r)); }
DEV float cmd_speed(const float* c) { return sqrtf(c[0] * c[0] + c[1] * c[1]) + fabsf(c[2]); }

__global__ void k_vel_apply_action(const b2s_velocity_mdp p) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= p.nworld * p.nu) return;
  const int j = i % p.nu;
  const float a = p.actions_in[i];
  p.prev_action[i] = p.action[i];
  p.action[i] = a;
  p.ctrl[i] = ADD(p.default_qpos[7 + j], MUL(a, p.action_scale[j]));
}

__global__ void k_vel_air_time(const b2s_velocity_mdp p) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= p.nworld * 2) return;
  const int w = i >> 1, k = i & 1;
  const float t = p.time[w];
  const float elapsed = t - p.air_last_time[w];
  const bool contact = p.sensordata[(size_t)w * p.nsensordata + p.s_found[k]] > 0.f;
  const float ca = p.air_cur[i], cc = p.con_cur[i];
  if (ca > 0.f && contact) p.air_last[i] = ca + elapsed;
  p.air_cur[i] = contact ? 0.f : ca + elapsed;
  if (cc > 0.f && !contact) p.con_last[i] = cc + elapsed;
  p.con_cur[i] = contact ? cc + elapsed : 0.f;
  __syncwarp();                 // both feet of a world sit in one warp: everybody has read last_time
  if (k == 0) p.air_last_time[w] = t;
}

// resample the velocity command of one env (lane 0): uniform ranges, heading / standing flags, timer
DEV void resample_command(const b2s_velocity_mdp& p, int w, const float* u) {
  float* c = p.command + 3 * (size_t)w;
  c[0] = uni(p.cmd_lo[0], p.cmd_span[0], u[0]);
  c[1] = uni(p.cmd_lo[1], p.cmd_span[1], u[1]);
  c[2] = uni(p.cmd_lo[2], p.cmd_span[2], u[2]);
  p.heading_target[w] = uni(p.cmd_lo[3], p.cmd_span[3], u[3]);
  p.is_heading[w] = u[4] <= 0.3f;
  p.is_standing[w] = u[5] <= 0.1f;
  p.cmd_time_left[w] = uni(3.0f, 5.0f, u[6]);
}

__global__ void __launch_bounds__(32 * MDP_WARPS) k_vel_post(const b2s_velocity_mdp p) {
  const int lane = threadIdx.x & 31, w = blockIdx.x * MDP_WARPS + (threadIdx.x >> 5);
  if (w >= p.nworld) return;
  const int nu = p.nu, nq = p.nq, nv = p.nv;
  float* qpos = p.qpos + (size_t)w * nq;
  float* qvel = p.qvel + (size_t)w * nv;
  const float* sd = p.sensordata + (size_t)w * p.nsensordata;
  const float* cmd = p.command + 3 * (size_t)w;
  const float* U = p.U + (size_t)w * p.nrand + 9 + 2 * nu;
  // ---- episode counter and terminations
  const long long ep = p.episode_length_buf[w] + 1;
  const v3 pg = quat_apply_inverse(qpos + 3, mk3(0.f, 0.f, -1.f));
  const bool time_out = ep >= (long long)p.max_episode_length;
  const bool fell = fabsf(acosf(clampf(-pg.z, -1.f, 1.f))) > p.fall_angle;
  const bool done = fell || time_out;
  // ---- reward terms (state before any reset)
  const float speed = cmd_speed(cmd);
  float term[B2S_VEL_NTERM];
  {
    const v3 lin_b = quat_apply_inverse(qpos + 3, ld3(qvel));
    const float ex = cmd[0] - lin_b.x, ey = cmd[1] - lin_b.y;
    term[B2S_VEL_TERM_TRACK_LIN] = expf(-(ex * ex + ey * ey + lin_b.z * lin_b.z) / 0.25f);
    const float ez = cmd[2] - qvel[5];
    term[B2S_VEL_TERM_TRACK_ANG] = expf(-(ez * ez + (qvel[3] * qvel[3] + qvel[4] * qvel[4])) / 0.5f);
    const v3 pgt = quat_apply_inverse(p.xquat + ((size_t)w * p.nbody + p.torso) * 4, mk3(0.f, 0.f, -1.f));
    term[B2S_VEL_TERM_UPRIGHT] = expf(-(pgt.x * pgt.x + pgt.y * pgt.y) / 0.2f);
    const float* tv = p.cvel + ((size_t)w * p.nbody + p.torso) * 6;
    term[B2S_VEL_TERM_BODY_ANG_VEL] = tv[0] * tv[0] + tv[1] * tv[1];
    const float* am = sd + p.s_angmom;
    term[B2S_VEL_TERM_ANGMOM] = am[0] * am[0] + am[1] * am[1] + am[2] * am[2];
    term[B2S_VEL_TERM_SELF_COLLISIONS] = sd[p.s_self];
  }
  {
    // joint-wise terms: one lane per joint
    const float* std = speed < 0.05f ? p.std_standing : (speed < 1.5f ? p.std_walking : p.std_running);
    float pose = 0.f, lim = 0.f, rate = 0.f;
    for (int j = lane; j < nu; j += 32) {
      const float q = qpos[7 + j];
      const float d = q - p.default_qpos[7 + j], s = std[j];
      pose += d * d / (s * s);
      lim += -fminf(q - p.soft_lo[j], 0.f) + fmaxf(q - p.soft_hi[j], 0.f);
      const float da = p.action[(size_t)w * nu + j] - p.prev_action[(size_t)w * nu + j];
      rate += da * da;
    }
    term[B2S_VEL_TERM_POSE] = expf(-warp_sum(pose) / (float)nu);
    term[B2S_VEL_TERM_DOF_LIMITS] = warp_sum(lim);
    term[B2S_VEL_TERM_ACTION_RATE] = warp_sum(rate);
  }
  {
    // foot terms (2 feet, evaluated by every lane)
    float air = 0.f, clear = 0.f, swing = 0.f, slip = 0.f, land = 0.f;
    float peaks[2];
#pragma unroll
    for (int k = 0; k < 2; k++) {
      const float found = sd[p.s_found[k]];
      const float* f = sd + p.s_force[k];
      const float* sp = p.site_xpos + ((size_t)w * p.nsite + p.feet_site[k]) * 3;
      const float* cv = p.cvel + ((size_t)w * p.nbody + p.feet_body[k]) * 6;
      const v3 off = ld3(sp) - ld3(p.subtree_com + ((size_t)w * p.nbody + p.feet_root[k]) * 3);
      const v3 vel = ld3(cv + 3) + cross(ld3(cv), off);
      const float vxy = sqrtf(vel.x * vel.x + vel.y * vel.y);
      const float cc = p.con_cur[2 * (size_t)w + k], ca = p.air_cur[2 * (size_t)w + k];
      const bool landed = cc > 0.f && cc < p.landed_dt;
      air += (ca > 0.05f && ca < 0.5f) ? 1.f : 0.f;
      clear += fabsf(sp[2] - 0.1f) * vxy;
      float pk = p.peak_heights[2 * (size_t)w + k];
      if (found == 0.f) pk = fmaxf(pk, sp[2]);
      const float err = pk / 0.1f - 1.0f;
      swing += landed ? err * err : 0.f;
      peaks[k] = landed ? 0.f : pk;
      slip += found > 0.f ? vxy * vxy : 0.f;
      land += landed ? sqrtf(f[0] * f[0] + f[1] * f[1] + f[2] * f[2]) : 0.f;
    }
    const float moving = speed > 0.05f ? 1.f : 0.f;
    term[B2S_VEL_TERM_AIR_TIME] = air * (speed > 0.5f ? 1.f : 0.f);
    term[B2S_VEL_TERM_FOOT_CLEARANCE] = clear * moving;
    term[B2S_VEL_TERM_FOOT_SWING] = swing * moving;
    term[B2S_VEL_TERM_FOOT_SLIP] = slip * moving;
    term[B2S_VEL_TERM_SOFT_LANDING] = land * moving;
    __syncwarp();
    if (lane < 2) p.peak_heights[2 * (size_t)w + lane] = done ? 0.f : peaks[lane];
  }
  float rew = 0.f;
#pragma unroll
  for (int k = 0; k < B2S_VEL_NTERM; k++) {
    if (p.weight[k] != 0.f) {
      const float val = term[k] * p.weight[k];
      rew += val;
      if (lane == 0) p.term_sums[(size_t)k * p.nworld + w] += val;
    }
  }
  __syncwarp();
  if (lane == 0) {
    p.rew[w] = rew; p.dones[w] = done; p.time_outs[w] = time_out;
    p.episode_length_buf[w] = done ? 0 : ep;
  }
  if (!done) return;
  // ---- masked reset: root pose uniform around the env origin with a random yaw, joints at the keyframe, zero velocity
  for (int j = 7 + lane; j < nq; j += 32) qpos[j] = p.default_qpos[j];
  for (int j = lane; j < nv; j += 32) { qvel[j] = 0.f; p.qacc_warmstart[(size_t)w * nv + j] = 0.f; }
  for (int j = lane; j < nu; j += 32) {
    p.ctrl[(size_t)w * nu + j] = p.default_qpos[7 + j];
    p.action[(size_t)w * nu + j] = 0.f; p.prev_action[(size_t)w * nu + j] = 0.f;
  }
  if (lane < 2) {
    const size_t i = 2 * (size_t)w + lane;
    p.air_cur[i] = 0.f; p.air_last[i] = 0.f; p.con_cur[i] = 0.f; p.con_last[i] = 0.f;
  }
  if (lane == 0) {
    const float* org = p.env_origins + 3 * (size_t)w;
    const float* dq = p.default_qpos;
    qpos[0] = ADD(dq[0], ADD(uni(-0.5f, 1.0f, U[0]), org[0]));
    qpos[1] = ADD(dq[1], ADD(uni(-0.5f, 1.0f, U[1]), org[1]));
    qpos[2] = ADD(dq[2], org[2]);
    const float yaw = uni(-3.14f, 6.28f, U[2]);
    const float bw = cosf(MUL(0.5f, yaw)), bz = sinf(MUL(0.5f, yaw));
    // quat_mul(default, (bw, 0, 0, bz)) in the restatement's term order
    const float aw = dq[3], ax = dq[4], ay = dq[5], az = dq[6];
    qpos[3] = SUB(SUB(SUB(MUL(aw, bw), MUL(ax, 0.f)), MUL(ay, 0.f)), MUL(az, bz));
    qpos[4] = SUB(ADD(ADD(MUL(aw, 0.f), MUL(ax, bw)), MUL(ay, bz)), MUL(az, 0.f));
    qpos[5] = ADD(ADD(SUB(MUL(aw, 0.f), MUL(ax, bz)), MUL(ay, bw)), MUL(az, 0.f));
    qpos[6] = ADD(SUB(ADD(MUL(aw, bz), MUL(ax, 0.f)), MUL(ay, 0.f)), MUL(az, bw));
    p.push_time_left[w] = uni(1.0f, 2.0f, U[3]);
    resample_command(p, w, U + 4);
  }
}

__global__ void __launch_bounds__(32 * MDP_WARPS) k_vel_observe(const b2s_velocity_mdp p) {
  const int lane = threadIdx.x & 31, w = blockIdx.x * MDP_WARPS + (threadIdx.x >> 5);
  if (w >= p.nworld) return;
  const int nu = p.nu, nq = p.nq, nv = p.nv;
  const float* qpos = p.qpos + (size_t)w * nq;
  float* qvel = p.qvel + (size_t)w * nv;
  const float* sd = p.sensordata + (size_t)w * p.nsensordata;
  const float* Un = p.U + (size_t)w * p.nrand;
  const float* U = Un + 9 + 2 * nu;
  float* cmd = p.command + 3 * (size_t)w;
  if (lane == 0) {
    if (p.dones[w]) p.air_last_time[w] = p.time[w];
    // command manager: timer-driven resample, heading control, standing envs
    const float tl = SUB(p.cmd_time_left[w], p.step_dt);
    p.cmd_time_left[w] = tl;
    if (tl <= 0.f) resample_command(p, w, U + 11);
    const float* q = qpos + 3;
    const float heading = atan2f(2.0f * (q[1] * q[2] + q[0] * q[3]), 1.0f - 2.0f * (q[2] * q[2] + q[3] * q[3]));
    const float two_pi = 6.283185307179586f;
    float r = fmodf(p.heading_target[w] - heading + 3.141592653589793f, two_pi);
    if (r != 0.f && r < 0.f) r += two_pi;
    const float err = r - 3.141592653589793f;
    if (p.is_heading[w]) cmd[2] = clampf(0.5f * err, p.cmd_lo[2], p.cmd_hi[2]);
    if (p.is_standing[w]) { cmd[0] = 0.f; cmd[1] = 0.f; cmd[2] = 0.f; }
    // interval event: push_by_setting_velocity
    const float pl = SUB(p.push_time_left[w], p.step_dt);
    if (pl <= 0.f) {
      qvel[0] = ADD(qvel[0], SUB(U[18], 0.5f));
      qvel[1] = ADD(qvel[1], SUB(U[19], 0.5f));
      p.push_time_left[w] = uni(1.0f, 2.0f, U[20]);
    } else p.push_time_left[w] = pl;
  }
  __syncwarp();
  // ---- observations: policy (noisy) and critic (clean + privileged)
  float* o = p.obs + (size_t)w * (12 + 3 * nu);
  float* c = p.critic_obs + (size_t)w * (12 + 3 * nu + 12);
  const v3 pg = quat_apply_inverse(qpos + 3, mk3(0.f, 0.f, -1.f));
  if (lane < 3) {
    const float lin = sd[p.s_linvel + lane], gy = sd[p.s_gyro + lane];
    const float g = lane == 0 ? pg.x : (lane == 1 ? pg.y : pg.z);
    o[lane] = lin + (2.0f * Un[lane] - 1.0f) * 0.5f;
    o[3 + lane] = (gy + (2.0f * Un[3 + lane] - 1.0f) * 0.2f) * 0.25f;
    o[6 + lane] = g + (2.0f * Un[6 + lane] - 1.0f) * 0.05f;
    c[lane] = lin; c[3 + lane] = gy * 0.25f; c[6 + lane] = g;
    o[9 + 3 * nu + lane] = cmd[lane]; c[9 + 3 * nu + lane] = cmd[lane];
  }
  for (int j = lane; j < nu; j += 32) {
    const float jp = qpos[7 + j] - p.default_qpos[7 + j], jv = qvel[6 + j], a = p.action[(size_t)w * nu + j];
    o[9 + j] = jp + (2.0f * Un[9 + j] - 1.0f) * 0.01f;
    o[9 + nu + j] = (jv + (2.0f * Un[9 + nu + j] - 1.0f) * 1.5f) * 0.05f;
    o[9 + 2 * nu + j] = a;
    c[9 + j] = jp; c[9 + nu + j] = jv * 0.05f; c[9 + 2 * nu + j] = a;
  }
  float* cx = c + 12 + 3 * nu;
  if (lane < 2) {
    cx[lane] = p.site_xpos[((size_t)w * p.nsite + p.feet_site[lane]) * 3 + 2];
    cx[2 + lane] = p.air_cur[2 * (size_t)w + lane];
    cx[4 + lane] = sd[p.s_found[lane]] > 0.f ? 1.f : 0.f;
  }
  if (lane < 6) {
    const float f = sd[p.s_force[lane / 3] + lane % 3];
    cx[6 + lane] = (f > 0.f ? 1.f : (f < 0.f ? -1.f : 0.f)) * log1pf(fabsf(f));
  }
}

static int check(const b2s_velocity_mdp* p) {
  if (!p || p->nworld < 1 || p->nu < 1 || p->nrand < 9 + 2 * p->nu + 21) return -1;
  return 0;
}
extern "C" int b2s_vel_apply_action(const b2s_velocity_mdp* p, void* stream) {
  if (check(p)) return -1;
  const int n = p->nworld * p->nu;
  k_vel_apply_action<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(*p);
  b2s_count_launch();
  return 0;
}
extern "C" int b2s_vel_air_time(const b2s_velocity_mdp* p, void* stream) {
  if (check(p)) return -1;
  const int n = p->nworld * 2;
  k_vel_air_time<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(*p);
  b2s_count_launch();
  return 0;
}
extern "C" int b2s_vel_post(const b2s_velocity_mdp* p, void* stream) {
  if (check(p)) return -1;
  k_vel_post<<<(p->nworld + MDP_WARPS - 1) / MDP_WARPS, 32 * MDP_WARPS, 0, (cudaStream_t)stream>>>(*p);
  b2s_count_launch();
  return 0;
}
extern "C" int b2s_vel_observe(const b2s_velocity_mdp* p, void* stream) {
  if (check(p)) return -1;
  k_vel_observe<<<(p->nworld + MDP_WARPS - 1) / MDP_WARPS, 32 * MDP_WARPS, 0, (cudaStream_t)stream>>>(*p);
  b2s_count_launch();
  return 0;
}
