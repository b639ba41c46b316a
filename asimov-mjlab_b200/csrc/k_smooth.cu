// k_smooth.cu -- position and smooth-dynamics stages, one warp per world.
//
//   k_position : kinematics + com_pos          
//   k_dynamics : crb, com_vel, rne, passive, actuation, qfrc_smooth,
//                in-shared-memory Cholesky of M, qacc_smooth   (rows K4, K9-K12)
//
// Tree recursions are turned into independent per-body sums: every spatial
// quantity is expressed about the root's subtree COM in world orientation, so
// cvel/cacc of a body is a sum over its ancestor dofs (body_chain) and
// composite inertia / bias force of a body is a sum over the contiguous id
// range of its subtree (bodies are stored in DFS pre-order).  No level-by-level
// launches, no atomics: lanes own bodies/dofs and meet only at __syncwarp().
#include "b2s_math.cuh"
#include "b2s_api.h"

extern __shared__ __align__(16) float b2s_smem[];

// ---------------------------------------------------------------------------------------------- k_position
// Shared memory per warp: body frames 22 nb | local poses 7 nb | joint locals 12 njnt | geom write-out stage 32 x 12.
// Everything a body contributes that does NOT depend on its parent -- its pose in the parent frame, the joint offsets, the
// joint rotation from qpos (sincos) -- is computed by one lane per body / joint BEFORE the walk down the tree, so the
// sequential part (one step per tree level: 11 for the G1) reads shared memory only; it used to chase
// level_body -> body_parentid -> body_jntadr -> jnt_qposadr -> qpos through global memory on every level (ncu r2d: 19 % of the
// kernel's samples were long-scoreboard stalls inside the level loop).  The arithmetic and its order are unchanged.
#define POS_STAGE_LD 13
__host__ __device__ inline int pos_smem_floats(int nb, int njnt) { return 22 * nb + 7 * nb + 12 * njnt + nb + 32 * POS_STAGE_LD; }
__global__ void __launch_bounds__(32 * B2S_WARPS_PER_CTA) k_position(const ModelDev m, const DataDev d) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int w = blockIdx.x * B2S_WARPS_PER_CTA + warp;
  // first kernel of step / forward: reset the solver's scheduling queues (k_constraint refills them)
  if (blockIdx.x == 0 && threadIdx.x < B2S_SOLVE_CLASSES) d.solve_count[threadIdx.x] = 0;
  if (blockIdx.x == 0 && threadIdx.x < 3 && m.col_defer) d.cvx_ntask[threadIdx.x] = 0;      // and the convex-pair task list (k_collision phase 1 refills it)
  if (w >= d.nworld) return;
  const int nb = m.nbody;
  float* sx = b2s_smem + (size_t)warp * pos_smem_floats(nb, m.njnt);   // xpos
  float* sq = sx + 3 * nb;                          // xquat
  float* sm = sq + 4 * nb;                          // xmat
  float* sip = sm + 9 * nb;                         // xipos
  float* scom = sip + 3 * nb;                       // subtree_com
  float* slp = scom + 3 * nb;                       // pose in the parent frame: pos 3, quat 4
  float* sj = slp + 7 * nb;                         // per joint: jpos 3, jaxis 3, (hinge / ball: local rotation 4 | slide: displacement 1 | free: pos 3 + quat 4 over jpos..)
  float* smx = sj + 12 * m.njnt;                    // body masses (the subtree-COM loop reads shared memory only)
  float* stg = smx + nb;                        // geom / site write-out stage
  const float* qpos = d.qpos + (size_t)w * m.nq;
  const float* qpos0 = m.qpos0.at(w);
  const float* bpos = m.body_pos.at(w);
  const float* bquat = m.body_quat.at(w);
  if (lane == 0) {
    sx[0] = sx[1] = sx[2] = 0.f; sq[0] = 1.f; sq[1] = sq[2] = sq[3] = 0.f;
    for (int k = 0; k < 9; k++) sm[k] = (k % 4 == 0) ? 1.f : 0.f;
  }
  for (int b = 1 + lane; b < nb; b += 32) {
    v3 lp; q4 lq;
    const int mid = m.body_mocapid[b];
    if (mid >= 0) {
      lp = ld3(d.mocap_pos + ((size_t)w * m.nmocap + mid) * 3);
      lq = qnormalize(ldq(d.mocap_quat + ((size_t)w * m.nmocap + mid) * 4));
    } else { lp = ld3(bpos + 3 * b); lq = ldq(bquat + 4 * b); }
    st3(slp + 7 * b, lp); stq(slp + 7 * b + 3, lq);
  }
  for (int j = lane; j < m.njnt; j += 32) {
    const int qa = m.jnt_qposadr[j], t = m.jnt_type[j];
    float* o = sj + 12 * j;
    if (t == B2S_JNT_FREE) {
      st3(o, ld3(qpos + qa)); stq(o + 3, qnormalize(ldq(qpos + qa + 3))); st3(o + 7, ld3(m.jnt_axis + 3 * j));
      continue;
    }
    const v3 jaxis = ld3(m.jnt_axis + 3 * j);
    st3(o, ld3(m.jnt_pos + 3 * j)); st3(o + 3, jaxis);
    if (t == B2S_JNT_SLIDE) o[6] = qpos[qa] - qpos0[qa];
    else stq(o + 6, (t == B2S_JNT_HINGE) ? axis_angle(jaxis, qpos[qa] - qpos0[qa]) : qnormalize(ldq(qpos + qa)));
  }
  for (int L = 1; L < m.nlevel; L++) {
    __syncwarp();
    for (int idx = m.level_adr[L] + lane; idx < m.level_adr[L + 1]; idx += 32) {
      const int b = m.level_body[idx], p = m.body_parentid[b];
      const v3 lp = ld3(slp + 7 * b); const q4 lq = ldq(slp + 7 * b + 3);
      m33 pm = ldm(sm + 9 * p);
      v3 pos = ld3(sx + 3 * p) + mulv(pm, lp);
      q4 quat = qmul(ldq(sq + 4 * p), lq);
      const int jn = m.body_jntnum[b], ja = m.body_jntadr[b];
      for (int jj = 0; jj < jn; jj++) {
        const int j = ja + jj, t = m.jnt_type[j];
        const float* o = sj + 12 * j;
        float* xan = d.xanchor + ((size_t)w * m.njnt + j) * 3;
        float* xax = d.xaxis + ((size_t)w * m.njnt + j) * 3;
        if (t == B2S_JNT_FREE) {
          pos = ld3(o);
          quat = ldq(o + 3);
          st3(xan, pos); st3(xax, mulv(q2m(quat), ld3(o + 7)));
          continue;
        }
        const v3 jpos = ld3(o), jaxis = ld3(o + 3);
        m33 R = q2m(quat);
        v3 anchor = pos + mulv(R, jpos), axis = mulv(R, jaxis);
        st3(xan, anchor); st3(xax, axis);
        if (t == B2S_JNT_SLIDE) pos = pos + axis * o[6];
        else {
          quat = qmul(quat, ldq(o + 6));
          pos = anchor - mulv(q2m(quat), jpos);
        }
      }
      quat = qnormalize(quat);
      st3(sx + 3 * b, pos); stq(sq + 4 * b, quat); stm(sm + 9 * b, q2m(quat));
    }
  }
  __syncwarp();
  // inertial frames
  const float* mass = m.body_mass.at(w);
  {
    const float* ipos = m.body_ipos.at(w);
    const float* iquat = m.body_iquat.at(w);
    for (int b = lane; b < nb; b += 32) {
      m33 R = ldm(sm + 9 * b);
      v3 xi = ld3(sx + 3 * b) + mulv(R, ld3(ipos + 3 * b));
      st3(sip + 3 * b, xi);
      smx[b] = mass[b];
      st3(d.xipos + ((size_t)w * nb + b) * 3, xi);
      stm(d.ximat + ((size_t)w * nb + b) * 9, q2m(qmul(ldq(sq + 4 * b), ldq(iquat + 4 * b))));
    }
  }
  // coalesced write-out of body frames
  for (int i = lane; i < 3 * nb; i += 32) d.xpos[(size_t)w * 3 * nb + i] = sx[i];
  for (int i = lane; i < 4 * nb; i += 32) d.xquat[(size_t)w * 4 * nb + i] = sq[i];
  for (int i = lane; i < 9 * nb; i += 32) d.xmat[(size_t)w * 9 * nb + i] = sm[i];
  // geoms and sites: 32 frames per pass through the stage; a pass whose ids are consecutive (the usual case: the robot's geoms)
  // leaves as 3 x 32 + 9 x 32 contiguous floats instead of 12 scattered 4-byte stores per lane
  {
    const float* gpos = m.geom_pos.at(w);
    const float* gquat = m.geom_quat.at(w);
    for (int k0 = 0; k0 < m.nposegeom; k0 += 32) {
      const int k = k0 + lane, cnt = min(32, m.nposegeom - k0);
      const int g0 = m.pose_geom[k0];
      int g = -1;
      if (k < m.nposegeom) {
        g = m.pose_geom[k];
        const int b = m.geom_bodyid[g];
        st3(stg + POS_STAGE_LD * lane, ld3(sx + 3 * b) + mulv(ldm(sm + 9 * b), ld3(gpos + 3 * g)));
        stm(stg + POS_STAGE_LD * lane + 3, q2m(qmul(ldq(sq + 4 * b), ldq(gquat + 4 * g))));
      }
      const bool contig = __all_sync(FULL, g < 0 || g == g0 + lane);
      __syncwarp();
      if (contig) {
        float* op = d.geom_xpos + ((size_t)w * m.ngeom + g0) * 3;
        float* om = d.geom_xmat + ((size_t)w * m.ngeom + g0) * 9;
        for (int i = lane; i < 3 * cnt; i += 32) op[i] = stg[POS_STAGE_LD * (i / 3) + i % 3];
        for (int i = lane; i < 9 * cnt; i += 32) om[i] = stg[POS_STAGE_LD * (i / 9) + 3 + i % 9];
      } else if (g >= 0) {
        for (int i = 0; i < 3; i++) d.geom_xpos[((size_t)w * m.ngeom + g) * 3 + i] = stg[POS_STAGE_LD * lane + i];
        for (int i = 0; i < 9; i++) d.geom_xmat[((size_t)w * m.ngeom + g) * 9 + i] = stg[POS_STAGE_LD * lane + 3 + i];
      }
      __syncwarp();
    }
    const float* spos = m.site_pos.at(w);
    const float* squat = m.site_quat.at(w);
    for (int k = lane; k < m.nposesite; k += 32) {
      const int s = m.pose_site[k];
      const int b = m.site_bodyid[s];
      const size_t si = b2s_site_index(m, d, w, s);
      st3(d.site_xpos + si * 3, ld3(sx + 3 * b) + mulv(ldm(sm + 9 * b), ld3(spos + 3 * s)));
      stm(d.site_xmat + si * 9, q2m(qmul(ldq(sq + 4 * b), ldq(squat + 4 * s))));
    }
  }
  __syncwarp();
  // subtree COM: direct sum over the subtree's contiguous id range (masses staged above: the loop reads shared memory only;
  // the multiply-add stays ONE fused operation per term: the parity tests resolve a product rounded on its own)
  for (int b = lane; b < nb; b += 32) {
    v3 acc = mk3(0.f, 0.f, 0.f);
    const int e = m.body_subtree_end[b];
#pragma unroll 4
    for (int j = b; j < e; j++) acc = acc + ld3(sip + 3 * j) * smx[j];
    const float smass = m.body_subtreemass[b];
    v3 com = (smass < B2S_MINVAL) ? ld3(sip + 3 * b) : acc * (1.0f / smass);
    st3(scom + 3 * b, com);
    st3(d.subtree_com + ((size_t)w * nb + b) * 3, com);
  }
  __syncwarp();
  // cinert: body inertia about the root subtree COM, world orientation
  {
    const float* inertia = m.body_inertia.at(w);
    for (int b = lane; b < nb; b += 32) {
      m33 R = ldm(d.ximat + ((size_t)w * nb + b) * 9);   // written by this lane above
      const float i0 = inertia[3 * b], i1 = inertia[3 * b + 1], i2 = inertia[3 * b + 2], ms = mass[b];
      v3 off = ld3(sip + 3 * b) - ld3(scom + 3 * m.body_rootid[b]);
      float* c = d.cinert + ((size_t)w * nb + b) * 10;
      const float* r = R.m;
      float Ixx = r[0] * i0 * r[0] + r[1] * i1 * r[1] + r[2] * i2 * r[2];
      float Iyy = r[3] * i0 * r[3] + r[4] * i1 * r[4] + r[5] * i2 * r[5];
      float Izz = r[6] * i0 * r[6] + r[7] * i1 * r[7] + r[8] * i2 * r[8];
      float Ixy = r[0] * i0 * r[3] + r[1] * i1 * r[4] + r[2] * i2 * r[5];
      float Ixz = r[0] * i0 * r[6] + r[1] * i1 * r[7] + r[2] * i2 * r[8];
      float Iyz = r[3] * i0 * r[6] + r[4] * i1 * r[7] + r[5] * i2 * r[8];
      float o2 = dot(off, off);
      c[0] = Ixx + ms * (o2 - off.x * off.x); c[1] = Iyy + ms * (o2 - off.y * off.y); c[2] = Izz + ms * (o2 - off.z * off.z);
      c[3] = Ixy - ms * off.x * off.y; c[4] = Ixz - ms * off.x * off.z; c[5] = Iyz - ms * off.y * off.z;
      c[6] = ms * off.x; c[7] = ms * off.y; c[8] = ms * off.z; c[9] = ms;
    }
  }
  // cdof: joint motion axes as spatial vectors about the root subtree COM
  for (int j = lane; j < m.njnt; j += 32) {
    const int b = m.jnt_bodyid[j], t = m.jnt_type[j];
    int da = m.jnt_dofadr[j];
    v3 off = ld3(scom + 3 * m.body_rootid[b]) - ld3(d.xanchor + ((size_t)w * m.njnt + j) * 3);
    float* cd = d.cdof + ((size_t)w * m.nv + da) * 6;
    if (t == B2S_JNT_FREE) {
      for (int a = 0; a < 3; a++) { for (int k = 0; k < 6; k++) cd[6 * a + k] = 0.f; cd[6 * a + 3 + a] = 1.f; }
      cd += 18;
    }
    if (t == B2S_JNT_FREE || t == B2S_JNT_BALL) {
      m33 R = ldm(sm + 9 * b);
      for (int a = 0; a < 3; a++) { v3 ax = col(R, a); st3(cd + 6 * a, ax); st3(cd + 6 * a + 3, cross(ax, off)); }
    } else {
      v3 ax = ld3(d.xaxis + ((size_t)w * m.njnt + j) * 3);
      if (t == B2S_JNT_SLIDE) { st3(cd, mk3(0.f, 0.f, 0.f)); st3(cd + 3, ax); }
      else { st3(cd, ax); st3(cd + 3, cross(ax, off)); }
    }
  }
}

// Frames of geoms and sites on static (world-welded, joint-less, non-mocap) bodies: they never move, so they are
// computed once (put_data, b2s_refresh_static) from the body frames k_position left in xpos/xquat.
__global__ void k_static_frames(const ModelDev m, const DataDev d) {
  // static geoms of every world, then static sites: of every world (dense layout) or once (split layout: the frames of world 0)
  const long long ng = (long long)d.nworld * m.nstatgeom;
  const long long total = ng + (long long)(m.site_split ? 1 : d.nworld) * m.nstatsite;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    if (idx < ng) {
      const int w = (int)(idx / m.nstatgeom), k = (int)(idx - (long long)w * m.nstatgeom);
      const int g = m.stat_geom[k], b = m.geom_bodyid[g];
      const v3 bp = ld3(d.xpos + ((size_t)w * m.nbody + b) * 3);
      const q4 bq = ldq(d.xquat + ((size_t)w * m.nbody + b) * 4);
      st3(d.geom_xpos + ((size_t)w * m.ngeom + g) * 3, bp + mulv(q2m(bq), ld3(m.geom_pos.at(w) + 3 * g)));
      stm(d.geom_xmat + ((size_t)w * m.ngeom + g) * 9, q2m(qmul(bq, ldq(m.geom_quat.at(w) + 4 * g))));
    } else {
      const long long r = idx - ng;
      const int w = (int)(r / m.nstatsite), k = (int)(r - (long long)w * m.nstatsite);
      const int s = m.stat_site[k], b = m.site_bodyid[s];
      const v3 bp = ld3(d.xpos + ((size_t)w * m.nbody + b) * 3);
      const q4 bq = ldq(d.xquat + ((size_t)w * m.nbody + b) * 4);
      const size_t si = b2s_site_index(m, d, w, s);
      st3(d.site_xpos + si * 3, bp + mulv(q2m(bq), ld3(m.site_pos.at(w) + 3 * s)));
      stm(d.site_xmat + si * 9, q2m(qmul(bq, ldq(m.site_quat.at(w) + 4 * s))));
    }
  }
}
void launch_static_frames(const ModelDev& m, const DataDev& d, cudaStream_t st) {
  const long long total = (long long)d.nworld * m.nstatgeom + (long long)(m.site_split ? 1 : d.nworld) * m.nstatsite;
  if (total == 0) return;
  const int grid = (int)((total + 255) / 256 > 148 * 16 ? 148 * 16 : (total + 255) / 256);
  k_static_frames<<<grid, 256, 0, st>>>(m, d);
  b2s_count_launch();
}

// ---------------------------------------------------------------------------------------------- k_dynamics
// shared memory per warp (floats): Cholesky scratch line 2*RC_LINE | cdof 6nv | A: cinert 10nb, later cvel 6nb + cdof_dot 6nv |
// B: crb 10nb, later cfrc 6nb + fext 6nb | M as a packed triangle N(N+1)/2 | invd 7T | vec 4*nvp.  The two aliased regions and
// the packed matrix bring a G1 world to 7.7 KB: 7 CTAs of 4 warps per SM, so 4096 worlds are ONE wave (was 11.4 KB, 1.73 waves).
__host__ __device__ inline int dyn_regA(int nv, int nb) { return 10 * nb > 6 * nb + 6 * nv ? 10 * nb : 6 * nb + 6 * nv; }
__host__ __device__ inline int dyn_regB(int nb) { return 12 * nb; }
__host__ __device__ inline int dyn_tri(int T) { return (b2s_hdim(T) * (b2s_hdim(T) + 1) / 2 + 3) & ~3; }
__host__ __device__ inline int dyn_smem_floats(int nv, int nvp, int nb, int T) {
  return (2 * RC_LINE + 6 * nv + dyn_regA(nv, nb) + dyn_regB(nb) + dyn_tri(T) + 7 * T + 4 * nvp + 3) & ~3;
}

template <int T>
__global__ void __launch_bounds__(32 * B2S_WARPS_PER_CTA) k_dynamics(const ModelDev m, const DataDev d) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int w = blockIdx.x * B2S_WARPS_PER_CTA + warp;
  if (w >= d.nworld) return;
  const int nb = m.nbody, nv = m.nv, nvp = m.nvp, ld = b2s_hld(T);
  float* s_col = b2s_smem + (size_t)warp * dyn_smem_floats(nv, nvp, nb, T);     // 16-byte aligned
  float* s_cdof = s_col + 2 * RC_LINE;
  float* s_cin = s_cdof + 6 * nv;            // region A
  float* s_cvel = s_cin;
  float* s_cdd = s_cvel + 6 * nb;
  float* s_crb = s_cin + dyn_regA(nv, nb);   // region B
  float* s_cfrc = s_crb;
  float* s_fext = s_cfrc + 6 * nb;
  float* s_M = s_crb + dyn_regB(nb);         // packed triangle: entry (i, j), j <= i, at i (i + 1) / 2 + j
  float* s_invd = s_M + dyn_tri(T);
  float* s_qvel = s_invd + 7 * T;
  float* s_act = s_qvel + nvp;      // qfrc_actuator
  float* s_rhs = s_act + nvp;       // qfrc_smooth
  float* s_tmp = s_rhs + nvp;
  const float* qvel = d.qvel + (size_t)w * nv;
  const float* qpos = d.qpos + (size_t)w * m.nq;
  // stage cdof, cinert, qvel
#pragma unroll 4
  for (int i = lane; i < 6 * nv; i += 32) s_cdof[i] = d.cdof[(size_t)w * 6 * nv + i];
  for (int i = lane; i < nv; i += 32) { const float v = qvel[i]; s_qvel[i] = v; s_act[i] = 0.f; d.snap_qvel[(size_t)w * nv + i] = v; }
  for (int i = lane; i < m.nq; i += 32) d.snap_qpos[(size_t)w * m.nq + i] = qpos[i];
  if (lane == 0) d.snap_time[w] = d.time[w];
  for (int i = lane; i < dyn_tri(T); i += 32) s_M[i] = 0.f;
  __syncwarp();
  // composite inertia: sum of cinert over the subtree range (same frame for all bodies).  cinert is staged once,
  // coalesced, into region A (cvel / cdof_dot are written there after the sums)
#pragma unroll 4
  for (int i = lane; i < 10 * nb; i += 32) s_cin[i] = d.cinert[(size_t)w * nb * 10 + i];
  __syncwarp();
  for (int b = lane; b < nb; b += 32) {
    float acc[10];
#pragma unroll
    for (int k = 0; k < 10; k++) acc[k] = 0.f;
    const int e = m.body_subtree_end[b];
    for (int j = b; j < e; j++) {
      const float2* c = reinterpret_cast<const float2*>(s_cin + 10 * j);      // 40-byte stride from an 8-byte aligned base
#pragma unroll
      for (int k = 0; k < 5; k++) { const float2 v = c[k]; acc[2 * k] += v.x; acc[2 * k + 1] += v.y; }
    }
#pragma unroll
    for (int k = 0; k < 10; k++) s_crb[10 * b + k] = acc[k];
  }
  __syncwarp();      // region A changes hands: cinert -> cvel | cdof_dot
  // com_vel: cvel(b) = sum over ancestor dofs; cdof_dot of the body's own dofs on the way
  for (int b = lane; b < nb; b += 32) {
    sv6 v = zero6();
    const int c0 = m.body_chain_adr[b], c1 = m.body_chain_adr[b + 1], own = m.body_dofadr[b];
    const int nown = m.body_dofnum[b];
    int c = c0;
    for (; c < c1 - nown; c++) { const int j = m.body_chain[c]; axpy6(v, s_qvel[j], ld6s(s_cdof + 6 * j)); }
    // own joints, in order
    const int jn = m.body_jntnum[b], ja = m.body_jntadr[b];
    for (int jj = 0; jj < jn; jj++) {
      const int t = m.jnt_type[ja + jj];
      int da = m.jnt_dofadr[ja + jj];
      if (t == B2S_JNT_FREE) {
        for (int a = 0; a < 3; a++) { st6s(s_cdd + 6 * (da + a), zero6()); axpy6(v, s_qvel[da + a], ld6s(s_cdof + 6 * (da + a))); }
        da += 3;
      }
      if (t == B2S_JNT_FREE || t == B2S_JNT_BALL) {
        for (int a = 0; a < 3; a++) st6s(s_cdd + 6 * (da + a), cross_motion(v, ld6s(s_cdof + 6 * (da + a))));
        for (int a = 0; a < 3; a++) axpy6(v, s_qvel[da + a], ld6s(s_cdof + 6 * (da + a)));
      } else {
        st6s(s_cdd + 6 * da, cross_motion(v, ld6s(s_cdof + 6 * da)));
        axpy6(v, s_qvel[da], ld6s(s_cdof + 6 * da));
      }
    }
    (void)own;
    st6s(s_cvel + 6 * b, v);
  }
  __syncwarp();
  for (int i = lane; i < 6 * nb; i += 32) d.cvel[(size_t)w * 6 * nb + i] = s_cvel[i];
  for (int i = lane; i < 6 * nv; i += 32) d.cdof_dot[(size_t)w * 6 * nv + i] = s_cdd[i];
  // mass matrix: M[i][j] = cdof_j . (crb[body_i] cdof_i) for j ancestor-or-self of i
  const float* armature = m.dof_armature.at(w);
  for (int i = lane; i < nv; i += 32) {
    const int b = m.dof_bodyid[i];
    sv6 buf = mul_inert(s_crb + 10 * b, ld6s(s_cdof + 6 * i));
    const int c0 = m.body_chain_adr[b], c1 = m.body_chain_adr[b + 1];
    for (int c = c0; c < c1; c++) {
      const int j = m.body_chain[c];
      if (j > i) break;
      float v = dot6(ld6s(s_cdof + 6 * j), buf);
      if (j == i) v += armature[i];
      s_M[i * (i + 1) / 2 + j] = v;
    }
  }
  __syncwarp();      // region B changes hands: crb -> cfrc | fext
  // rne: bias force per body, then subtree sums
  {
    const bool grav = !(m.disableflags & B2S_DSBL_GRAVITY);
    for (int b = lane; b < nb; b += 32) {
      sv6 a = zero6();
      if (grav) { a.v[3] = -m.gravity[0]; a.v[4] = -m.gravity[1]; a.v[5] = -m.gravity[2]; }
      sv6 f = zero6(), fe = zero6();
      if (b > 0) {
        const int c0 = m.body_chain_adr[b], c1 = m.body_chain_adr[b + 1];
        for (int c = c0; c < c1; c++) { const int j = m.body_chain[c]; axpy6(a, s_qvel[j], ld6s(s_cdd + 6 * j)); }
        const float* I = d.cinert + ((size_t)w * nb + b) * 10;
        sv6 v = ld6s(s_cvel + 6 * b);
        f = mul_inert(I, a);
        sv6 g = cross_force(v, mul_inert(I, v));
        axpy6(f, 1.0f, g);
        // external Cartesian wrench applied at xipos, as a spatial force about the root subtree COM
        const float* x = d.xfrc_applied + ((size_t)w * nb + b) * 6;
        v3 frc = ld3(x), trq = ld3(x + 3);
        if (frc.x != 0.f || frc.y != 0.f || frc.z != 0.f || trq.x != 0.f || trq.y != 0.f || trq.z != 0.f) {
          v3 off = ld3(d.xipos + ((size_t)w * nb + b) * 3) - ld3(d.subtree_com + ((size_t)w * nb + m.body_rootid[b]) * 3);
          v3 t = trq + cross(off, frc);
          fe.v[0] = t.x; fe.v[1] = t.y; fe.v[2] = t.z; fe.v[3] = frc.x; fe.v[4] = frc.y; fe.v[5] = frc.z;
        }
      }
      st6s(s_cfrc + 6 * b, f); st6s(s_fext + 6 * b, fe);
    }
  }
  __syncwarp();
  // write M (symmetric, dense, row stride nvp) to HBM
  // one lane per row, 16-byte stores (nvp is a multiple of 4, rows are 16-byte aligned); row r mirrors the lower triangle
  {
    constexpr int Q = (7 * T + 3) / 4;
    const int n4 = nvp >> 2;
#pragma unroll 1
    for (int r = lane; r < nv; r += 32) {
      float4* out = reinterpret_cast<float4*>(d.qM + ((size_t)w * nv + r) * nvp);
#pragma unroll
      for (int j = 0; j < Q; j++) {
        float e[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
          const int c = 4 * j + k;
          e[k] = c < nv ? (c <= r ? s_M[r * (r + 1) / 2 + c] : s_M[c * (c + 1) / 2 + r]) : 0.f;
        }
        if (j < n4) out[j] = make_float4(e[0], e[1], e[2], e[3]);
      }
    }
  }
  // fixed tendons: length / velocity; spring (deadband) and damper forces accumulate in s_tmp
  if (m.ntendon > 0) {
    for (int i = lane; i < nv; i += 32) s_tmp[i] = 0.f;
    __syncwarp();
    const bool passive = !(m.disableflags & B2S_DSBL_PASSIVE);
    for (int t = lane; t < m.ntendon; t += 32) {
      const int a0 = m.tendon_adr[t], a1 = a0 + m.tendon_num[t];
      float len = 0.f, vel = 0.f;
      for (int k = a0; k < a1; k++) { len += m.wrap_coef[k] * qpos[m.wrap_qpos[k]]; vel += m.wrap_coef[k] * s_qvel[m.wrap_dof[k]]; }
      d.ten_length[(size_t)w * m.ntendon + t] = len;
      d.ten_velocity[(size_t)w * m.ntendon + t] = vel;
      const float ks = m.tendon_stiffness[t], kb = m.tendon_damping[t];
      if (passive && (ks != 0.f || kb != 0.f)) {
        const float lo = m.tendon_lengthspring[2 * t], hi = m.tendon_lengthspring[2 * t + 1];
        float f = 0.f;
        if (len > hi) f = ks * (hi - len); else if (len < lo) f = ks * (lo - len);
        f -= kb * vel;
        for (int k = a0; k < a1; k++) atomicAdd(&s_tmp[m.wrap_dof[k]], m.wrap_coef[k] * f);
      }
    }
    __syncwarp();
  }
  // actuation (joint or fixed-tendon transmission)
  if (!(m.disableflags & B2S_DSBL_ACTUATION)) {
    const float* gainprm = m.actuator_gainprm.at(w);
    const float* biasprm = m.actuator_biasprm.at(w);
    for (int u = lane; u < m.nu; u += 32) {
      const int j = m.actuator_trnid[u];
      const float gear = m.actuator_gear[u];
      const bool tend = m.actuator_trntype[u] == B2S_TRN_TENDON;
      float len, vel;
      if (tend) { len = gear * d.ten_length[(size_t)w * m.ntendon + j]; vel = gear * d.ten_velocity[(size_t)w * m.ntendon + j]; }
      else { len = gear * qpos[m.jnt_qposadr[j]]; vel = gear * s_qvel[m.jnt_dofadr[j]]; }
      float c = d.ctrl[(size_t)w * m.nu + u];
      if (m.actuator_ctrllimited[u] && !(m.disableflags & B2S_DSBL_CLAMPCTRL)) c = clampf(c, m.actuator_ctrlrange[2 * u], m.actuator_ctrlrange[2 * u + 1]);
      float gain = gainprm[3 * u];
      if (m.actuator_gaintype[u] == B2S_GAIN_AFFINE) gain += gainprm[3 * u + 1] * len + gainprm[3 * u + 2] * vel;
      float bias = 0.f;
      if (m.actuator_biastype[u] == B2S_BIAS_AFFINE) bias = biasprm[3 * u] + biasprm[3 * u + 1] * len + biasprm[3 * u + 2] * vel;
      float f = gain * c + bias;
      if (m.actuator_forcelimited[u]) f = clampf(f, m.actuator_forcerange[2 * u], m.actuator_forcerange[2 * u + 1]);
      d.actuator_length[(size_t)w * m.nu + u] = len;
      d.actuator_velocity[(size_t)w * m.nu + u] = vel;
      d.actuator_force[(size_t)w * m.nu + u] = f;
      if (tend) {
        const int a0 = m.tendon_adr[j], a1 = a0 + m.tendon_num[j];
        for (int k = a0; k < a1; k++) atomicAdd(&s_act[m.wrap_dof[k]], gear * m.wrap_coef[k] * f);
      } else atomicAdd(&s_act[m.jnt_dofadr[j]], gear * f);
    }
  } else {
    for (int u = lane; u < m.nu; u += 32) d.actuator_force[(size_t)w * m.nu + u] = 0.f;
  }
  __syncwarp();
  // per-dof generalized forces: bias = cdof . sum_subtree(cfrc), applied = cdof . sum_subtree(fext)
  {
    const float* damping = m.dof_damping.at(w);
    const float* stiff = m.jnt_stiffness.at(w);
    const float* qspring = m.qpos_spring.at(w);
    const bool passive = !(m.disableflags & B2S_DSBL_PASSIVE);
    for (int i = lane; i < nv; i += 32) {
      const int b = m.dof_bodyid[i], e = m.body_subtree_end[b];
      sv6 fs = zero6(), es = zero6();
      for (int j = b; j < e; j++) { axpy6(fs, 1.0f, ld6s(s_cfrc + 6 * j)); axpy6(es, 1.0f, ld6s(s_fext + 6 * j)); }
      sv6 cd = ld6s(s_cdof + 6 * i);
      const float bias = dot6(cd, fs), ext = dot6(cd, es);
      float pas = 0.f;
      if (passive) {
        const int j = m.dof_jntid[i], t = m.jnt_type[j];
        if ((t == B2S_JNT_HINGE || t == B2S_JNT_SLIDE) && stiff[j] != 0.f) pas -= stiff[j] * (qpos[m.jnt_qposadr[j]] - qspring[m.jnt_qposadr[j]]);
        pas -= damping[i] * s_qvel[i];
        if (m.ntendon > 0) pas += s_tmp[i];
      }
      const float smooth = pas - bias + d.qfrc_applied[(size_t)w * nv + i] + s_act[i] + ext;
      d.qfrc_bias[(size_t)w * nv + i] = bias;
      d.qfrc_passive[(size_t)w * nv + i] = pas;
      d.qfrc_actuator[(size_t)w * nv + i] = s_act[i];
      d.qfrc_smooth[(size_t)w * nv + i] = smooth;
      s_rhs[i] = smooth;
    }
  }
  __syncwarp();
  // factor M in shared memory and solve for the unconstrained acceleration
  rowchol_smem<T, true>(s_M, nv, ld, lane, s_invd, s_col);
  float b0 = lane < nv ? s_rhs[lane] : 0.f, b1 = lane + 32 < nv ? s_rhs[lane + 32] : 0.f;
  rowchol_solve<T, true>(s_M, s_invd, ld, lane, b0, b1);
  if (lane < nv) d.qacc_smooth[(size_t)w * nv + lane] = b0;
  if (lane + 32 < nv) d.qacc_smooth[(size_t)w * nv + lane + 32] = b1;
}

// ---------------------------------------------------------------------------------------------- launchers
void launch_position(const ModelDev& m, const DataDev& d, cudaStream_t st) {
  const int grid = (d.nworld + B2S_WARPS_PER_CTA - 1) / B2S_WARPS_PER_CTA;
  const size_t smem = smem_position(m);
  k_position<<<grid, 32 * B2S_WARPS_PER_CTA, smem, st>>>(m, d);
  b2s_count_launch();
}
#define TILE_DISPATCH(NV, EXPR_T)                               \
  switch (b2s_tile_for(NV)) {                                   \
    case 1: { constexpr int T = 1; EXPR_T; } break;             \
    case 2: { constexpr int T = 2; EXPR_T; } break;             \
    case 3: { constexpr int T = 3; EXPR_T; } break;             \
    case 4: { constexpr int T = 4; EXPR_T; } break;             \
    case 5: { constexpr int T = 5; EXPR_T; } break;             \
    case 6: { constexpr int T = 6; EXPR_T; } break;             \
    default: { constexpr int T = 8; EXPR_T; } break;            \
  }
void launch_dynamics(const ModelDev& m, const DataDev& d, cudaStream_t st) {
  const int grid = (d.nworld + B2S_WARPS_PER_CTA - 1) / B2S_WARPS_PER_CTA;
  const size_t smem = smem_dynamics(m);
  TILE_DISPATCH(m.nv, (k_dynamics<T><<<grid, 32 * B2S_WARPS_PER_CTA, smem, st>>>(m, d)));
  b2s_count_launch();
}
size_t smem_position(const ModelDev& m) { return (size_t)B2S_WARPS_PER_CTA * pos_smem_floats(m.nbody, m.njnt) * sizeof(float); }
size_t smem_dynamics(const ModelDev& m) {
  return (size_t)B2S_WARPS_PER_CTA * dyn_smem_floats(m.nv, m.nvp, m.nbody, b2s_tile_for(m.nv)) * sizeof(float);
}
const void* fn_position() { return (const void*)k_position; }
const void* fn_dynamics(const ModelDev& m) {
  const void* f = nullptr;
  TILE_DISPATCH(m.nv, (f = (const void*)k_dynamics<T>));
  return f;
}
