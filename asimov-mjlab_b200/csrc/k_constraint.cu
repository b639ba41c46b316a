// k_constraint.cu -- joint-limit and contact constraint assembly, one warp per world.
//
// Row order follows MuJoCo: limits (joint order, lower side before upper) then contacts (contact order;
// pyramidal cone: 2(dim-1) rows per frictional contact, 1 row for condim 1).  Row addresses come from
// ballot / warp-scan compaction, so the layout is deterministic.  Each contact's rows are produced
// cooperatively: lanes own Jacobian columns (dofs), so efc_J rows are written with coalesced stores and
// efc_vel = J qvel is a warp reduction.
#include "b2s_math.cuh"
#include "b2s_api.h"

extern __shared__ float b2s_smem[];

enum { EFC_LIMIT = 3, EFC_CONTACT_FRICTIONLESS = 4, EFC_CONTACT_PYRAMIDAL = 5 };

DEV float impedance(const float* si, float pos, float margin) {
  const float d0 = si[0], dw = si[1], width = si[2], mid = si[3], power = si[4];
  if (d0 == dw || width <= B2S_MINVAL) return 0.5f * (d0 + dw);
  float x = fabsf((pos - margin) / width);
  if (x >= 1.f) return dw;
  if (x <= 0.f) return d0;
  float y;
  if (power == 1.f) y = x;
  else if (x <= mid) y = powf(x, power) / powf(mid, power - 1.f);
  else y = 1.f - powf(1.f - x, power) / powf(1.f - mid, power - 1.f);
  return d0 + y * (dw - d0);
}
// D and aref of one row (mj_makeImpedance): R = (1-imp)/imp * diagApprox * Rscale
DEV void finish_row(const ModelDev& m, const DataDev& d, size_t row, float pos, float margin, float vel,
                    float sr0, float sr1, const float* si, float diagApprox, float Rscale) {
  const float imp = impedance(si, pos, margin);
  const float R = fmaxf((1.f - imp) * diagApprox / imp * Rscale, B2S_MINVAL);
  float K, B;
  const float dmax = si[1];
  if (sr0 > 0.f) {
    float tc = sr0;
    if (!(m.disableflags & B2S_DSBL_REFSAFE)) tc = fmaxf(tc, 2.f * m.timestep);
    K = 1.f / fmaxf(dmax * dmax * tc * tc * sr1 * sr1, B2S_MINVAL);
    B = 2.f / fmaxf(dmax * tc, B2S_MINVAL);
  } else {
    K = -sr0 / fmaxf(dmax * dmax, B2S_MINVAL);
    B = -sr1 / fmaxf(dmax, B2S_MINVAL);
  }
  d.efc_D[row] = 1.f / R;
  d.efc_pos[row] = pos; d.efc_margin[row] = margin; d.efc_vel[row] = vel;
  d.efc_aref[row] = -B * vel - K * imp * (pos - margin);
  d.efc_force[row] = 0.f;
}

__global__ void __launch_bounds__(32 * B2S_WARPS_PER_CTA) k_constraint(const ModelDev m, const DataDev d) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int w = blockIdx.x * B2S_WARPS_PER_CTA + warp;
  if (w >= d.nworld) return;
  const int nv = m.nv, nvp = m.nvp, nb = m.nbody, njmax = d.njmax;
  float* s_cdof = b2s_smem + (size_t)warp * (6 * nv + nvp);
  float* s_qvel = s_cdof + 6 * nv;
  if (m.disableflags & B2S_DSBL_CONSTRAINT) { if (lane == 0) { d.nefc[w] = 0; d.nl[w] = 0; } return; }
  for (int i = lane; i < 6 * nv; i += 32) s_cdof[i] = d.cdof[(size_t)w * 6 * nv + i];
  for (int i = lane; i < nvp; i += 32) s_qvel[i] = i < nv ? d.qvel[(size_t)w * nv + i] : 0.f;
  __syncwarp();
  const size_t row0 = (size_t)w * njmax;
  float* J = d.efc_J + row0 * nvp;
  int nefc = 0;
  // ---- joint limits
  if (!(m.disableflags & B2S_DSBL_LIMIT)) {
    const float* range = m.jnt_range.at(w);
    const float* qpos = d.qpos + (size_t)w * m.nq;
    for (int base = 0; base < 2 * m.njnt; base += 32) {
      const int item = base + lane, j = item >> 1;
      const float side = (item & 1) ? 1.f : -1.f;
      bool act = false; float dist = 0.f, margin = 0.f;
      if (j < m.njnt && m.jnt_limited[j]) {
        const int t = m.jnt_type[j];
        if (t == B2S_JNT_HINGE || t == B2S_JNT_SLIDE) {
          margin = m.jnt_margin[j];
          dist = side * (range[2 * j + (item & 1)] - qpos[m.jnt_qposadr[j]]);
          act = dist < margin;
        }
      }
      const unsigned mask = __ballot_sync(FULL, act);
      const int r = nefc + __popc(mask & ((1u << lane) - 1u));
      if (act && r < njmax) {
        const int dof = m.jnt_dofadr[j];
        for (int i = 0; i < nvp; i++) J[(size_t)r * nvp + i] = 0.f;
        J[(size_t)r * nvp + dof] = -side;
        d.efc_type[row0 + r] = EFC_LIMIT; d.efc_id[row0 + r] = j;
        finish_row(m, d, row0 + r, dist, margin, -side * s_qvel[dof], m.jnt_solref[2 * j], m.jnt_solref[2 * j + 1],
                   m.jnt_solimp + 5 * j, m.dof_invweight0[dof], 1.f);
      }
      nefc = min(nefc + __popc(mask), njmax);
    }
  }
  const int nl = nefc;
  // ---- contact row addresses
  const int ncon = (m.disableflags & B2S_DSBL_CONTACT) ? 0 : d.ncon[w];
  const size_t con0 = (size_t)w * d.nconmax;
  bool full = false;
  for (int base = 0; base < ncon; base += 32) {
    const int c = base + lane;
    int nrow = 0;
    if (c < ncon && d.con_dist[con0 + c] < d.con_includemargin[con0 + c]) {
      const int dim = d.con_dim[con0 + c];
      nrow = dim == 1 ? 1 : 2 * (dim - 1);
    }
    int total;
    const int excl = warp_excl_scan(nrow, lane, &total);
    int adr = nefc + excl;
    const bool over = full || (adr + nrow > njmax);   // once the row budget overflows, drop this and all later contacts
    const unsigned omask = __ballot_sync(FULL, over && nrow > 0);
    if (omask) {
      const int first = __ffs(omask) - 1;
      if (lane >= first) { adr = -1; nrow = 0; }
      full = true;
      const int used = __shfl_sync(FULL, nefc + excl, first);
      nefc = used;
    } else nefc += total;
    if (c < ncon) d.con_efc_address[con0 + c] = nrow > 0 ? adr : -1;
  }
  __syncwarp();
  // ---- contact rows, one contact at a time, lanes own dof columns
  const float* com = d.subtree_com + (size_t)w * nb * 3;
  for (int c = 0; c < ncon; c++) {
    const int adr = d.con_efc_address[con0 + c];
    if (adr < 0) continue;
    const int dim = d.con_dim[con0 + c], nrow = dim == 1 ? 1 : 2 * (dim - 1);
    const int g1 = d.con_geom[2 * (con0 + c)], g2 = d.con_geom[2 * (con0 + c) + 1];
    const int b1 = m.geom_bodyid[g1], b2 = m.geom_bodyid[g2];
    const float* fr = d.con_frame + 9 * (con0 + c);
    const float* fri = d.con_friction + 5 * (con0 + c);
    const v3 pos = ld3(d.con_pos + 3 * (con0 + c));
    const v3 off1 = pos - ld3(com + 3 * m.body_rootid[b1]), off2 = pos - ld3(com + 3 * m.body_rootid[b2]);
    const v3 fn = ld3(fr), ft1 = ld3(fr + 3), ft2 = ld3(fr + 6);
    const unsigned char* k1 = m.body_dofmask + (size_t)b1 * nv;
    const unsigned char* k2 = m.body_dofmask + (size_t)b2 * nv;
    float vel[10];
#pragma unroll
    for (int k = 0; k < 10; k++) vel[k] = 0.f;
    for (int i = lane; i < nvp; i += 32) {
      float jc[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      if (i < nv) {
        const v3 ang = ld3(s_cdof + 6 * i), lin = ld3(s_cdof + 6 * i + 3);
        v3 jp = mk3(0.f, 0.f, 0.f), jr = mk3(0.f, 0.f, 0.f);
        if (k2[i]) { jp = jp + lin + cross(ang, off2); jr = jr + ang; }
        if (k1[i]) { jp = jp - (lin + cross(ang, off1)); jr = jr - ang; }
        jc[0] = dot(fn, jp); jc[1] = dot(ft1, jp); jc[2] = dot(ft2, jp);
        jc[3] = dot(fn, jr); jc[4] = dot(ft1, jr); jc[5] = dot(ft2, jr);
      }
      const float qv = s_qvel[i];
      if (dim == 1) { J[(size_t)adr * nvp + i] = jc[0]; vel[0] += jc[0] * qv; }
      else {
        for (int k = 0; k < nrow; k++) {
          const float v = jc[0] + ((k & 1) ? -1.f : 1.f) * fri[k >> 1] * jc[1 + (k >> 1)];
          J[(size_t)(adr + k) * nvp + i] = v;
          vel[k] += v * qv;
        }
      }
    }
    float myvel = 0.f;
    for (int k = 0; k < nrow; k++) { const float s = warp_sum(vel[k]); if (lane == k) myvel = s; }
    if (lane < nrow) {
      const float tran = m.body_invweight0[2 * b1] + m.body_invweight0[2 * b2];
      const float mu0 = fri[0];
      const float diag = dim == 1 ? tran : tran + mu0 * mu0 * tran;
      const float scale = dim == 1 ? 1.f : 2.f * mu0 * mu0 / m.impratio;
      d.efc_type[row0 + adr + lane] = dim == 1 ? EFC_CONTACT_FRICTIONLESS : EFC_CONTACT_PYRAMIDAL;
      d.efc_id[row0 + adr + lane] = c;
      finish_row(m, d, row0 + adr + lane, d.con_dist[con0 + c], d.con_includemargin[con0 + c], myvel,
                 d.con_solref[2 * (con0 + c)], d.con_solref[2 * (con0 + c) + 1], d.con_solimp + 5 * (con0 + c), diag, scale);
    }
  }
  if (lane == 0) { d.nefc[w] = nefc; d.nl[w] = nl; }
}

void launch_constraint(const ModelDev& m, const DataDev& d, cudaStream_t st) {
  const int grid = (d.nworld + B2S_WARPS_PER_CTA - 1) / B2S_WARPS_PER_CTA;
  k_constraint<<<grid, 32 * B2S_WARPS_PER_CTA, smem_constraint(m), st>>>(m, d);
  b2s_count_launch();
}
size_t smem_constraint(const ModelDev& m) { return (size_t)B2S_WARPS_PER_CTA * (6 * m.nv + m.nvp) * sizeof(float); }
const void* fn_constraint() { return (const void*)k_constraint; }
