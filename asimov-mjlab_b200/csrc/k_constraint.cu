// k_constraint.cu -- joint-limit and contact constraint assembly, one warp per world.
//
// Row order follows MuJoCo: limits (joint order, lower side before upper) then contacts (contact order;
// pyramidal cone: 2(dim-1) rows per frictional contact, 1 row for condim 1).  Row addresses come from
// ballot / warp-scan compaction, so the layout is deterministic.
//
// MATRIX-FREE ROWS.  The reference backend materialises the dense Jacobian efc_J (njmax x nv floats per
// world, 42 KB for G1) and streams it through every solver iteration.  Here a row is stored in its
// geometric form instead:
//   limit row   : (dof, sign)                                  J[r][dof] = sign
//   contact row : (body1, body2, g[6])  with g = [off x dir ; dir] the unit wrench of the row direction
//                 about the world's reference point c_ref, so that
//                 J[r][i] = (i in chain(body2)) g.cdof_i - (i in chain(body1)) g.cdof_i
// (cdof are MuJoCo's com-based spatial motion axes; within one kinematic tree they already share the
// reference point subtree_com[root], other trees are re-referenced to the first tree's).  J v, J^T f and
// J^T D J then cost O(6) per row plus O(chain^2) per body pair (k_solver.cu), efc_vel is g.(cvel2 - cvel1),
// and 28 bytes per row replace 144.  The dense efc_J is produced on demand by k_efc_jacobian for callers
// that want to look at it (tests do).
#include "b2s_math.cuh"
#include "b2s_api.h"

extern __shared__ float b2s_smem[];

enum { EFC_FRICTION_DOF = 1, EFC_LIMIT = 3, EFC_CONTACT_FRICTIONLESS = 4, EFC_CONTACT_PYRAMIDAL = 5, EFC_CONTACT_ELLIPTIC = 6 };

DEV float impedance(const float* si, float pos, float margin) {
  const float d0 = si[0], dw = si[1], width = si[2], mid = si[3], power = si[4];
  if (d0 == dw || width <= B2S_MINVAL) return 0.5f * (d0 + dw);
  float x = fabsf((pos - margin) / width);
  if (x >= 1.f) return dw;
  if (x <= 0.f) return d0;
  float y;
  if (power == 1.f) y = x;
  else if (power == 2.f) y = x <= mid ? x * x / mid : 1.f - (1.f - x) * (1.f - x) / (1.f - mid);   // MuJoCo's default solimp: two powf calls were 43 % of this kernel's instructions
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

__global__ void __launch_bounds__(32 * B2S_WARPS_PER_CTA, 7) k_constraint(const ModelDev m, const DataDev d) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int w = blockIdx.x * B2S_WARPS_PER_CTA + warp;
  if (w >= d.nworld) return;
  const int nv = m.nv, nb = m.nbody, njmax = d.njmax;
  float* s_cvel = b2s_smem + (size_t)warp * 6 * nb;     // body velocities about the world's reference point
  if (m.disableflags & B2S_DSBL_CONSTRAINT) {
    if (lane == 0) { d.nefc[w] = 0; d.nl[w] = 0; d.nf[w] = 0; d.solve_order[atomicAdd(&d.solve_count[0], 1)] = w; }
    return;
  }
  const float* com = d.subtree_com + (size_t)w * nb * 3;
  const v3 cref = ld3(com + 3 * m.ref_root);
  for (int b = lane; b < nb; b += 32) {
    sv6 cv = ld6(d.cvel + ((size_t)w * nb + b) * 6);
    const int rb = m.body_rootid[b];
    if (rb != m.ref_root) {
      const v3 lin = mk3(cv.v[3], cv.v[4], cv.v[5]) + cross(mk3(cv.v[0], cv.v[1], cv.v[2]), cref - ld3(com + 3 * rb));
      cv.v[3] = lin.x; cv.v[4] = lin.y; cv.v[5] = lin.z;
    }
    st6(s_cvel + 6 * b, cv);
  }
  __syncwarp();
  const size_t row0 = (size_t)w * njmax;
  int nefc = 0;
  // ---- dof friction loss (mj_instantiateFriction): one row per dof with frictionloss > 0, J = e_dof, pos = margin = 0,
  // default solreffriction / solimpfriction; these rows come first
  if (m.fric && !(m.disableflags & B2S_DSBL_FRICTIONLOSS)) {
    const float* floss = m.dof_frictionloss.at(w);
    const float* qvel = d.qvel + (size_t)w * nv;
    const float si[5] = {0.9f, 0.95f, 0.001f, 0.5f, 2.0f};
    for (int base = 0; base < nv; base += 32) {
      const int i = base + lane;
      const bool act = i < nv && floss[i] > 0.f;
      const unsigned mask = __ballot_sync(FULL, act);
      const int r = nefc + __popc(mask & ((1u << lane) - 1u));
      if (act && r < njmax) {
        float* g = d.efc_g + (row0 + r) * 6;
        g[0] = 1.f; g[1] = g[2] = g[3] = g[4] = g[5] = 0.f;
        d.efc_bp[row0 + r] = -1 - i;
        d.efc_type[row0 + r] = EFC_FRICTION_DOF; d.efc_id[row0 + r] = i;
        d.efc_frictionloss[row0 + r] = floss[i];
        finish_row(m, d, row0 + r, 0.f, 0.f, qvel[i], 0.02f, 1.0f, si, m.dof_invweight0[i], 1.f);
      }
      nefc = min(nefc + __popc(mask), njmax);
    }
  }
  if (lane == 0) d.nf[w] = nefc;
  // ---- joint limits
  if (!(m.disableflags & B2S_DSBL_LIMIT)) {
    const float* range = m.jnt_range.at(w);
    const float* qpos = d.qpos + (size_t)w * m.nq;
    const float* qvel = d.qvel + (size_t)w * nv;
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
        float* g = d.efc_g + (row0 + r) * 6;
        g[0] = -side; g[1] = g[2] = g[3] = g[4] = g[5] = 0.f;
        d.efc_bp[row0 + r] = -1 - dof;
        d.efc_type[row0 + r] = EFC_LIMIT; d.efc_id[row0 + r] = j;
        finish_row(m, d, row0 + r, dist, margin, -side * qvel[dof], m.jnt_solref[2 * j], m.jnt_solref[2 * j + 1],
                   m.jnt_solimp + 5 * j, m.dof_invweight0[dof], 1.f);
      }
      nefc = min(nefc + __popc(mask), njmax);
    }
  }
  const int nl = nefc;
  // ---- contact rows: one lane per contact (address by warp scan, then the lane writes its own rows)
  const int ncon = (m.disableflags & B2S_DSBL_CONTACT) ? 0 : d.ncon[w];
  const size_t con0 = (size_t)w * d.nconmax;
  bool full = false;
  for (int base = 0; base < ncon; base += 32) {
    const int c = base + lane;
    int nrow = 0, dim = 0;
    if (c < ncon && d.con_dist[con0 + c] < d.con_includemargin[con0 + c]) {
      dim = d.con_dim[con0 + c];
      nrow = dim == 1 ? 1 : (m.cone == B2S_CONE_ELLIPTIC ? dim : 2 * (dim - 1));
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
    if (nrow > 0) {
      B2S_ASSERT(adr >= 0 && adr + nrow <= njmax && c < d.nconmax);
      const size_t o = con0 + c;
      const int g1 = d.con_geom[2 * o], g2 = d.con_geom[2 * o + 1];
      const int b1 = m.geom_bodyid[g1], b2 = m.geom_bodyid[g2];
      const float* fr = d.con_frame + 9 * o;
      const float* fri = d.con_friction + 5 * o;
      const v3 off = ld3(d.con_pos + 3 * o) - cref;
      // relative spatial velocity of the two bodies about c_ref
      float vr[6];
#pragma unroll
      for (int k = 0; k < 6; k++) vr[k] = s_cvel[6 * b2 + k] - s_cvel[6 * b1 + k];
      // base wrenches: translational directions (normal, tangent1, tangent2), then rotational ones
      float gb[6][6], vb[6];
#pragma unroll
      for (int k = 0; k < 6; k++) {
        if (k < dim) {
          const v3 dir = ld3(fr + 3 * (k % 3));
          if (k < 3) { const v3 t = cross(off, dir); gb[k][0] = t.x; gb[k][1] = t.y; gb[k][2] = t.z; gb[k][3] = dir.x; gb[k][4] = dir.y; gb[k][5] = dir.z; }
          else { gb[k][0] = dir.x; gb[k][1] = dir.y; gb[k][2] = dir.z; gb[k][3] = gb[k][4] = gb[k][5] = 0.f; }
          float s = 0.f;
#pragma unroll
          for (int q = 0; q < 6; q++) s += gb[k][q] * vr[q];
          vb[k] = s;
        }
      }
      const float tran = m.body_invweight0[2 * b1] + m.body_invweight0[2 * b2];
      const float mu0 = fri[0];
      const float diag = dim == 1 ? tran : tran + mu0 * mu0 * tran;
      const float scale = dim == 1 ? 1.f : 2.f * mu0 * mu0 / m.impratio;
      const float dist = d.con_dist[o], inc = d.con_includemargin[o];
      const float sr0 = d.con_solref[2 * o], sr1 = d.con_solref[2 * o + 1];
      const int bp = b1 | (b2 << 16);
      if (dim > 1 && m.cone == B2S_CONE_ELLIPTIC) {
        // elliptic cone: rows = the contact-frame axes; friction rows have pos = 0, share the normal row's impedance and are
        // regularised by R_1 = R_n / impratio, R_j = R_1 mu_1^2 / mu_j^2; their aref is -B vel (oracle orc_make_constraint)
        const float rot = m.body_invweight0[2 * b1 + 1] + m.body_invweight0[2 * b2 + 1];
        const float* si = d.con_solimp + 5 * o;
        float Rn = 0.f;
        for (int k = 0; k < nrow; k++) {
          const size_t row = row0 + adr + k;
          float* g = d.efc_g + row * 6;
          float vel = 0.f;
#pragma unroll
          for (int kk = 0; kk < 6; kk++) if (kk == k) {
#pragma unroll
            for (int q = 0; q < 6; q++) g[q] = gb[kk][q];
            vel = vb[kk];
          }
          d.efc_bp[row] = bp; d.efc_type[row] = EFC_CONTACT_ELLIPTIC; d.efc_id[row] = c;
          finish_row(m, d, row, dist, inc, vel, sr0, sr1, si, k < 3 ? tran : rot, 1.f);
          if (k == 0) Rn = 1.f / d.efc_D[row];
          else {
            const float mk = fri[k - 1];
            d.efc_D[row] = 1.f / fmaxf(Rn / m.impratio * mu0 * mu0 / (mk * mk), B2S_MINVAL);
            float B;
            if (sr0 > 0.f) {
              float tc = sr0;
              if (!(m.disableflags & B2S_DSBL_REFSAFE)) tc = fmaxf(tc, 2.f * m.timestep);
              B = 2.f / fmaxf(si[1] * tc, B2S_MINVAL);
            } else B = -sr1 / fmaxf(si[1], B2S_MINVAL);
            d.efc_aref[row] = -B * vel; d.efc_pos[row] = 0.f; d.efc_margin[row] = 0.f;
          }
        }
      } else
      for (int k = 0; k < nrow; k++) {
        const size_t row = row0 + adr + k;
        float* g = d.efc_g + row * 6;
        float vel;
        if (dim == 1) {
#pragma unroll
          for (int q = 0; q < 6; q++) g[q] = gb[0][q];
          vel = vb[0];
        } else {
          const int ti = 1 + (k >> 1);
          const float mu = ((k & 1) ? -1.f : 1.f) * fri[k >> 1];
          // gb is indexed dynamically: select with a short unrolled scan so it stays in registers
          float gt[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f}; float vt = 0.f;
#pragma unroll
          for (int kk = 1; kk < 6; kk++) if (kk == ti) {
#pragma unroll
            for (int q = 0; q < 6; q++) gt[q] = gb[kk][q];
            vt = vb[kk];
          }
#pragma unroll
          for (int q = 0; q < 6; q++) g[q] = gb[0][q] + mu * gt[q];
          vel = vb[0] + mu * vt;
        }
        d.efc_bp[row] = bp;
        d.efc_type[row] = dim == 1 ? EFC_CONTACT_FRICTIONLESS : EFC_CONTACT_PYRAMIDAL;
        d.efc_id[row] = c;
        finish_row(m, d, row, dist, inc, vel, sr0, sr1, d.con_solimp + 5 * o, diag, scale);
      }
    }
  }
  if (lane == 0) {
    d.nefc[w] = nefc; d.nl[w] = nl;
    // queue this world for the solver by expected cost: the Newton iteration count of its previous solve (temporal
    // coherence) x a per-iteration cost that grows with the row count; 0 rows = cheapest.  Order inside a class is whatever the atomics give; results do not depend on it.
    const int cls = nefc == 0 ? 0 : min(B2S_SOLVE_CLASSES - 1, 1 + (1 + min(d.solver_niter[w], 15)) * (32 + nefc) / 64);
    const int slot = atomicAdd(&d.solve_count[cls], 1);
    B2S_ASSERT(cls >= 0 && cls < B2S_SOLVE_CLASSES && slot >= 0 && slot < d.nworld && nefc <= njmax && nl <= nefc);
    d.solve_order[(size_t)cls * d.nworld + slot] = w;
  }
}

// Dense efc_J from the matrix-free rows (on demand; not part of step/forward).
__global__ void __launch_bounds__(32 * B2S_WARPS_PER_CTA) k_efc_jacobian(const ModelDev m, const DataDev d, float* J) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int w = blockIdx.x * B2S_WARPS_PER_CTA + warp;
  if (w >= d.nworld) return;
  const int nv = m.nv, nvp = m.nvp, nb = m.nbody, njmax = d.njmax;
  float* s_cdof = b2s_smem + (size_t)warp * 6 * nv;
  const float* com = d.subtree_com + (size_t)w * nb * 3;
  const v3 cref = ld3(com + 3 * m.ref_root);
  for (int i = lane; i < nv; i += 32) {
    sv6 cd = ld6(d.cdof + ((size_t)w * nv + i) * 6);
    const int rb = m.body_rootid[m.dof_bodyid[i]];
    if (rb != m.ref_root) {
      const v3 lin = mk3(cd.v[3], cd.v[4], cd.v[5]) + cross(mk3(cd.v[0], cd.v[1], cd.v[2]), cref - ld3(com + 3 * rb));
      cd.v[3] = lin.x; cd.v[4] = lin.y; cd.v[5] = lin.z;
    }
    st6(s_cdof + 6 * i, cd);
  }
  __syncwarp();
  const int nefc = d.nefc[w];
  const size_t row0 = (size_t)w * njmax;
  for (int r = 0; r < njmax; r++) {
    float* Jr = J + (row0 + r) * nvp;
    if (r >= nefc) { for (int i = lane; i < nvp; i += 32) Jr[i] = 0.f; continue; }
    const int bp = d.efc_bp[row0 + r];
    const sv6 g = ld6(d.efc_g + (row0 + r) * 6);
    for (int i = lane; i < nvp; i += 32) {
      float v = 0.f;
      if (i < nv) {
        if (bp < 0) v = (i == -1 - bp) ? g.v[0] : 0.f;
        else {
          const int b1 = bp & 0xffff, b2 = bp >> 16;
          const int s = (int)m.body_dofmask[(size_t)b2 * nv + i] - (int)m.body_dofmask[(size_t)b1 * nv + i];
          if (s) v = (float)s * dot6(g, ld6(s_cdof + 6 * i));
        }
      }
      Jr[i] = v;
    }
  }
}

void launch_constraint(const ModelDev& m, const DataDev& d, cudaStream_t st) {
  const int grid = (d.nworld + B2S_WARPS_PER_CTA - 1) / B2S_WARPS_PER_CTA;
  k_constraint<<<grid, 32 * B2S_WARPS_PER_CTA, smem_constraint(m), st>>>(m, d);
  b2s_count_launch();
}
void launch_efc_jacobian(const ModelDev& m, const DataDev& d, float* J, cudaStream_t st) {
  const int grid = (d.nworld + B2S_WARPS_PER_CTA - 1) / B2S_WARPS_PER_CTA;
  k_efc_jacobian<<<grid, 32 * B2S_WARPS_PER_CTA, (size_t)B2S_WARPS_PER_CTA * 6 * m.nv * sizeof(float), st>>>(m, d, J);
  b2s_count_launch();
}
size_t smem_constraint(const ModelDev& m) { return (size_t)B2S_WARPS_PER_CTA * 6 * m.nbody * sizeof(float); }
const void* fn_constraint() { return (const void*)k_constraint; }
