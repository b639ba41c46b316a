// k_solver.cu -- Newton constraint solver, one warp per world.
//
// Minimises  1/2 (a-a_s)^T M (a-a_s) + s(J a - aref)  over the acceleration a (MuJoCo's primal problem).
// Everything iteration-local lives on chip: the Hessian H = M + J^T D J is accumulated in REGISTER tiles
// (7 x 7 blocks of T x T, one lower-triangular block per lane, fed from 8-row blocks of efc_J staged in
// shared memory with coalesced 16-byte loads), factorised in shared memory (warp Cholesky) and applied with
// register-resident substitution.  Jaref / Jv stay in shared memory across iterations; efc_J is streamed
// from L2 twice per iteration (J v and the fused J^T f + Hessian pass).  Line search: mjwarp's parallel
// variant (one lane per candidate step, reference sim.py:91 ls_parallel=True) or MuJoCo's exact 1-D Newton
// with bracketing (warp-collective evaluations).
#include "b2s_math.cuh"
#include "b2s_api.h"

extern __shared__ float b2s_smem[];
#define RB 8   // rows of efc_J staged per block

__host__ __device__ inline int solve_smem_floats(int nv, int nvp, int njmax, int T) {
  // staging tile first (16-byte aligned), total rounded up to a multiple of 4 floats
  const int n = RB * nvp + 7 * T + 2 * RB + b2s_hdim(T) * b2s_hld(T) + 7 * T + 8 * nvp + 2 * njmax;
  return (n + 3) & ~3;
}

struct SolveCtx {
  int lane, nv, nvp, nefc, ld;
  const float *M, *J, *D, *aref;
  float *H, *invd, *qacc, *Ma, *grad, *search, *Mv, *qfs, *qas, *qfc, *Jaref, *Jv, *stage, *sDe, *sf;
};

// y[i] = sum_j M[i][j] x[j]; M symmetric in HBM/L2, read column-wise so lanes hit contiguous addresses
DEV void mat_vec_M(const SolveCtx& c, const float* x, float* y) {
  float a0 = 0.f, a1 = 0.f;
  for (int j = 0; j < c.nv; j++) {
    const float xj = x[j];
    const float* row = c.M + (size_t)j * c.nvp;
    if (c.lane < c.nv) a0 += row[c.lane] * xj;
    if (c.lane + 32 < c.nv) a1 += row[c.lane + 32] * xj;
  }
  if (c.lane < c.nv) y[c.lane] = a0;
  if (c.lane + 32 < c.nv) y[c.lane + 32] = a1;
}
// out[r] = J[r] . x - sub[r]   (lane per row, 16-byte loads)
DEV void mat_vec_J(const SolveCtx& c, const float* x, const float* sub, float* out) {
  for (int r = c.lane; r < c.nefc; r += 32) {
    const float4* row = reinterpret_cast<const float4*>(c.J + (size_t)r * c.nvp);
    float s = 0.f;
    for (int q = 0; q < c.nvp / 4; q++) {
      const float4 v = row[q];
      s += v.x * x[4 * q] + v.y * x[4 * q + 1] + v.z * x[4 * q + 2] + v.w * x[4 * q + 3];
    }
    out[r] = sub ? s - sub[r] : s;
  }
}
DEV float gauss_term(const SolveCtx& c) {
  float g = 0.f;
  for (int i = c.lane; i < c.nv; i += 32) g += (c.Ma[i] - c.qfs[i]) * (c.qacc[i] - c.qas[i]);
  return 0.5f * warp_sum(g);
}
DEV float constraint_cost(const SolveCtx& c, const float* jar) {
  float s = 0.f;
  for (int r = c.lane; r < c.nefc; r += 32) { const float x = jar[r]; if (x < 0.f) s += 0.5f * c.D[r] * x * x; }
  return warp_sum(s);
}
struct LsPoint { float alpha, cost, d0, d1; };
DEV LsPoint ls_eval(const SolveCtx& c, const float* qg, float alpha) {
  float q0 = 0.f, q1 = 0.f, q2 = 0.f;
  for (int r = c.lane; r < c.nefc; r += 32) {
    const float ja = c.Jaref[r], jv = c.Jv[r];
    if (ja + alpha * jv < 0.f) { const float D = c.D[r]; q0 += 0.5f * D * ja * ja; q1 += D * ja * jv; q2 += 0.5f * D * jv * jv; }
  }
  q0 = warp_sum(q0) + qg[0]; q1 = warp_sum(q1) + qg[1]; q2 = warp_sum(q2) + qg[2];
  LsPoint p;
  p.alpha = alpha; p.cost = alpha * alpha * q2 + alpha * q1 + q0; p.d0 = 2.f * alpha * q2 + q1; p.d1 = fmaxf(2.f * q2, B2S_MINVAL);
  return p;
}
DEV float ls_iterative(const SolveCtx& c, const ModelDev& m, const float* qg) {
  float sn = 0.f;
  for (int i = c.lane; i < c.nv; i += 32) sn += c.search[i] * c.search[i];
  sn = sqrtf(warp_sum(sn));
  if (sn < B2S_MINVAL) return 0.f;
  const float gtol = m.tolerance * m.ls_tolerance * sn * m.meaninertia * (float)max(c.nv, 1);
  const int maxit = m.ls_iterations;
  int it = 0;
  const LsPoint p0 = ls_eval(c, qg, 0.f);
  LsPoint p1 = ls_eval(c, qg, p0.alpha - p0.d0 / p0.d1); it++;
  if (p0.cost < p1.cost) p1 = p0;
  if (fabsf(p1.d0) < gtol) return p1.alpha;
  const float dir = p1.d0 < 0.f ? 1.f : -1.f;
  LsPoint p2 = p0; bool p2upd = false;
  while (p1.d0 * dir <= -gtol && it < maxit) {
    p2 = p1; p2upd = true;
    p1 = ls_eval(c, qg, p1.alpha - p1.d0 / p1.d1); it++;
    if (fabsf(p1.d0) < gtol) return p1.alpha;
  }
  if (it >= maxit || !p2upd) return p1.alpha;
  LsPoint p2next = p1, p1next = ls_eval(c, qg, p1.alpha - p1.d0 / p1.d1); it++;
  while (it < maxit) {
    const LsPoint pmid = ls_eval(c, qg, 0.5f * (p1.alpha + p2.alpha)); it++;
    bool b1 = false, b2 = false;
#pragma unroll
    for (int k = 0; k < 3; k++) {
      const LsPoint cd = k == 0 ? p1next : (k == 1 ? p2next : pmid);
      if (fabsf(cd.d0) < gtol) return cd.alpha;
      const bool in = (cd.alpha - p2.alpha) * dir > 0.f && (p1.alpha - cd.alpha) * dir > 0.f;
      if (cd.d0 * dir < 0.f && in && !b2) { p2 = cd; b2 = true; }
      else if (cd.d0 * dir > 0.f && in && !b1) { p1 = cd; b1 = true; }
    }
    if (!b1 && !b2) return pmid.cost < p0.cost ? pmid.alpha : 0.f;
    if (b1) { p1next = ls_eval(c, qg, p1.alpha - p1.d0 / p1.d1); it++; }
    if (b2) { p2next = ls_eval(c, qg, p2.alpha - p2.d0 / p2.d1); it++; }
  }
  if (p1.cost <= p2.cost && p1.cost < p0.cost) return p1.alpha;
  if (p2.cost <= p1.cost && p2.cost < p0.cost) return p2.alpha;
  return 0.f;
}
// one lane per candidate step: alpha_k = min_step^(k/(n-1)), k = 0..n-1, plus alpha = 0; rows are summed in
// row order by every lane (same order as the oracle), the best candidate wins (ties: smallest k)
DEV float ls_parallel(const SolveCtx& c, const ModelDev& m, const float* qg) {
  const int n = m.ls_iterations;
  // efc_D of the active rows is re-read by every candidate: stage it in the (idle) J staging tile
  const float* Dr = c.D;
  if (c.nefc <= RB * c.nvp) {
    for (int r = c.lane; r < c.nefc; r += 32) c.stage[r] = c.D[r];
    __syncwarp();
    Dr = c.stage;
  }
  const float lg = logf(m.ls_parallel_min_step);
  float best = 0.f, balpha = 0.f; int bk = 0x7fffffff;
  {  // alpha = 0 baseline, evaluated the same way
    float s = qg[0];
    for (int r = 0; r < c.nefc; r++) { const float x = c.Jaref[r]; if (x < 0.f) s += 0.5f * Dr[r] * x * x; }
    best = s; balpha = 0.f; bk = -1;
  }
  for (int base = 0; base < n; base += 32) {
    const int k = base + c.lane;
    if (k < n) {
      const float alpha = n > 1 ? expf((float)k / (float)(n - 1) * lg) : 1.f;
      float q0 = qg[0], q1 = qg[1], q2 = qg[2];
      for (int r = 0; r < c.nefc; r++) {
        const float ja = c.Jaref[r], jv = c.Jv[r];
        if (ja + alpha * jv < 0.f) { const float D = Dr[r]; q0 += 0.5f * D * ja * ja; q1 += D * ja * jv; q2 += 0.5f * D * jv * jv; }
      }
      const float cost = alpha * alpha * q2 + alpha * q1 + q0;
      if (cost < best) { best = cost; balpha = alpha; bk = k; }
    }
  }
  // warp arg-min on (cost, k); bk = -1 (alpha = 0) wins ties so a step is only taken if it strictly improves
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float oc = __shfl_xor_sync(FULL, best, o), oa = __shfl_xor_sync(FULL, balpha, o);
    const int ok = __shfl_xor_sync(FULL, bk, o);
    if (oc < best || (oc == best && ok < bk)) { best = oc; balpha = oa; bk = ok; }
  }
  __syncwarp();
  return balpha;
}

// Fused pass over efc_J: forces / active set from Jaref, qfrc_constraint = J^T f, and (if withH) the
// register-tile accumulation of J^T D J.
template <int T>
DEV void pass_JT(const SolveCtx& c, float* efc_force, bool withH, int bi, int bj, float (&acc)[T * T]) {
  float qf0 = 0.f, qf1 = 0.f;
  if (withH) {
#pragma unroll
    for (int k = 0; k < T * T; k++) acc[k] = 0.f;
  }
  const int nvp = c.nvp;
  for (int r0 = 0; r0 < c.nefc; r0 += RB) {
    const int nr = min(RB, c.nefc - r0);
    const float4* src = reinterpret_cast<const float4*>(c.J + (size_t)r0 * nvp);
    float4* dst = reinterpret_cast<float4*>(c.stage);
    for (int i = c.lane; i < nr * nvp / 4; i += 32) dst[i] = src[i];
    if (c.lane < RB) {
      const int r = r0 + c.lane;
      float De = 0.f, f = 0.f;
      if (r < c.nefc) {
        const float jar = c.Jaref[r];
        if (jar < 0.f) { De = c.D[r]; f = -De * jar; }
        efc_force[r] = f;
      }
      c.sDe[c.lane] = De; c.sf[c.lane] = f;
    }
    __syncwarp();
    for (int rr = 0; rr < nr; rr++) {
      const float De = c.sDe[rr];
      if (De == 0.f) continue;
      const float f = c.sf[rr];
      const float* row = c.stage + rr * nvp;
      if (c.lane < nvp) qf0 += row[c.lane] * f;
      if (c.lane + 32 < nvp) qf1 += row[c.lane + 32] * f;
      if (withH && bi >= 0) {
        float a[T], b[T];
#pragma unroll
        for (int k = 0; k < T; k++) { a[k] = row[bi * T + k] * De; b[k] = row[bj * T + k]; }
#pragma unroll
        for (int x = 0; x < T; x++)
#pragma unroll
          for (int y = 0; y < T; y++) acc[x * T + y] += a[x] * b[y];
      }
    }
    __syncwarp();
  }
  if (c.lane < c.nv) c.qfc[c.lane] = qf0;
  if (c.lane + 32 < c.nv) c.qfc[c.lane + 32] = qf1;
}

template <int T>
__global__ void __launch_bounds__(32 * B2S_WARPS_PER_CTA, 4) k_solve(const ModelDev m, const DataDev d) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int w = blockIdx.x * B2S_WARPS_PER_CTA + warp;
  if (w >= d.nworld) return;
  const int nv = m.nv, nvp = m.nvp, njmax = d.njmax;
  SolveCtx c;
  c.lane = lane; c.nv = nv; c.nvp = nvp; c.ld = b2s_hld(T); c.nefc = d.nefc[w];
  c.M = d.qM + (size_t)w * nv * nvp;
  c.J = d.efc_J + (size_t)w * njmax * nvp;
  c.D = d.efc_D + (size_t)w * njmax;
  c.aref = d.efc_aref + (size_t)w * njmax;
  float* base = b2s_smem + (size_t)warp * solve_smem_floats(nv, nvp, njmax, T);
  c.stage = base; base += RB * nvp;          // read slightly past its end by the last Hessian block
  c.sDe = base + 7 * T; c.sf = c.sDe + RB; base += 7 * T + 2 * RB;
  c.H = base; base += b2s_hdim(T) * c.ld;
  c.invd = base; base += 7 * T;
  c.qacc = base; c.Ma = base + nvp; c.grad = base + 2 * nvp; c.search = base + 3 * nvp; c.Mv = base + 4 * nvp;
  c.qfs = base + 5 * nvp; c.qas = base + 6 * nvp; c.qfc = base + 7 * nvp; base += 8 * nvp;
  c.Jaref = base; c.Jv = base + njmax; base += 2 * njmax;
  float* efc_force = d.efc_force + (size_t)w * njmax;
  float* g_qacc = d.qacc + (size_t)w * nv;
  float* g_qfc = d.qfrc_constraint + (size_t)w * nv;
  for (int i = lane; i < nvp; i += 32) {
    c.qas[i] = i < nv ? d.qacc_smooth[(size_t)w * nv + i] : 0.f;
    c.qfs[i] = i < nv ? d.qfrc_smooth[(size_t)w * nv + i] : 0.f;
    c.search[i] = 0.f; c.qacc[i] = 0.f; c.qfc[i] = 0.f;
  }
  // the RB*nvp staging tile is read slightly past nvp by the last Hessian block: keep the tail finite
  for (int i = lane; i < RB * nvp + 7 * T; i += 32) c.stage[i] = 0.f;   // tail [RB*nvp, +7T) stays zero
  __syncwarp();
  if (c.nefc == 0) {
    for (int i = lane; i < nv; i += 32) { g_qacc[i] = c.qas[i]; g_qfc[i] = 0.f; }
    if (lane == 0) { d.solver_niter[w] = 0; d.solver_cost[w] = 0.f; }
    return;
  }
  // lower-triangular Hessian block owned by this lane
  int bi, bj;
  tile_owner(lane, bi, bj);
  // ---- warm start (mj_fwdConstraint): keep qacc_warmstart only if it beats qacc_smooth
  const bool warm = !(m.disableflags & B2S_DSBL_WARMSTART);
  for (int i = lane; i < nv; i += 32) c.qacc[i] = warm ? d.qacc_warmstart[(size_t)w * nv + i] : c.qas[i];
  __syncwarp();
  mat_vec_M(c, c.qacc, c.Ma);
  mat_vec_J(c, c.qacc, c.aref, c.Jaref);
  __syncwarp();
  float gauss = gauss_term(c);
  float cost = constraint_cost(c, c.Jaref) + gauss;
  if (warm) {
    mat_vec_J(c, c.qas, c.aref, c.Jv);   // Jv as scratch
    __syncwarp();
    const float cost_smooth = constraint_cost(c, c.Jv);
    if (cost > cost_smooth) {
      for (int i = lane; i < nv; i += 32) c.qacc[i] = c.qas[i];
      for (int r = lane; r < c.nefc; r += 32) c.Jaref[r] = c.Jv[r];
      __syncwarp();
      mat_vec_M(c, c.qacc, c.Ma);
      __syncwarp();
      gauss = gauss_term(c);
      cost = cost_smooth + gauss;
    }
  }
  const float scale = 1.0f / (m.meaninertia * (float)max(nv, 1));
  int iter = 0;
  float acc[T * T];
  while (true) {
    __syncwarp();
    pass_JT<T>(c, efc_force, true, bi, bj, acc);
    __syncwarp();
    float gn = 0.f;
    for (int i = lane; i < nv; i += 32) { const float g = c.Ma[i] - c.qfs[i] - c.qfc[i]; c.grad[i] = g; gn += g * g; }
    gn = sqrtf(warp_sum(gn));
    if (iter > 0 && scale * gn < m.tolerance) break;
    if (iter >= m.iterations) break;
    // H = M + J^T D J: add M to the register tiles (identity on the padding), blocked Cholesky in registers
    // (tiles exchanged through c.H, which ends up holding L), search = -H^-1 grad
    if (bi >= 0) {
#pragma unroll
      for (int x = 0; x < T; x++)
#pragma unroll
        for (int y = 0; y < T; y++) {
          const int i = bi * T + x, j = bj * T + y;
          if (i < nv && j < nv) acc[x * T + y] += c.M[(size_t)i * nvp + j];
          else acc[x * T + y] = (i == j) ? 1.0f : 0.0f;
        }
    }
    tile_cholesky_regs<T>(acc, bi, bj, c.H, c.ld, c.invd);
    float b0 = lane < nv ? c.grad[lane] : 0.f, b1 = lane + 32 < nv ? c.grad[lane + 32] : 0.f;
    warp_chol_solve_inv(c.H, c.invd, nv, c.ld, lane, b0, b1);
    if (lane < nv) c.search[lane] = -b0;
    if (lane + 32 < nv) c.search[lane + 32] = -b1;
    __syncwarp();
    // line-search set-up: Mv, Jv, quadratic Gauss coefficients
    mat_vec_M(c, c.search, c.Mv);
    mat_vec_J(c, c.search, nullptr, c.Jv);
    __syncwarp();
    float g1 = 0.f, g2 = 0.f;
    for (int i = lane; i < nv; i += 32) { g1 += c.search[i] * (c.Ma[i] - c.qfs[i]); g2 += c.search[i] * c.Mv[i]; }
    float qg[3] = {gauss, warp_sum(g1), 0.5f * warp_sum(g2)};
    const float alpha = m.ls_parallel ? ls_parallel(c, m, qg) : ls_iterative(c, m, qg);
    if (alpha == 0.f) break;
    for (int i = lane; i < nv; i += 32) { c.qacc[i] += alpha * c.search[i]; c.Ma[i] += alpha * c.Mv[i]; }
    for (int r = lane; r < c.nefc; r += 32) c.Jaref[r] += alpha * c.Jv[r];
    __syncwarp();
    const float old = cost;
    gauss = gauss_term(c);
    cost = constraint_cost(c, c.Jaref) + gauss;
    iter++;
    if (scale * (old - cost) < m.tolerance) {
      __syncwarp();
      pass_JT<T>(c, efc_force, false, bi, bj, acc);
      __syncwarp();
      break;
    }
  }
  for (int i = lane; i < nv; i += 32) { g_qacc[i] = c.qacc[i]; g_qfc[i] = c.qfc[i]; }
  if (lane == 0) { d.solver_niter[w] = iter; d.solver_cost[w] = cost; }
}

static int tile_for(int nv) { return b2s_tile_for(nv); }

#define SOLVE_DISPATCH(EXPR_T)                                  \
  switch (tile_for(m.nv)) {                                     \
    case 1: { constexpr int T = 1; EXPR_T; } break;             \
    case 2: { constexpr int T = 2; EXPR_T; } break;             \
    case 3: { constexpr int T = 3; EXPR_T; } break;             \
    case 4: { constexpr int T = 4; EXPR_T; } break;             \
    case 5: { constexpr int T = 5; EXPR_T; } break;             \
    case 6: { constexpr int T = 6; EXPR_T; } break;             \
    default: { constexpr int T = 8; EXPR_T; } break;            \
  }

void launch_solve(const ModelDev& m, const DataDev& d, cudaStream_t st) {
  const int grid = (d.nworld + B2S_WARPS_PER_CTA - 1) / B2S_WARPS_PER_CTA;
  const size_t smem = smem_solve(m, d);
  SOLVE_DISPATCH((k_solve<T><<<grid, 32 * B2S_WARPS_PER_CTA, smem, st>>>(m, d)));
  b2s_count_launch();
}
size_t smem_solve(const ModelDev& m, const DataDev& d) {
  return (size_t)B2S_WARPS_PER_CTA * solve_smem_floats(m.nv, m.nvp, d.njmax, tile_for(m.nv)) * sizeof(float);
}
const void* fn_solve(const ModelDev& m) {
  const void* f = nullptr;
  SOLVE_DISPATCH((f = (const void*)k_solve<T>));
  return f;
}
