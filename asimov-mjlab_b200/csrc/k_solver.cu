// k_solver.cu -- matrix-free Newton constraint solver, one warp per world.
//
// Minimises  1/2 (a-a_s)^T M (a-a_s) + s(J a - aref)  over the acceleration a (MuJoCo's primal problem).
// The Jacobian is never materialised.  A contact row is (body1, body2, g) with g the 6-D unit wrench of the
// row direction about the world's reference point (k_constraint.cu), a limit row is (dof, sign), so
//   J x      = g . (X[body2] - X[body1]),  X[b] = sum_{i in chain(b)} cdof_i x_i      (one lane per body / row)
//   J^T f    = per run of rows that share a body pair: W = sum_r f_r g_r, then  s_i cdof_i . W  on the
//              signed chain difference of the pair (<= 2 x tree depth dofs)
//   J^T D J  = per run: K = sum_r D_r g_r g_r^T (6x6, one lane per entry), then  (s_i cdof_i)^T K (s_j cdof_j)
// which is O(6) per row plus O(chain^2) per body pair instead of O(nv^2) per row, with everything in shared
// memory (cdof 6nv floats, H nv^2, per-row Jaref/Jv/D).  H = M + J^T D J is factorised by the blocked
// register-tile Cholesky (b2s_math.cuh) and applied by shuffle substitution.  Line search: mjwarp's parallel
// variant (one lane per candidate step, reference sim.py:91 ls_parallel=True) or MuJoCo's exact 1-D Newton
// with bracketing (warp-collective evaluations).
#include "b2s_math.cuh"
#include "b2s_api.h"

extern __shared__ __align__(16) float b2s_smem[];

// The solver has its own CTA shape: ONE warp (= one world) per CTA.  Newton iteration counts differ from world to
// world, and a multi-warp CTA keeps its shared memory and registers until its slowest warp is done; measured on
// B200 at 4096 G1 worlds: 4 warps x 4 CTAs/SM 0.480 ms, 3 x 5 0.473, 2 x 7 0.467, 1 x 14..18 0.449-0.456 (occupancy
// between 12 and 18 warps per SM makes no difference; 96 registers with 18 warps of 3 or 6 per CTA is slower, 0.54).
#ifndef SOLVE_WARPS
#define SOLVE_WARPS 1
#endif
#ifndef SOLVE_MINB
#define SOLVE_MINB 16
#endif
// small per-dof / per-row loops stay rolled: the kernel's loop body is several times the 32 KB instruction cache
#ifdef SOLVE_UNROLL_SMALL
#define ROLL
#else
#define ROLL _Pragma("unroll 1")
#endif

__host__ __device__ inline int solve_scratch_floats(int nb) { return 6 * (nb > 32 ? nb : 32) + 32 + 32; }
// Staged per-world / per-model data (north_star: "TMA or shared-memory staging"): the world's mass matrix (nv x nvp floats,
// one cp.async.bulk completed on an mbarrier), then the barrier itself; at the end of the allocation the model's ancestor-dof
// chains as bytes + their CSR offsets as 16-bit words (read by every J v product and run flush of every iteration).
__host__ __device__ inline int solve_stage_front_floats(int T) { const int N = 7 * T; return N * ((N + 3) & ~3) + 4; }
__host__ __device__ inline int solve_chain_floats(int nb, int nchain) { return ((nchain + 3) >> 2) + ((nb + 2) >> 1); }
__host__ __device__ inline int solve_soff(int njmax, int nconmax, int T) {      // offset of the scratch area
  const int N = 7 * T, LD = (7 * T + 1) | 1, HF = (N * LD + N + 3) & ~3, NVP = (N + 3) & ~3;
  return (HF + 8 * NVP + 6 * N + 3 * njmax + nconmax + 2 + (njmax + 31) / 32 + 3) & ~3;
}
// nopt: optional vectors at the very end (CG: gradold, Mgradold; friction-loss rows: fl, frf) -- 4 in every variant but the default one
__host__ __device__ inline int solve_opt_off(int nb, int njmax, int nconmax, int T, int nchain, int elliptic) {
  return (solve_soff(njmax, nconmax, T) + solve_scratch_floats(nb) + (elliptic ? 2 * njmax : 0) + solve_chain_floats(nb, nchain) + 3) & ~3;
}
__host__ __device__ inline int solve_smem_floats(int nv, int nvp, int nb, int njmax, int nconmax, int T, int nchain, int elliptic = 0, int nopt = 4) {
  const int NVP = (7 * T + 3) & ~3;
  return (solve_stage_front_floats(T) + solve_opt_off(nb, njmax, nconmax, T, nchain, elliptic) + nopt * NVP + 3) & ~3;
}
// Row capacity of the shared-memory layout.  The default variant (V = 0) is launched with room for ModelDev::solve_rowcap (128) rows instead of
// njmax: with 8 vectors and 128 rows a G1 world needs 15.2 KB instead of 18.2 KB, which is 14 CTAs per SM instead of 12 -- 4096
// worlds are 27.7 per SM, i.e. two full rounds instead of two and a third.  Worlds with more rows are solved by a second launch
// with the full layout (k_constraint queues them separately).
#define SOLVE_BIG_SPAN 1       // worlds one CTA of the second launch looks at (16 was tried: a model whose worlds often exceed the capacity -- Asimov-toe -- then solves them 16 deep)

// Per-warp shared-memory layout with compile-time offsets (one base register instead of ~18 pointers):
//   H hdim*ld | invd 7T (padded to a float4 multiple) | 8 vectors of NVP | cdof 6*hdim | Jaref, Jv, D rowcap each |
//   run list nconmax+2 | active-set bit mask njmax/32 | scratch X/Y 6*max(nb,32), KW 32, Li 32
template <int T>
struct SolveCtx {
  static constexpr int N = 7 * T, LD = (7 * T + 1) | 1, HF = (N * LD + N + 3) & ~3, NVP = (N + 3) & ~3, FIX = HF + 8 * NVP + 6 * N;
  int lane, nv, nvp, nb, nefc, nl, nf, nrun, njmax, nconmax, soff;
  static constexpr int ld = LD, hfloats = HF;
  const float *M, *G, *aref;      // M: the staged copy in shared memory (row stride nvp)
  const int* bp;
  const unsigned char* chain;     // staged body_chain (dof ids) and body_chain_adr
  const unsigned short* cadr;
  float* sm;
  float* opt;                     // optional vectors (variants with CG / friction-loss rows only)
  DEV float* H() const { return sm; }
  DEV float* invd() const { return sm + N * LD; }
  DEV float* qacc() const { return sm + HF; }
  DEV float* Ma() const { return sm + HF + NVP; }
  DEV float* grad() const { return sm + HF + 2 * NVP; }
  DEV float* search() const { return sm + HF + 3 * NVP; }
  DEV float* Mv() const { return sm + HF + 4 * NVP; }
  DEV float* qfs() const { return sm + HF + 5 * NVP; }
  DEV float* qas() const { return sm + HF + 6 * NVP; }
  DEV float* qfc() const { return sm + HF + 7 * NVP; }
  DEV float* gradold() const { return opt; }                   // CG only
  DEV float* Mgradold() const { return opt + NVP; }            // CG only
  DEV float* fl() const { return opt + 2 * NVP; }              // friction-loss rows: f of row r < nf
  DEV float* frf() const { return opt + 3 * NVP; }             // and R f = f / D, the half width of the quadratic zone
  DEV float* cdof() const { return sm + HF + 8 * NVP; }
  DEV float* Jaref() const { return sm + FIX; }
  DEV float* Jv() const { return sm + FIX + njmax; }
  DEV float* D() const { return sm + FIX + 2 * njmax; }
  DEV int* run() const { return (int*)(sm + FIX + 3 * njmax); }
  DEV unsigned* amask() const { return (unsigned*)(sm + FIX + 3 * njmax) + nconmax + 2; }
  DEV float* X() const { return sm + soff; }
  DEV float* KW() const { return sm + soff + 6 * max(nb, 32); }
  DEV int* Li() const { return (int*)(sm + soff + 6 * max(nb, 32) + 32); }
  // elliptic cones only (the arrays exist only in that variant's allocation): per row, 0 = not part of an elliptic contact,
  // 1 = first (normal) row of one, 2 = one of its friction rows; and the row's scale (regularised mu for the normal row,
  // the friction coefficient for the others)
  DEV int* ekind() const { return (int*)(sm + soff + 6 * max(nb, 32) + 64); }
  DEV float* efri() const { return sm + soff + 6 * max(nb, 32) + 64 + njmax; }
};

// y[i] = sum_j M[i][j] x[j].  M is dense and symmetric, staged in shared memory by the kernel prologue's bulk copy, row stride
// nvp (a multiple of 4, padding columns are zero): lane = row, the row is read as float4, x as float4 broadcasts (zero-padded
// to NVP).  One accumulator, columns in ascending order.
template <int T>
DEV void mat_vec_M(const SolveCtx<T>& c, const float* x, float* y) {
  constexpr int N = 7 * T, Q = SolveCtx<T>::NVP / 4;
  const int n4 = c.nvp >> 2;
  const float4* x4 = reinterpret_cast<const float4*>(x);
#pragma unroll 1
  for (int half = 0; half < (N > 32 ? 2 : 1); half++) {
    const int row = c.lane + 32 * half;
    if (half == 1 && __all_sync(FULL, row >= c.nv)) break;
    const float4* r = reinterpret_cast<const float4*>(c.M + (size_t)min(row, c.nv - 1) * c.nvp);
    float4 mrow[Q];
#pragma unroll
    for (int j = 0; j < Q; j++) mrow[j] = r[min(j, n4 - 1)];     // j >= n4 meets x == 0
    float a = 0.f;
#pragma unroll
    for (int j = 0; j < Q; j++) {
      const float4 xv = x4[j];
      a += mrow[j].x * xv.x; a += mrow[j].y * xv.y; a += mrow[j].z * xv.z; a += mrow[j].w * xv.w;
    }
    if (row < c.nv) y[row] = a;
  }
}
// out[r] = J[r] . x - sub[r], matrix-free: body spatial vectors X[b] first (one lane per body), then one lane per row
template <int T>
DEV void mat_vec_J(const SolveCtx<T>& c, const ModelDev& m, const float* x, const float* sub, float* out) {
  __syncwarp();
  ROLL
  for (int b = c.lane; b < c.nb; b += 32) {
    sv6 v = zero6();
    ROLL
    for (int k = c.cadr[b], ke = c.cadr[b + 1]; k < ke; k++) { const int i = c.chain[k]; axpy6(v, x[i], ld6(c.cdof() + 6 * i)); }
    st6(c.X() + 6 * b, v);
  }
  __syncwarp();
  ROLL
  for (int r = c.lane; r < c.nefc; r += 32) {
    const int bp = c.bp[r];
    const float2* gp = reinterpret_cast<const float2*>(c.G + (size_t)r * 6);
    const float2 g01 = gp[0];
    float s;
    if (bp < 0) s = g01.x * x[-1 - bp];
    else {
      const float2 g23 = gp[1], g45 = gp[2];
      const float* x2 = c.X() + 6 * (bp >> 16);
      const float* x1 = c.X() + 6 * (bp & 0xffff);
      s = g01.x * (x2[0] - x1[0]) + g01.y * (x2[1] - x1[1]) + g23.x * (x2[2] - x1[2]) + g23.y * (x2[3] - x1[3]) +
          g45.x * (x2[4] - x1[4]) + g45.y * (x2[5] - x1[5]);
    }
    out[r] = sub ? s - sub[r] : s;
  }
  __syncwarp();
}
template <int T>
DEV float gauss_term(const SolveCtx<T>& c) {
  float g = 0.f;
  ROLL
  for (int i = c.lane; i < c.nv; i += 32) g += (c.Ma()[i] - c.qfs()[i]) * (c.qacc()[i] - c.qas()[i]);
  return 0.5f * warp_sum(g);
}
// One elliptic contact (3 rows starting at r) at residuals x[0..2] (mj_constraintUpdate, elliptic branch; oracle cone_eval):
// U = x scaled by (mu, friction), N = U_0, T = |U_1, U_2|; zone 0 top (no force), 1 bottom (all rows quadratic), 2 middle.
struct Cone { int zone; float mu, Dm, N, T, U[3], fri[3]; };
template <int T>
DEV Cone cone_eval(const SolveCtx<T>& c, int r, float x0, float x1, float x2) {
  Cone z;
  z.fri[0] = c.efri()[r]; z.fri[1] = c.efri()[r + 1]; z.fri[2] = c.efri()[r + 2];
  z.mu = z.fri[0];
  z.U[0] = x0 * z.fri[0]; z.U[1] = x1 * z.fri[1]; z.U[2] = x2 * z.fri[2];
  z.N = z.U[0]; z.T = sqrtf(z.U[1] * z.U[1] + z.U[2] * z.U[2]);
  z.Dm = c.D()[r] / fmaxf(z.mu * z.mu * (1.f + z.mu * z.mu), B2S_MINVAL);
  if (z.N >= z.mu * z.T || (z.T <= 0.f && z.N >= 0.f)) z.zone = 0;
  else if (z.mu * z.N + z.T <= 0.f || (z.T <= 0.f && z.N < 0.f)) z.zone = 1;
  else z.zone = 2;
  return z;
}
// forces of the three rows of an elliptic contact
template <int T>
DEV void cone_force(const SolveCtx<T>& c, int r, const Cone& z, float x0, float x1, float x2, float* f) {
  f[0] = f[1] = f[2] = 0.f;
  if (z.zone == 1) { f[0] = -c.D()[r] * x0; f[1] = -c.D()[r + 1] * x1; f[2] = -c.D()[r + 2] * x2; }
  else if (z.zone == 2) {
    f[0] = -z.Dm * (z.N - z.mu * z.T) * z.mu;
    f[1] = -f[0] / z.T * z.U[1] * z.fri[1]; f[2] = -f[0] / z.T * z.U[2] * z.fri[2];
  }
}
// Zone of row r at residual x: 0 inactive, 1 quadratic (the active set of the Hessian), and for friction-loss rows
// (r < nf, only in the FRIC instantiation) 2 / 3 = saturated at force +f / -f outside |x| < R f (mj_constraintUpdate)
template <int T, int V>
DEV int row_zone(const SolveCtx<T>& c, int r, float x) {
  constexpr bool FRIC = (V & 1) != 0, ELL = (V & 2) != 0; (void)FRIC; (void)ELL;   // variant bits: friction-loss rows, elliptic cones
  if (FRIC && r < c.nf) { const float rf = c.frf()[r]; return x <= -rf ? 2 : (x >= rf ? 3 : 1); }
  return x < 0.f ? 1 : 0;
}
template <int T, int V>
DEV float row_force(const SolveCtx<T>& c, int r, float x) {
  constexpr bool FRIC = (V & 1) != 0, ELL = (V & 2) != 0; (void)FRIC; (void)ELL;   // variant bits: friction-loss rows, elliptic cones
  const int z = row_zone<T, V>(c, r, x);
  if (FRIC && z >= 2) return z == 2 ? c.fl()[r] : -c.fl()[r];
  return z == 1 ? -c.D()[r] * x : 0.f;
}
template <int T, int V>
DEV float constraint_cost(const SolveCtx<T>& c, const float* jar) {
  constexpr bool FRIC = (V & 1) != 0, ELL = (V & 2) != 0; (void)FRIC; (void)ELL;   // variant bits: friction-loss rows, elliptic cones
  float s = 0.f;
  ROLL
  for (int r = c.lane; r < c.nefc; r += 32) {
    const float x = jar[r];
    if (ELL && c.ekind()[r]) {
      if (c.ekind()[r] == 1) {        // the lane of the normal row evaluates the whole contact
        const float x1 = jar[r + 1], x2 = jar[r + 2];
        const Cone z = cone_eval<T>(c, r, x, x1, x2);
        if (z.zone == 1) s += 0.5f * (c.D()[r] * x * x + c.D()[r + 1] * x1 * x1 + c.D()[r + 2] * x2 * x2);
        else if (z.zone == 2) { const float nmt = z.N - z.mu * z.T; s += 0.5f * z.Dm * nmt * nmt; }
      }
      continue;
    }
    const int z = row_zone<T, V>(c, r, x);
    if (z == 1) s += 0.5f * c.D()[r] * x * x;
    else if (FRIC && z >= 2) s += -c.fl()[r] * (0.5f * c.frf()[r] + (z == 2 ? x : -x));
  }
  return warp_sum(s);
}
struct LsPoint { float alpha, cost, d0, d1; };
// arguments of the exact (iterative) line search, passed BY VALUE so that this cold, out-of-line path does not
// force the solver context into local memory
struct LsArgs { const float *Jaref, *Jv, *D, *search, *fl, *frf, *efri; const int* ekind; int nefc, nf, nv, lane, maxit; float gtol_unit, q0, q1, q2; };
// E: the elliptic-cone branch is compiled in (the default instantiation keeps the smaller function: code size matters here)
template <bool E>
__device__ __noinline__ LsPoint ls_eval(const LsArgs a, float alpha) {
  float q0 = 0.f, q1 = 0.f, q2 = 0.f, xc = 0.f, xd0 = 0.f, xd1 = 0.f;
#pragma unroll 1
  for (int r = a.lane; r < a.nefc; r += 32) {
    const float ja = a.Jaref[r], jv = a.Jv[r], x = ja + alpha * jv;
    if (E && a.ekind[r]) {      // elliptic contact, evaluated by the lane of its normal row (PrimalEval)
      if (a.ekind[r] != 1) continue;
      const float mu = a.efri[r], f1 = a.efri[r + 1], f2 = a.efri[r + 2];
      const float ja1 = a.Jaref[r + 1], ja2 = a.Jaref[r + 2], jv1 = a.Jv[r + 1], jv2 = a.Jv[r + 2];
      const float U0 = ja * mu, U1 = ja1 * f1, U2 = ja2 * f2, V0 = jv * mu, V1 = jv1 * f1, V2 = jv2 * f2;
      const float UU = U1 * U1 + U2 * U2, UV = U1 * V1 + U2 * V2, VV = V1 * V1 + V2 * V2;
      const float N = U0 + alpha * V0, Tsqr = UU + alpha * (2.f * UV + alpha * VV), Tn = Tsqr > 0.f ? sqrtf(Tsqr) : 0.f;
      if (N >= mu * Tn || (Tsqr <= 0.f && N >= 0.f)) continue;
      if (mu * N + Tn <= 0.f || (Tsqr <= 0.f && N < 0.f)) {
        const float D0 = a.D[r], D1 = a.D[r + 1], D2 = a.D[r + 2];
        q0 += 0.5f * (D0 * ja * ja + D1 * ja1 * ja1 + D2 * ja2 * ja2);
        q1 += D0 * ja * jv + D1 * ja1 * jv1 + D2 * ja2 * jv2;
        q2 += 0.5f * (D0 * jv * jv + D1 * jv1 * jv1 + D2 * jv2 * jv2);
      } else {
        const float Dm = a.D[r] / fmaxf(mu * mu * (1.f + mu * mu), B2S_MINVAL);
        const float N1 = V0, T1 = (UV + alpha * VV) / Tn, T2 = VV / Tn - (UV + alpha * VV) * (UV + alpha * VV) / (Tn * Tn * Tn);
        const float nmt = N - mu * Tn;
        xc += 0.5f * Dm * nmt * nmt; xd0 += Dm * nmt * (N1 - mu * T1); xd1 += Dm * ((N1 - mu * T1) * (N1 - mu * T1) - nmt * mu * T2);
      }
      continue;
    }
    if (r < a.nf) {       // friction-loss row (nf == 0 unless the model has joint friction)
      const float f = a.fl[r], rf = a.frf[r];
      if (x <= -rf) { q0 += f * (-0.5f * rf - ja); q1 += -f * jv; continue; }
      if (x >= rf) { q0 += f * (-0.5f * rf + ja); q1 += f * jv; continue; }
      const float D = a.D[r]; q0 += 0.5f * D * ja * ja; q1 += D * ja * jv; q2 += 0.5f * D * jv * jv;
      continue;
    }
    if (x < 0.f) { const float D = a.D[r]; q0 += 0.5f * D * ja * ja; q1 += D * ja * jv; q2 += 0.5f * D * jv * jv; }
  }
  q0 = warp_sum(q0) + a.q0; q1 = warp_sum(q1) + a.q1; q2 = warp_sum(q2) + a.q2;
  if (E) { xc = warp_sum(xc); xd0 = warp_sum(xd0); xd1 = warp_sum(xd1); }
  LsPoint p;
  p.alpha = alpha; p.cost = alpha * alpha * q2 + alpha * q1 + q0 + xc; p.d0 = 2.f * alpha * q2 + q1 + xd0; p.d1 = fmaxf(2.f * q2 + xd1, B2S_MINVAL);
  return p;
}
template <bool E>
__device__ __noinline__ float ls_iterative(const LsArgs a) {
  float sn = 0.f;
#pragma unroll 1
  for (int i = a.lane; i < a.nv; i += 32) sn += a.search[i] * a.search[i];
  sn = sqrtf(warp_sum(sn));
  if (sn < B2S_MINVAL) return 0.f;
  const float gtol = a.gtol_unit * sn;
  const int maxit = a.maxit;
  int it = 0;
  const LsPoint p0 = ls_eval<E>(a, 0.f);
  LsPoint p1 = ls_eval<E>(a, p0.alpha - p0.d0 / p0.d1); it++;
  if (p0.cost < p1.cost) p1 = p0;
  if (fabsf(p1.d0) < gtol) return p1.alpha;
  const float dir = p1.d0 < 0.f ? 1.f : -1.f;
  LsPoint p2 = p0; bool p2upd = false;
  while (p1.d0 * dir <= -gtol && it < maxit) {
    p2 = p1; p2upd = true;
    p1 = ls_eval<E>(a, p1.alpha - p1.d0 / p1.d1); it++;
    if (fabsf(p1.d0) < gtol) return p1.alpha;
  }
  if (it >= maxit || !p2upd) return p1.alpha;
  LsPoint p2next = p1, p1next = ls_eval<E>(a, p1.alpha - p1.d0 / p1.d1); it++;
  while (it < maxit) {
    const LsPoint pmid = ls_eval<E>(a, 0.5f * (p1.alpha + p2.alpha)); it++;
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
    if (b1) { p1next = ls_eval<E>(a, p1.alpha - p1.d0 / p1.d1); it++; }
    if (b2) { p2next = ls_eval<E>(a, p2.alpha - p2.d0 / p2.d1); it++; }
  }
  if (p1.cost <= p2.cost && p1.cost < p0.cost) return p1.alpha;
  if (p2.cost <= p1.cost && p2.cost < p0.cost) return p2.alpha;
  return 0.f;
}
// one lane per candidate step: alpha_k = min_step^(k/(n-1)), k = 0..n-1, and alpha = 0 on one more lane; rows are
// summed in row order by every lane (same order as the oracle), the best candidate wins (ties: alpha = 0, then
// the smallest k), so a step is only taken if it strictly improves on alpha = 0
template <int T, int V>
DEV float ls_parallel(const SolveCtx<T>& c, const ModelDev& m, const float* qg) {
  constexpr bool FRIC = (V & 1) != 0, ELL = (V & 2) != 0; (void)FRIC; (void)ELL;   // variant bits: friction-loss rows, elliptic cones
  const int n = m.ls_iterations;
  const float lg = logf(m.ls_parallel_min_step);
  float best = 3.0e38f, balpha = 0.f; int bk = 0x7fffffff;
  for (int base = 0; base <= n; base += 32) {
    const int k = base + c.lane;
    if (k <= n) {
      const float alpha = k == n ? 0.f : (n > 1 ? expf((float)k / (float)(n - 1) * lg) : 1.f);
      float q0 = qg[0], q1 = qg[1], q2 = qg[2], xc = 0.f;
      for (int r = 0; r < c.nefc; r++) {
        const float ja = c.Jaref()[r], jv = c.Jv()[r];
        if (ELL && c.ekind()[r]) {      // elliptic contact: N(alpha), T(alpha) and the zone at this candidate (PrimalEval)
          const float ja1 = c.Jaref()[r + 1], ja2 = c.Jaref()[r + 2], jv1 = c.Jv()[r + 1], jv2 = c.Jv()[r + 2];
          const Cone z = cone_eval<T>(c, r, ja + alpha * jv, ja1 + alpha * jv1, ja2 + alpha * jv2);
          if (z.zone == 1) {
            const float D0 = c.D()[r], D1 = c.D()[r + 1], D2 = c.D()[r + 2];
            q0 += 0.5f * (D0 * ja * ja + D1 * ja1 * ja1 + D2 * ja2 * ja2);
            q1 += D0 * ja * jv + D1 * ja1 * jv1 + D2 * ja2 * jv2;
            q2 += 0.5f * (D0 * jv * jv + D1 * jv1 * jv1 + D2 * jv2 * jv2);
          } else if (z.zone == 2) { const float nmt = z.N - z.mu * z.T; xc += 0.5f * z.Dm * nmt * nmt; }
          r += 2;
          continue;
        }
        const int z = row_zone<T, V>(c, r, ja + alpha * jv);
        if (z == 1) { const float D = c.D()[r]; q0 += 0.5f * D * ja * ja; q1 += D * ja * jv; q2 += 0.5f * D * jv * jv; }
        else if (FRIC && z >= 2) { const float f = c.fl()[r], sg = z == 2 ? -1.f : 1.f; q0 += f * (-0.5f * c.frf()[r] + sg * ja); q1 += sg * f * jv; }
      }
      const float cost = ELL ? alpha * alpha * q2 + alpha * q1 + q0 + xc : alpha * alpha * q2 + alpha * q1 + q0;
      const int kk = k == n ? -1 : k;
      if (cost < best || (cost == best && kk < bk)) { best = cost; balpha = alpha; bk = kk; }
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float oc = __shfl_xor_sync(FULL, best, o), oa = __shfl_xor_sync(FULL, balpha, o);
    const int ok = __shfl_xor_sync(FULL, bk, o);
    if (oc < best || (oc == best && ok < bk)) { best = oc; balpha = oa; bk = ok; }
  }
  __syncwarp();
  return balpha;
}

// A run of contact rows that share the body pair (b1, b2) has been reduced to K (21 lanes) and W (6 lanes):
// apply it to qfrc_constraint and (withH) to the lower triangle of H.
template <int T>
DEV void flush_run(const SolveCtx<T>& c, const ModelDev& m, int bp, float acc, bool withH) {
  c.KW()[c.lane] = acc;
  const int b1 = bp & 0xffff, b2 = bp >> 16;
  const int a1 = c.cadr[b1], n1f = c.cadr[b1 + 1] - a1;
  const int a2 = c.cadr[b2], n2f = c.cadr[b2 + 1] - a2;
  // common ancestors are a common prefix of the two root-first chains: they cancel
  int cp = 0;
  {
    const int nmin = min(n1f, n2f);
    for (int base = 0; base < nmin; base += 32) {
      const int k = base + c.lane;
      const bool same = k < nmin && c.chain[a1 + k] == c.chain[a2 + k];
      const unsigned mk = __ballot_sync(FULL, same);
      if (mk != FULL) { cp = base + __ffs(~mk) - 1; break; }
      cp = base + 32;
    }
    cp = min(cp, nmin);
  }
  const int n2 = n2f - cp, n1 = n1f - cp, ns = n1 + n2;
  B2S_ASSERT(b1 >= 0 && b1 < c.nb && b2 >= 0 && b2 < c.nb && ns >= 0 && ns <= 2 * m.maxchain);
  __syncwarp();
  const float* K = c.KW(); const float* W = c.KW() + 21;
  float* Y = c.X();    // 6 floats per support entry (the body-vector scratch is idle during assembly)
  for (int base = 0; base < ns; base += 32) {
    // supports longer than 32 dofs are processed in chunks; Y/Li then hold one chunk at a time, so the cross
    // terms between chunks are handled by the (rare) slow path below
    const int a = base + c.lane;
    if (a < ns) {
      const int i = a < n2 ? c.chain[a2 + cp + a] : c.chain[a1 + cp + a - n2];
      const float s = a < n2 ? 1.f : -1.f;
      const sv6 cd = ld6(c.cdof() + 6 * i);
      c.qfc()[i] += s * (cd.v[0] * W[0] + cd.v[1] * W[1] + cd.v[2] * W[2] + cd.v[3] * W[3] + cd.v[4] * W[4] + cd.v[5] * W[5]);
      if (withH && ns <= 32) {
        // Y = s K cd with K symmetric, entries ordered (0,0) (1,0) (1,1) (2,0) ...
        float y[6];
        int e = 0;
#pragma unroll
        for (int p = 0; p < 6; p++) y[p] = 0.f;
#pragma unroll
        for (int p = 0; p < 6; p++)
#pragma unroll
          for (int q = 0; q <= p; q++) {
            const float kv = K[e++];
            y[p] += kv * cd.v[q];
            if (q != p) y[q] += kv * cd.v[p];
          }
#pragma unroll
        for (int p = 0; p < 6; p++) Y[6 * a + p] = s * y[p];
        c.Li()[a] = i | (a < n2 ? 0 : 0x10000);
      }
    }
  }
  __syncwarp();
  if (withH) {
    if (ns <= 32) {
      const int T2 = ns * (ns + 1) / 2;
      for (int t = c.lane; t < T2; t += 32) {
        int a = (int)((sqrtf(8.f * (float)t + 1.f) - 1.f) * 0.5f);
        while (a * (a + 1) / 2 > t) a--;
        while ((a + 1) * (a + 2) / 2 <= t) a++;
        const int b = t - a * (a + 1) / 2;
        const int la = c.Li()[a], lb = c.Li()[b];
        const int i = la & 0xffff, j = lb & 0xffff;
        const float s = (la & 0x10000) ? -1.f : 1.f;
        const float* cd = c.cdof() + 6 * i;
        const float* y = Y + 6 * b;
        const float v = s * (cd[0] * y[0] + cd[1] * y[1] + cd[2] * y[2] + cd[3] * y[3] + cd[4] * y[4] + cd[5] * y[5]);
        c.H()[max(i, j) * c.ld + min(i, j)] += v;
      }
    } else {
      // slow path (support > 32 dofs): every lane walks whole (i, j) pairs recomputing K cd_j on the fly
      const int T2 = ns * (ns + 1) / 2;
      for (int t = c.lane; t < T2; t += 32) {
        int a = (int)((sqrtf(8.f * (float)t + 1.f) - 1.f) * 0.5f);
        while (a * (a + 1) / 2 > t) a--;
        while ((a + 1) * (a + 2) / 2 <= t) a++;
        const int b = t - a * (a + 1) / 2;
        const int i = a < n2 ? c.chain[a2 + cp + a] : c.chain[a1 + cp + a - n2];
        const int j = b < n2 ? c.chain[a2 + cp + b] : c.chain[a1 + cp + b - n2];
        const float s = ((a < n2) == (b < n2)) ? 1.f : -1.f;
        const float* ci = c.cdof() + 6 * i; const float* cj = c.cdof() + 6 * j;
        float v = 0.f; int e = 0;
        for (int p = 0; p < 6; p++) for (int q = 0; q <= p; q++) {
          const float kv = K[e++];
          v += kv * ci[p] * cj[q];
          if (q != p) v += kv * ci[q] * cj[p];
        }
        c.H()[max(i, j) * c.ld + min(i, j)] += s * v;
      }
    }
  }
  __syncwarp();
}

// Forces / active set from Jaref, qfrc_constraint = J^T f, and (withH) J^T D J into the (zeroed) lower triangle
// of H in shared memory; M is added when the tiles are loaded for the factorisation.
template <int T, int V>
DEV void assemble(const SolveCtx<T>& c, const ModelDev& m, bool withH, int ia, int ib) {
  constexpr bool FRIC = (V & 1) != 0, ELL = (V & 2) != 0; (void)FRIC; (void)ELL;   // variant bits: friction-loss rows, elliptic cones
  __syncwarp();
  ROLL
  for (int i = c.lane; i < c.nvp; i += 32) c.qfc()[i] = 0.f;
  if (withH) {
    float4* h4 = reinterpret_cast<float4*>(c.H());
    _Pragma("unroll 2")
    for (int i = c.lane; i < c.hfloats / 4; i += 32) h4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  __syncwarp();
  // limit rows (they come first): one lane per row, J is +-1 on one dof
  ROLL
  for (int r = c.lane; r < c.nl; r += 32) {
    const float jar = c.Jaref()[r];
    const int z = row_zone<T, V>(c, r, jar);
    if (z) {
      const float g0 = c.G[(size_t)r * 6];
      const int dof = -1 - c.bp[r];
      atomicAdd(&c.qfc()[dof], g0 * row_force<T, V>(c, r, jar));
      if (withH && z == 1) atomicAdd(&c.H()[dof * c.ld + dof], c.D()[r]);
    }
  }
  __syncwarp();
  // contact rows: each run of rows with the same body pair is reduced to K (lanes 0..20) and W (lanes 21..26)
  const float* gA = c.G + ia;
  const float* gB = c.G + ib;
  const bool isK = c.lane < 21;
  for (int k = 0; k < c.nrun; k++) {
    const int r0 = c.run()[k], r1 = c.run()[k + 1];
    float acc = 0.f;
    bool any = false;
#pragma unroll 4
    for (int r = r0; r < r1; r++) {
      const float jar = c.Jaref()[r];
      if (ELL && c.ekind()[r]) {
        // elliptic contact (3 rows): W += sum_a f_a g_a, K += sum_ab Hc[a][b] g_a g_b^T with Hc = diag(D) in the bottom zone and
        // the cone Hessian Dm diag(fri) [1, -mu U'/T; ., mu N/T^3 U U' + (mu^2 - mu N/T) I] diag(fri) in the middle zone
        const float x1 = c.Jaref()[r + 1], x2 = c.Jaref()[r + 2];
        const Cone z = cone_eval<T>(c, r, jar, x1, x2);
        if (z.zone) {
          float f[3], ga3[3], gb3[3];
          cone_force<T>(c, r, z, jar, x1, x2, f);
#pragma unroll
          for (int a = 0; a < 3; a++) { ga3[a] = gA[(size_t)(r + a) * 6]; gb3[a] = gB[(size_t)(r + a) * 6]; }
          if (!isK) acc += f[0] * ga3[0] + f[1] * ga3[1] + f[2] * ga3[2];
          else if (z.zone == 1) acc += c.D()[r] * ga3[0] * gb3[0] + c.D()[r + 1] * ga3[1] * gb3[1] + c.D()[r + 2] * ga3[2] * gb3[2];
          else {
            const float it = 1.f / z.T, s3 = z.mu * z.N * it * it * it, dg = z.mu * z.mu - z.mu * z.N * it;
#pragma unroll
            for (int a = 0; a < 3; a++)
#pragma unroll
              for (int b = 0; b < 3; b++) {
                float h;
                if (a == 0 && b == 0) h = 1.f;
                else if (a == 0 || b == 0) h = -z.mu * z.U[a + b] * it;
                else h = s3 * z.U[a] * z.U[b] + (a == b ? dg : 0.f);
                acc += h * z.Dm * z.fri[a] * z.fri[b] * ga3[a] * gb3[b];
              }
          }
          any = true;
        }
        r += 2;
        continue;
      }
      const float De = jar < 0.f ? c.D()[r] : 0.f;
      const float ga = gA[(size_t)r * 6], gb = gB[(size_t)r * 6];
      acc += (isK ? De * gb : -De * jar) * ga;
      any |= jar < 0.f;
    }
    if (any) flush_run(c, m, c.bp[r0], acc, withH);
  }
  __syncwarp();
}

// Active set (rows with Jaref < 0) as a bit mask in shared memory; returns true when it equals the mask of the previous
// call.  H = M + J^T D J depends on the state only through this set (pyramidal cones: D is fixed per row), so an
// unchanged set means the factor of the previous iteration is still the factor of H (MuJoCo's Newton solver also
// refactors only on a change of the active set).
template <int T, int V>
DEV bool active_set_unchanged(const SolveCtx<T>& c) {
  constexpr bool FRIC = (V & 1) != 0, ELL = (V & 2) != 0; (void)FRIC; (void)ELL;   // variant bits: friction-loss rows, elliptic cones
  bool same = true;
  ROLL
  for (int base = 0; base < c.nefc; base += 32) {
    const int r = base + c.lane;
    bool on = false;
    if (r < c.nefc) {
      if (ELL && c.ekind()[r]) {
        const int r0 = r - (c.ekind()[r] == 1 ? 0 : (c.ekind()[r - 1] == 1 ? 1 : 2));      // normal row of this contact
        const Cone z = cone_eval<T>(c, r0, c.Jaref()[r0], c.Jaref()[r0 + 1], c.Jaref()[r0 + 2]);
        on = z.zone == 1;
        if (z.zone == 2) same = false;         // the cone Hessian changes with the state: never reuse the factor
      } else on = row_zone<T, V>(c, r, c.Jaref()[r]) == 1;
    }
    const unsigned mk = __ballot_sync(FULL, on);
    same &= c.amask()[base >> 5] == mk;
    __syncwarp();
    if (c.lane == 0) c.amask()[base >> 5] = mk;
  }
  __syncwarp();
  return ELL ? __all_sync(FULL, same) : same;
}

// Load the lower-triangular register tiles of H = M + (J^T D J in shared memory) and factorise.
template <int T>
DEV void factor_H(const SolveCtx<T>& c, int bi, int bj) {
  float a[T * T];
  if (bi >= 0) {
    const bool diag = bi == bj;
    const float* hs = c.H() + (bi * T) * c.ld + bj * T;
    const float* mg = c.M + (size_t)(bi * T) * c.nvp + bj * T;
    if (c.nv == 7 * T) {     // no padding: every (i, j) of the tile is a real entry
#pragma unroll
      for (int x = 0; x < T; x++)
#pragma unroll
        for (int y = 0; y < T; y++) a[x * T + y] = hs[x * c.ld + y] + mg[x * c.nvp + y];
    } else {
#pragma unroll
      for (int x = 0; x < T; x++)
#pragma unroll
        for (int y = 0; y < T; y++) {
          const int i = bi * T + x, j = bj * T + y;
          a[x * T + y] = (i < c.nv && j < c.nv) ? hs[x * c.ld + y] + mg[x * c.nvp + y] : (i == j ? 1.0f : 0.0f);
        }
    }
    // diagonal tiles: only the lower half was assembled (M is symmetric, H's upper triangle is zero); mirror it
#pragma unroll
    for (int x = 0; x < T; x++)
#pragma unroll
      for (int y = x + 1; y < T; y++) if (diag) a[x * T + y] = a[y * T + x];
  }
  __syncwarp();
  tile_cholesky_regs<T>(a, bi, bj, c.H(), c.ld, c.invd());
}

// big = 0: shared-memory layout with rows_cap(njmax, V) rows, worlds with more rows are left to the second launch (big = 1: the
// full layout, only worlds the first launch skipped)
__host__ __device__ inline int rows_cap(int njmax, int V, int big, int cap) { return (V == 0 && !big && njmax > cap) ? cap : njmax; }
template <int T, int V, bool BIG = false>
__global__ void __launch_bounds__(32 * SOLVE_WARPS, SOLVE_MINB) k_solve(const ModelDev m, const DataDev d) {
  constexpr int big = BIG ? 1 : 0;
  constexpr bool FRIC = (V & 1) != 0, ELL = (V & 2) != 0; (void)FRIC; (void)ELL;   // variant bits: friction-loss rows, elliptic cones
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#ifdef SOLVE_TIMING
  const long long t_start = clock64();      // development builds only (tools/exp_solve_time.py): per-world solve time in cycles -> solver_cost
#endif
  // big = 0: worlds are taken from the cost-ordered queues, most expensive class first (k_constraint filled them), one per CTA.
  // big = 1 (a small grid, normally nothing to do): CTA b looks at worlds 16 b .. 16 b + 15 and solves those with more than
  // solve_rowcap rows one after the other.
#pragma unroll 1
  for (int slot = 0; slot < (big ? SOLVE_BIG_SPAN : 1); slot++) {
  int w;
  if (big) {
    w = blockIdx.x * SOLVE_BIG_SPAN + slot;
    if (w >= d.nworld) break;
    if (d.nefc[w] <= m.solve_rowcap) continue;
  } else {
    w = blockIdx.x * SOLVE_WARPS + warp;
    if (w >= d.nworld) return;
    int cls = B2S_SOLVE_CLASSES - 1, before = 0;
#pragma unroll 1
    for (; cls > 0; cls--) { const int n = d.solve_count[cls]; if (w < before + n) break; before += n; }
    B2S_ASSERT(w - before >= 0 && w - before < d.solve_count[cls]);      // every queue slot below the class count was filled
    w = d.solve_order[(size_t)cls * d.nworld + (w - before)];
    B2S_ASSERT(w >= 0 && w < d.nworld);
  }
  const int nv = m.nv, nvp = m.nvp, nb = m.nbody, njmax = d.njmax;
  const int rcap = rows_cap(njmax, V, big, m.solve_rowcap);       // rows the shared-memory layout holds (global arrays keep the stride njmax)
  SolveCtx<T> c;
  c.lane = lane; c.nv = nv; c.nvp = nvp; c.nb = nb; c.nefc = d.nefc[w]; c.nl = d.nl[w]; c.nf = FRIC ? d.nf[w] : 0; c.njmax = rcap; c.nconmax = d.nconmax;
  if (V == 0 && njmax > m.solve_rowcap && (c.nefc > m.solve_rowcap) != (big != 0)) continue;      // the other launch's world
  c.soff = solve_soff(rcap, d.nconmax, T);
  c.G = d.efc_g + (size_t)w * njmax * 6;
  c.bp = d.efc_bp + (size_t)w * njmax;
  c.aref = d.efc_aref + (size_t)w * njmax;
  float* const wbase = b2s_smem + (size_t)warp * solve_smem_floats(nv, nvp, nb, rcap, d.nconmax, T, m.nchain, ELL ? 1 : 0, V == 0 ? 0 : 4);
  float* const Msm = wbase;                                                    // staged mass matrix (16-byte aligned)
  unsigned long long* const mbar = reinterpret_cast<unsigned long long*>(wbase + solve_stage_front_floats(T) - 4);
  c.M = Msm;
  c.sm = wbase + solve_stage_front_floats(T);
  {
    float* tail = c.sm + c.soff + solve_scratch_floats(nb) + (ELL ? 2 * rcap : 0);
    c.opt = c.sm + solve_opt_off(nb, rcap, d.nconmax, T, m.nchain, ELL ? 1 : 0);
    c.chain = reinterpret_cast<const unsigned char*>(tail);
    c.cadr = reinterpret_cast<const unsigned short*>(tail + ((m.nchain + 3) >> 2));
  }
  float* efc_force = d.efc_force + (size_t)w * njmax;
  float* g_qacc = d.qacc + (size_t)w * nv;
  float* g_qfc = d.qfrc_constraint + (size_t)w * nv;
  const bool warm = !(m.disableflags & B2S_DSBL_WARMSTART);
#pragma unroll 1
  for (int i = lane; i < SolveCtx<T>::NVP; i += 32) {
    c.qas()[i] = i < nv ? d.qacc_smooth[(size_t)w * nv + i] : 0.f;
    c.qfs()[i] = i < nv ? d.qfrc_smooth[(size_t)w * nv + i] : 0.f;
    c.search()[i] = (i < nv && warm) ? d.qacc_warmstart[(size_t)w * nv + i] : 0.f;   // candidate start point
    c.qacc()[i] = 0.f; c.qfc()[i] = 0.f;
  }
  if (c.nefc == 0) {
    __syncwarp();
#pragma unroll 1
    for (int i = lane; i < nv; i += 32) { g_qacc[i] = c.qas()[i]; g_qfc[i] = 0.f; }
    if (lane == 0) { d.solver_niter[w] = 0; d.solver_cost[w] = 0.f; }
    continue;
  }
  // ---- staging.  The world's mass matrix comes in by ONE bulk async copy (TMA, cp.async.bulk): lane 0 arms the mbarrier with
  // the byte count and issues it, the copy engine fills shared memory while the warp runs the rest of the prologue, and
  // the warp waits on the barrier just before the first M product.  (Issued after the nefc == 0 exit: a CTA must not retire
  // with a copy into its shared memory in flight.)  The model's chain tables are tiny and identical for every world: plain
  // loads (L1 / L2 hits), narrowed to bytes / 16-bit words.
  if (lane == 0) mbar_init(mbar, 1);
  __syncwarp();
  if (lane == 0) tma_load_1d(Msm, d.qM + (size_t)w * nv * nvp, (unsigned)(nv * nvp * sizeof(float)), mbar);
  {
    unsigned char* ch = const_cast<unsigned char*>(c.chain);
    unsigned short* ca = const_cast<unsigned short*>(c.cadr);
#pragma unroll 1
    for (int k = lane; k < m.nchain; k += 32) ch[k] = (unsigned char)m.body_chain[k];
#pragma unroll 1
    for (int b = lane; b <= nb; b += 32) ca[b] = (unsigned short)m.body_chain_adr[b];
  }
  // cdof about the world's reference point (other kinematic trees are re-referenced), efc_D of the rows
  {
    const float* com = d.subtree_com + (size_t)w * nb * 3;
    const v3 cref = ld3(com + 3 * m.ref_root);
#pragma unroll 1
    for (int i = lane; i < nv; i += 32) {
      sv6 cd = ld6(d.cdof + ((size_t)w * nv + i) * 6);
      const int rb = m.body_rootid[m.dof_bodyid[i]];
      if (rb != m.ref_root) {
        const v3 lin = mk3(cd.v[3], cd.v[4], cd.v[5]) + cross(mk3(cd.v[0], cd.v[1], cd.v[2]), cref - ld3(com + 3 * rb));
        cd.v[3] = lin.x; cd.v[4] = lin.y; cd.v[5] = lin.z;
      }
      st6(c.cdof() + 6 * i, cd);
    }
#pragma unroll 1
    for (int r = lane; r < c.nefc; r += 32) c.D()[r] = d.efc_D[(size_t)w * njmax + r];
    if (ELL) {
      // rows of elliptic contacts: kind (normal / friction row) and scale (regularised mu = friction_1 sqrt(R_1 / R_n); friction_j)
#pragma unroll 1
      for (int r = lane; r < c.nefc; r += 32) {
        int kind = 0; float fr = 0.f;
        if (d.efc_type[(size_t)w * njmax + r] == 6) {
          const int con = d.efc_id[(size_t)w * njmax + r];
          const int k = r - d.con_efc_address[(size_t)w * d.nconmax + con];
          const float* cf = d.con_friction + ((size_t)w * d.nconmax + con) * 5;
          kind = k == 0 ? 1 : 2;
          fr = k == 0 ? cf[0] * sqrtf(d.efc_D[(size_t)w * njmax + r] / d.efc_D[(size_t)w * njmax + r + 1]) : cf[k - 1];
        }
        c.ekind()[r] = kind; c.efri()[r] = fr;
      }
    }
    if (FRIC) {
#pragma unroll 1
      for (int r = lane; r < c.nf; r += 32) {
        const float f = d.efc_frictionloss[(size_t)w * njmax + r];
        c.fl()[r] = f; c.frf()[r] = f / d.efc_D[(size_t)w * njmax + r];
      }
    }
  }
  // runs of contact rows that share a body pair (fixed for the whole solve), in row order
  {
    int nrun = 0;
#pragma unroll 1
    for (int base_r = c.nl; base_r < c.nefc; base_r += 32) {
      const int r = base_r + lane;
      const bool start = r < c.nefc && (r == c.nl || c.bp[r] != c.bp[r - 1]);
      const unsigned mk = __ballot_sync(FULL, start);
      if (start) c.run()[nrun + __popc(mk & ((1u << lane) - 1u))] = r;
      nrun += __popc(mk);
    }
    B2S_ASSERT(nrun <= c.nconmax && c.nefc <= njmax && c.nl <= c.nefc);      // run list holds nconmax + 2 entries
    if (lane == 0) c.run()[nrun] = c.nefc;
    c.nrun = nrun;
  }
  int bi, bj;
  tile_owner(lane, bi, bj);
  // K / W entry owned by this lane in the run reduction: lanes 0..20 -> (ia, ib) of the symmetric 6x6, 21..26 -> W
  int ia = 0, ib = 0;
  if (lane < 21) { int e = lane; while (e > ia) { e -= ia + 1; ia++; } ib = e; }
  else if (lane < 27) { ia = lane - 21; ib = 0; }
  __syncwarp();
  // ---- warm start (mj_fwdConstraint): evaluate the cost at qacc_warmstart (held in c.search()) and at qacc_smooth;
  // keep the warm start only if it is not worse.  One code path evaluates both points (pass 0: smooth, pass 1: warm).
  float gauss = 0.f, cost = 0.f;
  mbar_wait(mbar, 0);       // the staged mass matrix has landed (phase 0 of the barrier)
#pragma unroll 1
  for (int pass = warm ? 0 : 1; pass < 2; pass++) {
    const float* x = (pass == 0 || !warm) ? c.qas() : c.search();
    mat_vec_J(c, m, x, c.aref, c.Jv());
    float cst = constraint_cost<T, V>(c, c.Jv());
    mat_vec_M(c, x, c.Mv());
    __syncwarp();
    float g = 0.f;
#pragma unroll 1
    for (int i = lane; i < nv; i += 32) g += (c.Mv()[i] - c.qfs()[i]) * (x[i] - c.qas()[i]);
    const float gs = 0.5f * warp_sum(g);      // exactly 0 at qacc_smooth
    // pass 0 evaluates qacc_smooth and is always adopted; pass 1 (the warm start) replaces it unless it is worse.
    // The comparison is on the constraint cost + Gauss term of the warm start vs the constraint cost at qacc_smooth.
    if (pass == 0 || !warm || !(cst + gs > cost)) {
#pragma unroll 1
      for (int i = lane; i < nv; i += 32) { c.qacc()[i] = x[i]; c.Ma()[i] = c.Mv()[i]; }
#pragma unroll 1
      for (int r = lane; r < c.nefc; r += 32) c.Jaref()[r] = c.Jv()[r];
      gauss = gs; cost = cst + gs;
    }
    __syncwarp();
  }
  const float scale = 1.0f / (m.meaninertia * (float)max(nv, 1));
  int iter = 0;
  bool withH = true, haveL = false;
  // variant bit 4: the exact line search and the CG solver are compiled in; without it the kernel is Newton + parallel
  // line search only (the reference's default, sim.py:55,91) and ~1 k instructions of cold code smaller
  constexpr bool FULL_OPTS = (V & 4) != 0;
  const bool newton = !FULL_OPTS || m.solver == B2S_SOL_NEWTON;
  if (!newton) {
    // CG (mj_solPrimal without the Hessian): the preconditioner is M^-1, so M is factorised once (H storage = 0 + M)
    float4* h4 = reinterpret_cast<float4*>(c.H());
    _Pragma("unroll 2")
    for (int i = lane; i < c.hfloats / 4; i += 32) h4[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    __syncwarp();
    factor_H<T>(c, bi, bj);
    haveL = true;
  }
#pragma unroll 1
  while (true) {
    // the factor in shared memory stays valid while the active set does not change
    const bool reuse = (active_set_unchanged<T, V>(c) && haveL) || !newton;
    assemble<T, V>(c, m, withH && !reuse, ia, ib);
    if (!withH) break;
    float gn = 0.f;
#pragma unroll 1
    for (int i = lane; i < nv; i += 32) { const float g = c.Ma()[i] - c.qfs()[i] - c.qfc()[i]; c.grad()[i] = g; gn += g * g; }
    gn = sqrtf(warp_sum(gn));
    if (iter > 0 && scale * gn < m.tolerance) break;
    if (iter >= m.iterations) break;
    // blocked Cholesky of H on register tiles (c.H() ends up holding L), search = -H^-1 grad
    if (!reuse) factor_H<T>(c, bi, bj);
    haveL = true;
    float b0 = lane < nv ? c.grad()[lane] : 0.f, b1 = lane + 32 < nv ? c.grad()[lane + 32] : 0.f;
    tile_chol_solve<T>(c.H(), c.invd(), c.ld, lane, b0, b1);
    if (newton) {
      if (lane < nv) c.search()[lane] = -b0;
      if (lane + 32 < nv) c.search()[lane + 32] = -b1;
    } else {
      // Polak-Ribiere: beta = grad . (Mgrad - Mgrad_old) / max(MINVAL, grad_old . Mgrad_old), clamped at 0; first pass beta = 0
      float num = 0.f, den = 0.f;
      if (iter > 0) {
        if (lane < nv) { num += c.grad()[lane] * (b0 - c.Mgradold()[lane]); den += c.gradold()[lane] * c.Mgradold()[lane]; }
        if (lane + 32 < nv) { num += c.grad()[lane + 32] * (b1 - c.Mgradold()[lane + 32]); den += c.gradold()[lane + 32] * c.Mgradold()[lane + 32]; }
      }
      num = warp_sum(num); den = warp_sum(den);
      const float beta = iter > 0 ? fmaxf(num / fmaxf(den, B2S_MINVAL), 0.f) : 0.f;
      if (lane < nv) { c.search()[lane] = -b0 + beta * c.search()[lane]; c.Mgradold()[lane] = b0; c.gradold()[lane] = c.grad()[lane]; }
      if (lane + 32 < nv) { c.search()[lane + 32] = -b1 + beta * c.search()[lane + 32]; c.Mgradold()[lane + 32] = b1; c.gradold()[lane + 32] = c.grad()[lane + 32]; }
    }
    __syncwarp();
    // line-search set-up: Mv, Jv, quadratic Gauss coefficients
    mat_vec_M(c, c.search(), c.Mv());
    mat_vec_J(c, m, c.search(), nullptr, c.Jv());
    float g1 = 0.f, g2 = 0.f;
#pragma unroll 1
    for (int i = lane; i < nv; i += 32) { g1 += c.search()[i] * (c.Ma()[i] - c.qfs()[i]); g2 += c.search()[i] * c.Mv()[i]; }
    float qg[3] = {gauss, warp_sum(g1), 0.5f * warp_sum(g2)};
    float alpha;
    if (!FULL_OPTS || m.ls_parallel) alpha = ls_parallel<T, V>(c, m, qg);
    else {
      LsArgs la;
      la.Jaref = c.Jaref(); la.Jv = c.Jv(); la.D = c.D(); la.search = c.search(); la.nefc = c.nefc; la.nf = c.nf; la.fl = c.fl(); la.frf = c.frf(); la.ekind = ELL ? c.ekind() : nullptr; la.efri = ELL ? c.efri() : nullptr; la.nv = nv; la.lane = lane;
      la.maxit = m.ls_iterations; la.gtol_unit = m.tolerance * m.ls_tolerance * m.meaninertia * (float)max(nv, 1);
      la.q0 = qg[0]; la.q1 = qg[1]; la.q2 = qg[2];
      alpha = ls_iterative<ELL>(la);
    }
    if (alpha == 0.f) break;
#pragma unroll 1
    for (int i = lane; i < nv; i += 32) { c.qacc()[i] += alpha * c.search()[i]; c.Ma()[i] += alpha * c.Mv()[i]; }
#pragma unroll 1
    for (int r = lane; r < c.nefc; r += 32) c.Jaref()[r] += alpha * c.Jv()[r];
    __syncwarp();
    const float old = cost;
    gauss = gauss_term(c);
    cost = constraint_cost<T, V>(c, c.Jaref()) + gauss;
    iter++;
    if (scale * (old - cost) < m.tolerance) withH = false;     // converged: one more pass for the final forces only
  }
  __syncwarp();
#pragma unroll 1
  for (int i = lane; i < nv; i += 32) { g_qacc[i] = c.qacc()[i]; g_qfc[i] = c.qfc()[i]; }
#pragma unroll 1
  for (int r = lane; r < c.nefc; r += 32) {
    if (ELL && c.ekind()[r]) {
      if (c.ekind()[r] == 1) {
        const float x0 = c.Jaref()[r], x1 = c.Jaref()[r + 1], x2 = c.Jaref()[r + 2];
        float f[3];
        cone_force<T>(c, r, cone_eval<T>(c, r, x0, x1, x2), x0, x1, x2, f);
        efc_force[r] = f[0]; efc_force[r + 1] = f[1]; efc_force[r + 2] = f[2];
      }
      continue;
    }
    efc_force[r] = row_force<T, V>(c, r, c.Jaref()[r]);
  }
  if (lane == 0) { d.solver_niter[w] = iter; d.solver_cost[w] = cost; }
#ifdef SOLVE_TIMING
  if (lane == 0) d.solver_cost[w] = (float)(clock64() - t_start);
#endif
  if (big) {      // the next world of this CTA re-arms the barrier and reuses the shared memory
    __syncwarp();
    if (lane == 0) asm volatile("mbarrier.inval.shared::cta.b64 [%0];" ::"r"(smem_u32(mbar)) : "memory");
    __syncwarp();
  }
  }
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

// the friction-loss instantiation is only used by models that can have such rows (ModelDev::fric): the default kernel
// carries none of that code
static int solve_variant(const ModelDev& m) {
  if (!m.ls_parallel || m.solver != B2S_SOL_NEWTON) return 7;
  if (m.cone == B2S_CONE_ELLIPTIC) return 3;
  return m.fric ? 1 : 0;
}
static size_t smem_solve_c(const ModelDev& m, const DataDev& d, int V, int rows) {
  // the elliptic per-row arrays exist in every instantiation that has the elliptic code (variants 3 and 7)
  return (size_t)SOLVE_WARPS * solve_smem_floats(m.nv, m.nvp, m.nbody, rows, d.nconmax, tile_for(m.nv), m.nchain, V >= 3 ? 1 : 0, V == 0 ? 0 : 4) * sizeof(float);
}
// CTAs per SM at a given shared-memory request: 228 KB per SM, 1 KB reserved per CTA, and the register bound of __launch_bounds__
static int solve_slots(size_t smem) { const int s = (int)(233472 / (smem + 1024)); return s < SOLVE_MINB ? s : SOLVE_MINB; }
// Row capacity the default variant's main launch really uses: the requested one (Model.opt.solve_rowcap) where the smaller layout
// buys CTAs per SM (G1: 12 -> 14), njmax (one launch, no split) where it does not (a 20-dof Asimov is register-bound at 16 either
// way, and its few worlds with > 128 rows would only pay for the second launch).
static int solve_eff_cap(const ModelDev& m, const DataDev& d) {
  if (solve_variant(m) != 0 || d.njmax <= m.solve_rowcap) return d.njmax;
  return solve_slots(smem_solve_c(m, d, 0, m.solve_rowcap)) > solve_slots(smem_solve_c(m, d, 0, d.njmax)) ? m.solve_rowcap : d.njmax;
}
bool solve_has_second_launch(const ModelDev& m, const DataDev& d) { return solve_eff_cap(m, d) < d.njmax; }
// `side` (optional): the stream the second launch goes to; the caller has already made it wait for k_constraint and joins it after
// the call.  The second launch normally finds nothing to do, so beside the main launch it costs nothing; behind it, it cost 7 us.
void launch_solve(const ModelDev& m0, const DataDev& d, cudaStream_t st, cudaStream_t side) {
  const int grid = (d.nworld + SOLVE_WARPS - 1) / SOLVE_WARPS;
  const int V = solve_variant(m0);
  ModelDev m = m0;
  m.solve_rowcap = solve_eff_cap(m0, d);      // the kernels read the capacity from their copy of the model
  const size_t smem = smem_solve_c(m, d, V, V == 0 ? m.solve_rowcap : d.njmax);
  if (V == 7) { SOLVE_DISPATCH((k_solve<T, 7><<<grid, 32 * SOLVE_WARPS, smem, st>>>(m, d))); }
  else if (V == 3) { SOLVE_DISPATCH((k_solve<T, 3><<<grid, 32 * SOLVE_WARPS, smem, st>>>(m, d))); }
  else if (V == 1) { SOLVE_DISPATCH((k_solve<T, 1><<<grid, 32 * SOLVE_WARPS, smem, st>>>(m, d))); }
  else {
    // the second launch goes FIRST: its few worlds are the most expensive ones of the step (most rows), and behind the main launch's
    // 4096 CTAs they would start half a kernel late (Asimov-toe: 25 such worlds cost 55 us per sub-step that way)
    if (m.solve_rowcap < d.njmax) {     // worlds with more than solve_rowcap rows: same kernel, full layout
      SOLVE_DISPATCH((k_solve<T, 0, true><<<(d.nworld + SOLVE_BIG_SPAN - 1) / SOLVE_BIG_SPAN, 32, smem_solve_c(m, d, 0, d.njmax) / SOLVE_WARPS, side ? side : st>>>(m, d)));
      b2s_count_launch();
    }
    SOLVE_DISPATCH((k_solve<T, 0><<<grid, 32 * SOLVE_WARPS, smem, st>>>(m, d)));
  }
  b2s_count_launch();
}
size_t smem_solve(const ModelDev& m, const DataDev& d) { return smem_solve_c(m, d, solve_variant(m), d.njmax); }
// Raise the dynamic shared-memory limit of EVERY solver instantiation of this model's tile size to the largest request any
// option combination can make (the elliptic variants carry 2 njmax extra floats): Model.opt.cone / solver / ls_parallel and a
// per-world dof_frictionloss can switch the instantiation after put_data.
template <int T>
static int solve_set_attr(int device, size_t smem) {
  int rc = 0;
  rc |= b2s_raise_smem(device, (const void*)k_solve<T, 0>, smem);
  rc |= b2s_raise_smem(device, (const void*)k_solve<T, 0, true>, smem);
  rc |= b2s_raise_smem(device, (const void*)k_solve<T, 1>, smem);
  rc |= b2s_raise_smem(device, (const void*)k_solve<T, 3>, smem);
  rc |= b2s_raise_smem(device, (const void*)k_solve<T, 7>, smem);
  return rc;
}
int solve_prepare_all(const ModelDev& m, const DataDev& d, size_t* need, int device) {
  const size_t smem = (size_t)SOLVE_WARPS * solve_smem_floats(m.nv, m.nvp, m.nbody, d.njmax, d.nconmax, tile_for(m.nv), m.nchain, 1) * sizeof(float);
  *need = smem;
  if (smem > 227 * 1024) return -1;
  int rc = 0;
  SOLVE_DISPATCH(rc = solve_set_attr<T>(device, smem));
  return rc ? -2 : 0;
}
const void* fn_solve(const ModelDev& m) {
  const void* f = nullptr;
  if (!m.ls_parallel || m.solver != B2S_SOL_NEWTON) { SOLVE_DISPATCH((f = (const void*)k_solve<T, 7>)); }
  else if (m.cone == B2S_CONE_ELLIPTIC) { SOLVE_DISPATCH((f = (const void*)k_solve<T, 3>)); }
  else if (m.fric) { SOLVE_DISPATCH((f = (const void*)k_solve<T, 1>)); }
  else { SOLVE_DISPATCH((f = (const void*)k_solve<T, 0>)); }
  return f;
}
