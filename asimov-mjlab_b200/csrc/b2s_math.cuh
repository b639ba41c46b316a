// b2s_math.cuh -- small fp32 device math used by every stage kernel.
#pragma once
#include <cuda_runtime.h>
#include "b2s_types.h"

#define FULL 0xffffffffu
// Device-side bounds / invariant checks of a CHECK build (tools/build_variant.sh check -DB2S_CHECK; compute-sanitizer is closed on
// this pool): a failed check traps the kernel and the next synchronisation reports it.  Compiled out of the shipped library.
#ifdef B2S_CHECK
#include <cassert>
#define B2S_ASSERT(c) assert(c)
#else
#define B2S_ASSERT(c) ((void)0)
#endif
#define DEV __device__ __forceinline__

struct v3 { float x, y, z; };
struct q4 { float w, x, y, z; };
struct m33 { float m[9]; };   // row-major
struct sv6 { float v[6]; };   // spatial vector [angular, linear]

DEV v3 mk3(float x, float y, float z) { v3 r = {x, y, z}; return r; }
DEV v3 ld3(const float* p) { return mk3(p[0], p[1], p[2]); }
DEV void st3(float* p, v3 a) { p[0] = a.x; p[1] = a.y; p[2] = a.z; }
DEV v3 operator+(v3 a, v3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
DEV v3 operator-(v3 a, v3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
DEV v3 operator*(v3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
DEV float dot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
DEV v3 cross(v3 a, v3 b) { return mk3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
DEV float norm(v3 a) { return sqrtf(dot(a, a)); }
DEV v3 normalized(v3 a) {
  float n = norm(a);
  if (n < B2S_MINVAL) return mk3(1.f, 0.f, 0.f);
  float s = 1.0f / n;
  return a * s;
}
DEV q4 ldq(const float* p) { q4 r = {p[0], p[1], p[2], p[3]}; return r; }
DEV void stq(float* p, q4 a) { p[0] = a.w; p[1] = a.x; p[2] = a.y; p[3] = a.z; }
DEV q4 qnormalize(q4 q) {
  float n = sqrtf(q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z);
  if (n < B2S_MINVAL) { q4 r = {1.f, 0.f, 0.f, 0.f}; return r; }
  float s = 1.0f / n;
  q4 r = {q.w * s, q.x * s, q.y * s, q.z * s};
  return r;
}
DEV q4 qmul(q4 a, q4 b) {
  q4 r = {a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z, a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y,
          a.w * b.y - a.x * b.z + a.y * b.w + a.z * b.x, a.w * b.z + a.x * b.y - a.y * b.x + a.z * b.w};
  return r;
}
DEV q4 axis_angle(v3 axis, float angle) {
  float s, c;
  sincosf(0.5f * angle, &s, &c);
  q4 r = {c, axis.x * s, axis.y * s, axis.z * s};
  return r;
}
DEV m33 q2m(q4 q) {
  m33 r;
  float q00 = q.w * q.w, q11 = q.x * q.x, q22 = q.y * q.y, q33 = q.z * q.z;
  float q01 = q.w * q.x, q02 = q.w * q.y, q03 = q.w * q.z, q12 = q.x * q.y, q13 = q.x * q.z, q23 = q.y * q.z;
  r.m[0] = q00 + q11 - q22 - q33; r.m[4] = q00 - q11 + q22 - q33; r.m[8] = q00 - q11 - q22 + q33;
  r.m[1] = 2.f * (q12 - q03); r.m[2] = 2.f * (q13 + q02);
  r.m[3] = 2.f * (q12 + q03); r.m[5] = 2.f * (q23 - q01);
  r.m[6] = 2.f * (q13 - q02); r.m[7] = 2.f * (q23 + q01);
  return r;
}
DEV m33 ldm(const float* p) { m33 r;
#pragma unroll
  for (int k = 0; k < 9; k++) r.m[k] = p[k];
  return r; }
DEV void stm(float* p, const m33& a) {
#pragma unroll
  for (int k = 0; k < 9; k++) p[k] = a.m[k]; }
DEV v3 mulv(const m33& a, v3 v) {
  return mk3(a.m[0] * v.x + a.m[1] * v.y + a.m[2] * v.z, a.m[3] * v.x + a.m[4] * v.y + a.m[5] * v.z,
             a.m[6] * v.x + a.m[7] * v.y + a.m[8] * v.z);
}
DEV v3 mulTv(const m33& a, v3 v) {
  return mk3(a.m[0] * v.x + a.m[3] * v.y + a.m[6] * v.z, a.m[1] * v.x + a.m[4] * v.y + a.m[7] * v.z,
             a.m[2] * v.x + a.m[5] * v.y + a.m[8] * v.z);
}
DEV v3 col(const m33& a, int c) { return mk3(a.m[c], a.m[3 + c], a.m[6 + c]); }
DEV q4 m2q(const m33& R) {
  const float* m = R.m;
  q4 q;
  float tr = m[0] + m[4] + m[8];
  if (tr > 0.f) { float s = sqrtf(tr + 1.f) * 2.f; q.w = 0.25f * s; q.x = (m[7] - m[5]) / s; q.y = (m[2] - m[6]) / s; q.z = (m[3] - m[1]) / s; }
  else if (m[0] > m[4] && m[0] > m[8]) { float s = sqrtf(1.f + m[0] - m[4] - m[8]) * 2.f; q.w = (m[7] - m[5]) / s; q.x = 0.25f * s; q.y = (m[1] + m[3]) / s; q.z = (m[2] + m[6]) / s; }
  else if (m[4] > m[8]) { float s = sqrtf(1.f + m[4] - m[0] - m[8]) * 2.f; q.w = (m[2] - m[6]) / s; q.x = (m[1] + m[3]) / s; q.y = 0.25f * s; q.z = (m[5] + m[7]) / s; }
  else { float s = sqrtf(1.f + m[8] - m[0] - m[4]) * 2.f; q.w = (m[3] - m[1]) / s; q.x = (m[2] + m[6]) / s; q.y = (m[5] + m[7]) / s; q.z = 0.25f * s; }
  return qnormalize(q);
}

// ---- spatial algebra (MuJoCo's com-based convention: [angular, linear] about the root subtree COM)
DEV sv6 ld6(const float* p) { sv6 r;
#pragma unroll
  for (int k = 0; k < 6; k++) r.v[k] = p[k];
  return r; }
DEV void st6(float* p, const sv6& a) {
#pragma unroll
  for (int k = 0; k < 6; k++) p[k] = a.v[k]; }
// the same for 8-byte aligned addresses (shared-memory arrays of 6-vectors whose base is an even float offset): three 8-byte
// accesses instead of six 4-byte ones -- k_dynamics ran at 80 % of the L1 / shared-memory pipe with the scalar form
DEV sv6 ld6s(const float* p) {
  const float2* q = reinterpret_cast<const float2*>(p);
  const float2 a = q[0], b = q[1], c = q[2];
  sv6 r; r.v[0] = a.x; r.v[1] = a.y; r.v[2] = b.x; r.v[3] = b.y; r.v[4] = c.x; r.v[5] = c.y;
  return r; }
DEV void st6s(float* p, const sv6& a) {
  float2* q = reinterpret_cast<float2*>(p);
  q[0] = make_float2(a.v[0], a.v[1]); q[1] = make_float2(a.v[2], a.v[3]); q[2] = make_float2(a.v[4], a.v[5]); }
DEV sv6 zero6() { sv6 r;
#pragma unroll
  for (int k = 0; k < 6; k++) r.v[k] = 0.f;
  return r; }
DEV void axpy6(sv6& y, float a, const sv6& x) {
#pragma unroll
  for (int k = 0; k < 6; k++) y.v[k] += a * x.v[k]; }
DEV float dot6(const sv6& a, const sv6& b) {
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < 6; k++) s += a.v[k] * b.v[k];
  return s; }
// 10-number inertia about the common frame: [Ixx Iyy Izz Ixy Ixz Iyz mx my mz m]
DEV sv6 mul_inert(const float* I, const sv6& x) {
  const float* v = x.v; sv6 r;
  r.v[0] = I[0] * v[0] + I[3] * v[1] + I[4] * v[2] - I[8] * v[4] + I[7] * v[5];
  r.v[1] = I[3] * v[0] + I[1] * v[1] + I[5] * v[2] + I[8] * v[3] - I[6] * v[5];
  r.v[2] = I[4] * v[0] + I[5] * v[1] + I[2] * v[2] - I[7] * v[3] + I[6] * v[4];
  r.v[3] = I[8] * v[1] - I[7] * v[2] + I[9] * v[3];
  r.v[4] = I[6] * v[2] - I[8] * v[0] + I[9] * v[4];
  r.v[5] = I[7] * v[0] - I[6] * v[1] + I[9] * v[5];
  return r;
}
DEV sv6 cross_motion(const sv6& a, const sv6& b) {
  const float *vel = a.v, *v = b.v; sv6 r;
  r.v[0] = -vel[2] * v[1] + vel[1] * v[2];
  r.v[1] = vel[2] * v[0] - vel[0] * v[2];
  r.v[2] = -vel[1] * v[0] + vel[0] * v[1];
  r.v[3] = -vel[2] * v[4] + vel[1] * v[5] - vel[5] * v[1] + vel[4] * v[2];
  r.v[4] = vel[2] * v[3] - vel[0] * v[5] + vel[5] * v[0] - vel[3] * v[2];
  r.v[5] = -vel[1] * v[3] + vel[0] * v[4] - vel[4] * v[0] + vel[3] * v[1];
  return r;
}
DEV sv6 cross_force(const sv6& a, const sv6& b) {
  const float *vel = a.v, *f = b.v; sv6 r;
  r.v[0] = -vel[2] * f[1] + vel[1] * f[2] - vel[5] * f[4] + vel[4] * f[5];
  r.v[1] = vel[2] * f[0] - vel[0] * f[2] + vel[5] * f[3] - vel[3] * f[5];
  r.v[2] = -vel[1] * f[0] + vel[0] * f[1] - vel[4] * f[3] + vel[3] * f[4];
  r.v[3] = -vel[2] * f[4] + vel[1] * f[5];
  r.v[4] = vel[2] * f[3] - vel[0] * f[5];
  r.v[5] = -vel[1] * f[3] + vel[0] * f[4];
  return r;
}

// ---- TMA bulk copy (cp.async.bulk, the non-tensor 1-D form) completed on an mbarrier: one elected lane arms the barrier with
// the byte count and issues the copy, the copy engine writes shared memory asynchronously while the warp goes on with its
// prologue, and every lane waits on the barrier's phase before the first read.  dst / src 16-byte aligned, bytes % 16 == 0.
DEV unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
DEV void mbar_init(unsigned long long* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
DEV void tma_load_1d(void* dst, const void* src, unsigned bytes, unsigned long long* bar) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
DEV void mbar_wait(unsigned long long* bar, unsigned phase) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "B2S_WAIT:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra B2S_DONE;\n"
      "bra B2S_WAIT;\n"
      "B2S_DONE:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(phase) : "memory");
}

// ---- warp helpers
DEV float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
  return v;
}
DEV int warp_excl_scan(int v, int lane, int* total) {
  int x = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(FULL, x, o); if (lane >= o) x += y; }
  *total = __shfl_sync(FULL, x, 31);
  return x - v;
}
DEV float clampf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }

// In-place dense Cholesky A = L L^T on the lower triangle of an n x n matrix held in shared memory with row
// stride `ld` (ld odd => conflict-free column walks).  One warp; right-looking.
DEV void warp_cholesky(float* A, int n, int ld, int lane) {
  for (int k = 0; k < n; k++) {
    __syncwarp();
    float dkk = A[k * ld + k];
    dkk = sqrtf(fmaxf(dkk, B2S_MINVAL));
    float inv = 1.0f / dkk;
    __syncwarp();
    if (lane == 0) A[k * ld + k] = dkk;
    for (int i = k + 1 + lane; i < n; i += 32) A[i * ld + k] *= inv;
    __syncwarp();
    for (int i = k + 1 + lane; i < n; i += 32) {
      float lik = A[i * ld + k];
      for (int j = k + 1; j <= i; j++) A[i * ld + j] -= lik * A[j * ld + k];
    }
  }
  __syncwarp();
}
// Solve L L^T x = b with the factor above.  b is distributed: lane l holds b[l] in b0 and b[l+32] in b1.
DEV void warp_chol_solve(const float* L, int n, int ld, int lane, float& b0, float& b1) {
  for (int k = 0; k < n; k++) {   // forward: L y = b
    float bk = __shfl_sync(FULL, k < 32 ? b0 : b1, k & 31) / L[k * ld + k];
    if (lane == k) b0 = bk;
    if (lane + 32 == k) b1 = bk;
    if (lane > k && lane < n) b0 -= L[lane * ld + k] * bk;
    if (lane + 32 > k && lane + 32 < n) b1 -= L[(lane + 32) * ld + k] * bk;
  }
  for (int k = n - 1; k >= 0; k--) {   // backward: L^T x = y
    float xk = __shfl_sync(FULL, k < 32 ? b0 : b1, k & 31) / L[k * ld + k];
    if (lane == k) b0 = xk;
    if (lane + 32 == k) b1 = xk;
    if (lane < k) b0 -= L[k * ld + lane] * xk;
    if (lane + 32 < k) b1 -= L[k * ld + lane + 32] * xk;
  }
}

// ---------------------------------------------------------------------------------------------------------
// Blocked Cholesky on REGISTER tiles.  The symmetric n x n matrix (n <= 7T) is cut into 7 x 7 blocks of T x T;
// lane l < 28 owns lower-triangular block (bi, bj), bj <= bi, as a[T*T] in registers.  Per block column kb:
// the diagonal owner factors its tile, panel owners solve against it, everybody else applies the rank-T
// update; tiles are exchanged through the shared-memory matrix S (row stride ld), which ends up holding L
// (lower triangle) for the substitution step, with 1/L[k][k] in invd[k].
DEV void tile_owner(int lane, int& bi, int& bj) {
  bi = -1; bj = 0;
  if (lane < 28) { bi = 0; int l = lane; while (l > bi) { l -= bi + 1; bi++; } bj = l; }
}
template <int T>
DEV void tile_cholesky_regs(float (&a)[T * T], int bi, int bj, float* S, int ld, float* invd) {
#pragma unroll 1
  for (int kb = 0; kb < 7; kb++) {
    if (bi == kb && bj == kb) {
      float* out = S + (kb * T) * ld + kb * T;
#pragma unroll
      for (int y = 0; y < T; y++) {
        float sdiag = a[y * T + y];
#pragma unroll
        for (int z = 0; z < y; z++) sdiag -= a[y * T + z] * a[y * T + z];
        // 1/sqrt by the SFU approximation + one Newton step (full fp32 accuracy), sqrt = s * rsqrt(s): this runs on ONE
        // lane, 7T times per factorisation, on the critical path; IEEE sqrt + divide were 10 % of the solver's instructions
        const float sd = fmaxf(sdiag, B2S_MINVAL);
        float inv = rsqrtf(sd);
        inv = inv * (1.5f - 0.5f * sd * inv * inv);
        const float dd = sd * inv;
        a[y * T + y] = dd;
        out[y * ld + y] = dd;
        invd[kb * T + y] = inv;
#pragma unroll
        for (int x = y + 1; x < T; x++) {
          float t = a[x * T + y];
#pragma unroll
          for (int z = 0; z < y; z++) t -= a[x * T + z] * a[y * T + z];
          t *= inv;
          a[x * T + y] = t;
          out[x * ld + y] = t;
        }
      }
    }
    __syncwarp();
    if (bj == kb && bi > kb) {
      const float* lkk = S + (kb * T) * ld + kb * T;
      float* out = S + (bi * T) * ld + kb * T;
#pragma unroll
      for (int y = 0; y < T; y++) {
        const float inv = invd[kb * T + y];
        float l[T];
#pragma unroll
        for (int z = 0; z < y; z++) l[z] = lkk[y * ld + z];
#pragma unroll
        for (int x = 0; x < T; x++) {
          float t = a[x * T + y];
#pragma unroll
          for (int z = 0; z < y; z++) t -= a[x * T + z] * l[z];
          t *= inv;
          a[x * T + y] = t;
          out[x * ld + y] = t;
        }
      }
    }
    __syncwarp();
    if (bj > kb && bi >= bj) {
      const float* li = S + (bi * T) * ld + kb * T;
      const float* lj = S + (bj * T) * ld + kb * T;
#pragma unroll
      for (int z = 0; z < T; z++) {
        float ci[T], cj[T];
#pragma unroll
        for (int x = 0; x < T; x++) { ci[x] = li[x * ld + z]; cj[x] = lj[x * ld + z]; }
#pragma unroll
        for (int x = 0; x < T; x++)
#pragma unroll
          for (int y = 0; y < T; y++) a[x * T + y] -= ci[x] * cj[y];
      }
    }
  }
  __syncwarp();
}
// Factor the lower triangle of the n x n matrix stored in S (row stride ld >= 7T) in place.
template <int T>
DEV void tile_cholesky_smem(float* S, int n, int ld, int lane, float* invd) {
  int bi, bj;
  tile_owner(lane, bi, bj);
  float a[T * T];
  if (bi >= 0) {
#pragma unroll
    for (int x = 0; x < T; x++)
#pragma unroll
      for (int y = 0; y < T; y++) {
        const int i = bi * T + x, j = bj * T + y;
        a[x * T + y] = (i < n && j < n) ? (j <= i ? S[i * ld + j] : S[j * ld + i]) : (i == j ? 1.0f : 0.0f);
      }
  }
  __syncwarp();
  tile_cholesky_regs<T>(a, bi, bj, S, ld, invd);
}
// Substitution with the factor above: lane l holds b[l] in b0 and b[l+32] in b1 (n <= 64).
// REQUIRES the strict upper triangle of L's storage to be zero (rows < hdim): the running residual
// t_l = b_l - sum_j L[l][j] y_j is then updated by every lane without predicates.
DEV void warp_chol_solve_inv(const float* L, const float* invd, int n, int ld, int lane, float& b0, float& b1) {
  const bool hi = n > 32;
  const int row0 = min(lane, n - 1), row1 = min(lane + 32, n - 1);   // clamp: lanes without a row re-read row n-1 (result unused)
  const int col0 = min(lane, ld - 1), col1 = min(lane + 32, ld - 1);
  float y0 = 0.f, y1 = 0.f;
  for (int k = 0; k < n; k++) {                // forward: L y = b
    const float yk = __shfl_sync(FULL, k < 32 ? b0 : b1, k & 31) * invd[k];
    if (lane == (k & 31)) { if (k < 32) y0 = yk; else y1 = yk; }
    b0 -= L[row0 * ld + k] * yk;
    if (hi) b1 -= L[row1 * ld + k] * yk;
  }
  // backward: L^T x = y
  if (lane >= n) y0 = 0.f;
  if (lane + 32 >= n) y1 = 0.f;
  for (int k = n - 1; k >= 0; k--) {
    const float xk = __shfl_sync(FULL, k < 32 ? y0 : y1, k & 31) * invd[k];
    if (lane == (k & 31)) { if (k < 32) b0 = xk; else b1 = xk; }
    y0 -= L[k * ld + col0] * xk;
    if (hi) y1 -= L[k * ld + col1] * xk;
  }
}
// Same substitution with the PADDED dimension N = 7T as a compile-time constant: fully unrolled, shared-memory
// offsets become immediates.  Rows/cols n..N-1 of L are the identity padding written by the tile Cholesky and
// the matching entries of b must be 0 on entry.  REQUIRES a zero strict upper triangle in L's storage.
// Partially unrolled (8 or 7 steps), not fully: the fully unrolled form is ~900 SASS instructions per call site and the
// solver's loop body is already larger than the 32 KB instruction cache.  Rows 0..31 (held in b0) and rows 32..N-1
// (held in b1 on lanes 0..N-33) are separate loops so that no step has to select between the two.
template <int T>
DEV void tile_chol_solve(const float* L, const float* invd, int ld, int lane, float& b0, float& b1) {
  constexpr int N = 7 * T;
  constexpr bool HI = N > 32;
  constexpr int N0 = HI ? 32 : N, U = HI ? 8 : 7;
  const float* Lr0 = L + min(lane, N - 1) * ld;
  const float* Lr1 = L + min(lane + 32, N - 1) * ld;
  const float* Lc0 = L + min(lane, N - 1);
  const float* Lc1 = L + min(lane + 32, N - 1);
  float y0 = 0.f, y1 = 0.f;
#pragma unroll U
  for (int k = 0; k < N0; k++) {                 // forward: L y = b
    const float yk = __shfl_sync(FULL, b0, k) * invd[k];
    if (lane == k) y0 = yk;
    b0 -= Lr0[k] * yk;
    if (HI) b1 -= Lr1[k] * yk;
  }
  if (HI) {
#pragma unroll
    for (int k = 32; k < N; k++) {
      const float yk = __shfl_sync(FULL, b1, k - 32) * invd[k];
      if (lane == k - 32) y1 = yk;
      b1 -= Lr1[k] * yk;                         // rows < 32 have zeros in columns >= 32
    }
#pragma unroll
    for (int k = N - 1; k >= 32; k--) {          // backward: L^T x = y
      const float xk = __shfl_sync(FULL, y1, k - 32) * invd[k];
      if (lane == k - 32) b1 = xk;
      y0 -= Lc0[k * ld] * xk;
      y1 -= Lc1[k * ld] * xk;
    }
  }
#pragma unroll U
  for (int k = N0 - 1; k >= 0; k--) {
    const float xk = __shfl_sync(FULL, y0, k) * invd[k];
    if (lane == k) b0 = xk;
    y0 -= Lc0[k * ld] * xk;                      // column entries above the diagonal are zero: rows >= 32 need none
  }
}

// ---------------------------------------------------------------------------------------------------------
// Row-owner Cholesky (round 2; replaces the 28-lane tile version in every stage kernel).  The symmetric matrix A (N = 7T
// padded rows, n <= N real ones, identity padding) is factored as A = U U^T with U UPPER triangular, i.e. eliminating the LAST
// dof first (MuJoCo's own order, which is also the order in which a kinematic tree has no fill-in).  Lane i owns row i of the
// upper triangle in REGISTERS (a[j] = A_ij, j >= i; lanes 0 .. N-33 also own row i + 32 in b[]), the loop over pivots is fully
// unrolled so every register index is static.  Step k: the pivot comes from its owner by shuffle, every lane scales its own
// column entry l_i = a[k] / U_kk, the column goes through a double-buffered shared-memory line (one contiguous store, then
// broadcast LDS.128 reads), and each lane applies a[j] -= l_i l_j for j < k.  About 17 instructions of overhead per pivot and
// k FMAs: ~1.2 k warp instructions for N = 35 (the tile version: ~2.5 k, at 28/32 lanes and 7 x 3 phases).
// Output: U row-major in S (entry (i, k), k >= i, at S[i * ld + k]; the strict lower triangle of S is NOT cleared and the
// substitution below does not need it to be), 1 / U_kk in invd[k].  colbuf: 2 x RC_LINE floats, 16-byte aligned.
#define RC_LINE 64
template <int T>
struct RowChol {
  static constexpr int N = 7 * T;
  static constexpr bool HI = N > 32;
  static constexpr int NB = HI ? N - 32 : 1;      // second rows (32 .. N-1), one per lane 0 .. NB-1
};
// rows of the symmetric matrix whose LOWER triangle is stored in S (row stride ld; entry (i, j), j >= i, is read at S[j * ld + i],
// which is contiguous across lanes); rows / columns >= n are identity padding
// PK (packed): S holds only a triangle, entry (max, min) at S[max (max + 1) / 2 + min] -- N (N + 1) / 2 floats instead of N x ld.
// M's lower triangle and the factor U (entry (i, k), i <= k, at k (k + 1) / 2 + i) then share the same slots; the reads of "junk"
// positions the algorithms rely on being harmless stay in bounds (they land on other entries of the triangle).
template <int T, bool PK = false>
DEV void rowchol_load_lower(const float* S, int ld, int n, int lane, float (&a)[7 * T], float (&b)[RowChol<T>::NB]) {
  constexpr int N = RowChol<T>::N;
  const int i0 = min(lane, N - 1);
#pragma unroll
  for (int j = 0; j < N; j++) a[j] = (lane < n && j < n) ? S[PK ? j * (j + 1) / 2 + i0 : j * ld + i0] : (j == lane ? 1.0f : 0.0f);
  if (RowChol<T>::HI) {
    const int i1 = min(lane + 32, N - 1);
#pragma unroll
    for (int j = 32; j < N; j++) b[j - 32] = (lane + 32 < n && j < n) ? S[PK ? j * (j + 1) / 2 + i1 : j * ld + i1] : (j == lane + 32 ? 1.0f : 0.0f);
  }
}
// steps k = k_hi-1 .. 0 of the elimination (k_hi = N: the whole factorisation)
template <int T, bool PK = false>
DEV void rowchol_factor(float (&a)[7 * T], float (&b)[RowChol<T>::NB], int lane, float* S, int ld, float* invd, float* colbuf, int k_hi = 7 * T) {
  constexpr int N = RowChol<T>::N;
  constexpr bool HI = RowChol<T>::HI;
  float* const Ur0 = S + (PK ? min(lane, N - 1) : min(lane, N - 1) * ld);      // entry (row, k) at Ur[PK ? k (k + 1) / 2 : k]
  float* const Ur1 = S + (PK ? min(lane + 32, N - 1) : min(lane + 32, N - 1) * ld);
#pragma unroll
  for (int k = N - 1; k >= 0; k--) {
    if (k < k_hi) {
      float* const line = colbuf + (k & 1) * RC_LINE;
      // pivot: 1/sqrt by the SFU approximation + one Newton step (full fp32 accuracy), sqrt = s * rsqrt(s)
      const float p = __shfl_sync(FULL, k < 32 ? a[k] : b[k >= 32 ? k - 32 : 0], k & 31);
      const float sd = fmaxf(p, B2S_MINVAL);
      float inv = rsqrtf(sd);
      inv = inv * (1.5f - 0.5f * sd * inv * inv);
      float la = a[k] * inv;
      if (k < 32 && lane == k) la = sd * inv;
      a[k] = la;
      line[lane] = la;
      if (lane <= k && lane < N) Ur0[PK ? k * (k + 1) / 2 : k] = la;
      if (HI && k >= 32) {
        float lb = b[k >= 32 ? k - 32 : 0] * inv;
        if (lane == k - 32) lb = sd * inv;
        b[k >= 32 ? k - 32 : 0] = lb;
        if (lane < RowChol<T>::NB) { line[32 + lane] = lb; if (lane + 32 <= k) Ur1[PK ? k * (k + 1) / 2 : k] = lb; }
      }
      if (lane == 0) invd[k] = inv;
      __syncwarp();
      const float4* l4 = reinterpret_cast<const float4*>(line);
#pragma unroll
      for (int q = 0; q < (k + 3) / 4; q++) {
        const float4 v = l4[q];
        if (4 * q + 0 < k) a[4 * q + 0] -= la * v.x;
        if (4 * q + 1 < k) a[4 * q + 1] -= la * v.y;
        if (4 * q + 2 < k) a[4 * q + 2] -= la * v.z;
        if (4 * q + 3 < k) a[4 * q + 3] -= la * v.w;
      }
      if (HI && k > 32) {
        const float lb = b[k >= 32 ? k - 32 : 0];
#pragma unroll
        for (int j = 32; j < k; j++) b[j - 32] -= lb * line[j];
      }
    }
  }
  __syncwarp();
}
// Solve U U^T x = g with the factor above: lane l holds g[l] in b0 and g[l + 32] in b1 (entries of padding rows must be 0).
// U y = g runs from the last row up, U^T x = y from the first down; a finished lane keeps its result in a second register, so
// the junk in S's strict lower triangle (and the clamped rows of lanes >= N) only ever lands in values that are not read again.
template <int T, bool PK = false>
DEV void rowchol_solve(const float* S, const float* invd, int ld, int lane, float& b0, float& b1) {
  constexpr int N = RowChol<T>::N;
  constexpr bool HI = RowChol<T>::HI;
  constexpr int N0 = HI ? 32 : N, UN = HI ? 8 : 7;
  const int r0 = min(lane, N - 1), r1 = min(lane + 32, N - 1);
  const float* Ur0 = S + (PK ? r0 : r0 * ld);                    // entry (row, k) at Ur[PK ? k (k + 1) / 2 : k]
  const float* Ur1 = S + (PK ? r1 : r1 * ld);
  const float* Uc0 = S + (PK ? r0 * (r0 + 1) / 2 : r0);          // entry (k, col) at Uc[PK ? k : k ld]
  const float* Uc1 = S + (PK ? r1 * (r1 + 1) / 2 : r1);
#define RC_ROW(k) (PK ? (k) * ((k) + 1) / 2 : (k))
#define RC_COL(k) (PK ? (k) : (k) * ld)
  float y0 = 0.f, y1 = 0.f;
  if (HI) {
#pragma unroll
    for (int k = N - 1; k >= 32; k--) {
      const float yk = __shfl_sync(FULL, b1, k - 32) * invd[k];
      if (lane == k - 32) y1 = yk;
      b0 -= Ur0[RC_ROW(k)] * yk;
      b1 -= Ur1[RC_ROW(k)] * yk;
    }
  }
#pragma unroll UN
  for (int k = N0 - 1; k >= 0; k--) {
    const float yk = __shfl_sync(FULL, b0, k) * invd[k];
    if (lane == k) y0 = yk;
    b0 -= Ur0[RC_ROW(k)] * yk;
  }
#pragma unroll UN
  for (int k = 0; k < N0; k++) {
    const float xk = __shfl_sync(FULL, y0, k) * invd[k];
    if (lane == k) b0 = xk;
    y0 -= Uc0[RC_COL(k)] * xk;
    if (HI) y1 -= Uc1[RC_COL(k)] * xk;
  }
  if (HI) {
#pragma unroll
    for (int k = 32; k < N; k++) {
      const float xk = __shfl_sync(FULL, y1, k - 32) * invd[k];
      if (lane == k - 32) b1 = xk;
      y1 -= Uc1[RC_COL(k)] * xk;
    }
  }
}
#undef RC_ROW
#undef RC_COL
// factor the symmetric matrix whose lower triangle is in S (n real rows) in place: S ends up holding U (see above)
template <int T, bool PK = false>
DEV void rowchol_smem(float* S, int n, int ld, int lane, float* invd, float* colbuf) {
  float a[7 * T], b[RowChol<T>::NB];
  rowchol_load_lower<T, PK>(S, ld, n, lane, a, b);
  __syncwarp();
  rowchol_factor<T, PK>(a, b, lane, S, ld, invd, colbuf);
}
__host__ __device__ inline int b2s_hdim(int T) { return 7 * T; }                         // padded matrix dimension
__host__ __device__ inline int b2s_hld(int T) { return (7 * T + 1) | 1; }                 // odd row stride > 7T
__host__ __device__ inline int b2s_tile_for(int nv) { int T = (nv + 6) / 7; if (T < 1) T = 1; if (T == 7) T = 8; return T; }
