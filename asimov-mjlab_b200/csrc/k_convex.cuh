// k_convex.cuh -- convex / height-field narrowphase, one lane per geom pair.
//
// GJK distance + EPA penetration between convex CORES (point, segment, box, vertex cloud); spheres and capsules are a
// core plus a radius, so their results are exact for the rounded shape.  plane-mesh scans the hull vertices and keeps a
// 4-point manifold; a height field is visited as triangular prisms over the sub-grid under the other geom's bounding
// sphere.  Needed by the fork's Asimov robots (mesh collision geoms, asimov_toe.xml:12,49) and by height-field terrain
// (terrains/heightfield_terrains.py:213-242).  Everything is per-thread state: the simplex lives in registers, the EPA
// polytope (28 vertices, 56 faces) in local memory; none of it is instantiated for models without mesh / hfield pairs.
#pragma once
#include "b2s_math.cuh"

#define CVX_POINT 0
#define CVX_SEGMENT 1
#define CVX_BOX 2
#define CVX_VERTS 3
#define CVX_CYLINDER 4      // size.x radius, size.y half height along local z
#define CVX_ELLIPSOID 5     // size = the three radii
#define GJK_MAXIT 48
#define EPA_MAXIT 48
#define EPA_MAXV 28
#define EPA_MAXF 56
#define CVX_TOL 1e-6f

struct Cvx {
  int type;
  v3 pos;
  m33 mat;
  v3 size;             // SEGMENT: size.y = half length along local z; BOX: half sizes
  const float* vert;   // VERTS: local frame, 4 floats per vertex (mesh hull; world == false), or world frame, 3 floats per vertex (prism)
  int nvert;
  bool world;
  bool coop;           // true: all 32 lanes are in this call with identical arguments and share the vertex scan
  float radius;
};
struct CvxRes { float dist; v3 n, pa, pb;
#ifdef COL_TIMING
  int dbg_gjk, dbg_epa; long long dbg_epa_cyc;      // development builds: GJK iterations, EPA rounds and cycles of this pair
#endif
};
struct Simplex { v3 Y[4], A[4]; float lam[4]; int n; };

// support point of the core in world direction dir (ties: lowest vertex index)
__device__ __noinline__ v3 cvx_support(const Cvx& s, v3 dir) {
  if (s.type == CVX_POINT) return s.pos;
  if (s.type == CVX_VERTS && s.world) {
    int best = 0; float bd = dot(ld3(s.vert), dir);
    for (int i = 1; i < s.nvert; i++) { const float t = dot(ld3(s.vert + 3 * i), dir); if (t > bd) { bd = t; best = i; } }
    return ld3(s.vert + 3 * best);
  }
  const v3 dl = mulTv(s.mat, dir);
  v3 loc;
  if (s.type == CVX_SEGMENT) loc = mk3(0.f, 0.f, dl.z >= 0.f ? s.size.y : -s.size.y);
  else if (s.type == CVX_BOX) loc = mk3(dl.x >= 0.f ? s.size.x : -s.size.x, dl.y >= 0.f ? s.size.y : -s.size.y, dl.z >= 0.f ? s.size.z : -s.size.z);
  else if (s.type == CVX_CYLINDER) {       // rim point of the cap facing dir
    const float h = sqrtf(dl.x * dl.x + dl.y * dl.y);
    loc = mk3(h > B2S_MINVAL ? s.size.x * dl.x / h : 0.f, h > B2S_MINVAL ? s.size.x * dl.y / h : 0.f, dl.z >= 0.f ? s.size.y : -s.size.y);
  } else if (s.type == CVX_ELLIPSOID) {    // S^2 d / |S d|
    const v3 e = mk3(s.size.x * dl.x, s.size.y * dl.y, s.size.z * dl.z);
    const float h = sqrtf(dot(e, e));
    loc = h > B2S_MINVAL ? mk3(s.size.x * e.x / h, s.size.y * e.y / h, s.size.z * e.z / h) : mk3(s.size.x, 0.f, 0.f);
  } else if (s.coop) {
    // warp-cooperative scan: lane-strided vertices, then (max value, lowest index) across the warp
    const int lane = threadIdx.x & 31;
    int best = 0x7fffffff; float bd = -3.0e38f;
    const float4* v4 = reinterpret_cast<const float4*>(s.vert);        // mesh hulls are stored 4 floats per vertex
    // four independent 16-byte loads in flight per lane before the first compare (the scan is latency-bound: L1 / L2 hits);
    // candidates are still visited in ascending index order per lane, so ties keep the lowest index
#pragma unroll 1
    for (int i0 = lane; i0 < s.nvert; i0 += 128) {
      float4 p[4];
#pragma unroll
      for (int u = 0; u < 4; u++) p[u] = __ldg(v4 + min(i0 + 32 * u, s.nvert - 1));
#pragma unroll
      for (int u = 0; u < 4; u++) {
        const int i = i0 + 32 * u;
        const float t = p[u].x * dl.x + p[u].y * dl.y + p[u].z * dl.z;
        if (i < s.nvert && t > bd) { bd = t; best = i; }
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ob = __shfl_xor_sync(0xffffffffu, bd, o);
      const int oi = __shfl_xor_sync(0xffffffffu, best, o);
      if (ob > bd || (ob == bd && oi < best)) { bd = ob; best = oi; }
    }
    loc = ld3(s.vert + 4 * best);
  } else {
    const float4* v4 = reinterpret_cast<const float4*>(s.vert);
    int best = 0; float bd = -3.0e38f;
    for (int i = 0; i < s.nvert; i++) {
      const float4 p = __ldg(v4 + i);
      const float t = p.x * dl.x + p.y * dl.y + p.z * dl.z;
      if (t > bd) { bd = t; best = i; }
    }
    loc = ld3(s.vert + 4 * best);
  }
  return mulv(s.mat, loc) + s.pos;
}
DEV void mink_support(const Cvx& A, const Cvx& B, v3 dir, v3& y, v3& a) {
  a = cvx_support(A, dir);
  const v3 b = cvx_support(B, dir * -1.f);
  y = a - b;
}

// ---- closest point to the origin on a simplex (Voronoi regions); the simplex is reduced in place
DEV void sx_keep1(Simplex& s, int i) { s.Y[0] = s.Y[i]; s.A[0] = s.A[i]; s.n = 1; s.lam[0] = 1.f; }
DEV void sx_keep2(Simplex& s, int i, int j, float t) {
  const v3 y0 = s.Y[i], y1 = s.Y[j], a0 = s.A[i], a1 = s.A[j];
  s.Y[0] = y0; s.Y[1] = y1; s.A[0] = a0; s.A[1] = a1; s.n = 2; s.lam[0] = 1.f - t; s.lam[1] = t;
}
DEV void closest_segment(Simplex& s) {
  const v3 ab = s.Y[1] - s.Y[0];
  float t = -dot(s.Y[0], ab);
  const float den = dot(ab, ab);
  if (t <= 0.f || den <= 0.f) { sx_keep1(s, 0); return; }
  if (t >= den) { sx_keep1(s, 1); return; }
  t /= den; s.lam[0] = 1.f - t; s.lam[1] = t;
}
DEV void closest_triangle(Simplex& s) {
  const v3 a = s.Y[0], b = s.Y[1], c = s.Y[2];
  const v3 ab = b - a, ac = c - a;
  const float d1 = -dot(ab, a), d2 = -dot(ac, a);
  if (d1 <= 0.f && d2 <= 0.f) { sx_keep1(s, 0); return; }
  const float d3 = -dot(ab, b), d4 = -dot(ac, b);
  if (d3 >= 0.f && d4 <= d3) { sx_keep1(s, 1); return; }
  const float vc = d1 * d4 - d3 * d2;
  if (vc <= 0.f && d1 >= 0.f && d3 <= 0.f) { sx_keep2(s, 0, 1, d1 / (d1 - d3)); return; }
  const float d5 = -dot(ab, c), d6 = -dot(ac, c);
  if (d6 >= 0.f && d5 <= d6) { sx_keep1(s, 2); return; }
  const float vb = d5 * d2 - d1 * d6;
  if (vb <= 0.f && d2 >= 0.f && d6 <= 0.f) { sx_keep2(s, 0, 2, d2 / (d2 - d6)); return; }
  const float va = d3 * d6 - d5 * d4;
  if (va <= 0.f && (d4 - d3) >= 0.f && (d5 - d6) >= 0.f) { sx_keep2(s, 1, 2, (d4 - d3) / ((d4 - d3) + (d5 - d6))); return; }
  const float den = va + vb + vc;
  if (den <= 0.f) { s.n = 2; closest_segment(s); return; }      // collinear
  s.lam[1] = vb / den; s.lam[2] = vc / den; s.lam[0] = 1.f - s.lam[1] - s.lam[2];
}
DEV v3 sx_point(const v3* P, const Simplex& s) {
  v3 r = mk3(0.f, 0.f, 0.f);
  for (int i = 0; i < s.n; i++) r = r + P[i] * s.lam[i];
  return r;
}
// true: the origin is inside the tetrahedron (simplex untouched); else reduced to the closest face feature
__device__ __noinline__ bool closest_tetrahedron(Simplex& s, float scale2) {
  float bestd = -1.f;
  Simplex best;
  best.n = 0;
#pragma unroll 1
  for (int f = 0; f < 4; f++) {
    const int i0 = f == 3 ? 1 : 0, i1 = f == 0 ? 1 : (f == 1 ? 2 : 3), i2 = f == 0 ? 2 : (f == 1 ? 3 : (f == 2 ? 1 : 2)), i3 = f == 0 ? 3 : (f == 1 ? 1 : (f == 2 ? 2 : 0));
    const v3 a = s.Y[i0], b = s.Y[i1], c = s.Y[i2], dd = s.Y[i3];
    const v3 nn = cross(b - a, c - a);
    const float sp = -dot(a, nn), sd = dot(dd - a, nn);
    // outer side of this face, or a (nearly) flat tetrahedron: resolved through its faces, never certified as enclosing
    if (sp * sd < 0.f || sd * sd <= 1e-10f * dot(nn, nn) * scale2) {
      Simplex t;
      t.Y[0] = a; t.Y[1] = b; t.Y[2] = c; t.A[0] = s.A[i0]; t.A[1] = s.A[i1]; t.A[2] = s.A[i2]; t.n = 3;
      closest_triangle(t);
      const v3 v = sx_point(t.Y, t);
      const float d2 = dot(v, v);
      if (bestd < 0.f || d2 < bestd) { bestd = d2; best = t; }
    }
  }
  if (bestd < 0.f) return true;
  s = best;
  return false;
}

// The same for warp-level callers (all 32 lanes here with the same simplex): face f is evaluated by the lanes with lane % 4 == f,
// the winner -- smallest distance, lowest face index on ties, exactly the serial loop's choice -- is broadcast.  A GJK run that
// cycles to its iteration cap spends most of its time here (measured: 5 us per iteration, 250 us per capped pair).
__device__ __noinline__ bool closest_tetrahedron_warp(Simplex& s, float scale2) {
  const int f = threadIdx.x & 3;
  const int i0 = f == 3 ? 1 : 0, i1 = f == 0 ? 1 : (f == 1 ? 2 : 3), i2 = f == 0 ? 2 : (f == 1 ? 3 : (f == 2 ? 1 : 2)), i3 = f == 0 ? 3 : (f == 1 ? 1 : (f == 2 ? 2 : 0));
  const v3 a = s.Y[i0], b = s.Y[i1], c = s.Y[i2], dd = s.Y[i3];
  const v3 nn = cross(b - a, c - a);
  const float sp = -dot(a, nn), sd = dot(dd - a, nn);
  const bool cand = sp * sd < 0.f || sd * sd <= 1e-10f * dot(nn, nn) * scale2;
  Simplex t;
  t.n = 0;
  float d2 = 0.f;
  if (cand) {
    t.Y[0] = a; t.Y[1] = b; t.Y[2] = c; t.A[0] = s.A[i0]; t.A[1] = s.A[i1]; t.A[2] = s.A[i2]; t.n = 3;
    closest_triangle(t);
    const v3 v = sx_point(t.Y, t);
    d2 = dot(v, v);
  }
  __syncwarp();
  int bestf = -1; float bestd = -1.f;
#pragma unroll
  for (int g = 0; g < 4; g++) {
    const float dg = __shfl_sync(0xffffffffu, d2, g);
    const int cg = __shfl_sync(0xffffffffu, cand ? 1 : 0, g);
    if (cg && (bestd < 0.f || dg < bestd)) { bestd = dg; bestf = g; }
  }
  if (bestf < 0) return true;
  s.n = __shfl_sync(0xffffffffu, t.n, bestf);
#pragma unroll
  for (int i = 0; i < 3; i++) {
    s.Y[i].x = __shfl_sync(0xffffffffu, t.Y[i].x, bestf); s.Y[i].y = __shfl_sync(0xffffffffu, t.Y[i].y, bestf); s.Y[i].z = __shfl_sync(0xffffffffu, t.Y[i].z, bestf);
    s.A[i].x = __shfl_sync(0xffffffffu, t.A[i].x, bestf); s.A[i].y = __shfl_sync(0xffffffffu, t.A[i].y, bestf); s.A[i].z = __shfl_sync(0xffffffffu, t.A[i].z, bestf);
    s.lam[i] = __shfl_sync(0xffffffffu, t.lam[i], bestf);
  }
  return false;
}

// ---- EPA: triangle faces with outward normals; the face closest to the origin is expanded along its normal
struct EpaFace { unsigned char v[3], alive; v3 n; float d; };
DEV void epa_make_face(EpaFace& f, const v3* V, int a, int b, int c) {
  f.v[0] = (unsigned char)a; f.v[1] = (unsigned char)b; f.v[2] = (unsigned char)c; f.alive = 1;
  v3 n = cross(V[b] - V[a], V[c] - V[a]);
  const float len = norm(n);
  if (len < 1e-20f) { f.n = mk3(0.f, 0.f, 0.f); f.d = 1e30f; return; }      // sliver: never chosen
  f.n = mk3(n.x / len, n.y / len, n.z / len);
  f.d = dot(f.n, V[a]);
}
// polytope storage of one EPA run.  Per lane in local memory by default; warp-level callers (all 32 lanes run the SAME polytope
// redundantly) pass one copy in shared memory instead: every lane stores the same value to the same address and reads it back
// itself, so no synchronisation is needed and the 32 x 2.7 KB of local-memory traffic per warp disappear.
struct EpaWork { v3 V[EPA_MAXV], VA[EPA_MAXV]; EpaFace F[EPA_MAXF]; unsigned char E[EPA_MAXF * 3][2]; };
__device__ __noinline__ bool epa_penetration(const Cvx& A, const Cvx& B, const Simplex& s, CvxRes& out, EpaWork* wk = nullptr) {
  EpaWork local_wk;
  if (!wk) wk = &local_wk;
  v3* const V = wk->V; v3* const VA = wk->VA;
  EpaFace* const F = wk->F;
  unsigned char (*const E)[2] = wk->E;
  int nv = 4, nf = 0;
  for (int i = 0; i < 4; i++) { V[i] = s.Y[i]; VA[i] = s.A[i]; }
  if (dot(cross(V[1] - V[0], V[2] - V[0]), V[3] - V[0]) > 0.f) {      // make (0,1,2) face away from vertex 3
    v3 t = V[1]; V[1] = V[2]; V[2] = t; t = VA[1]; VA[1] = VA[2]; VA[2] = t;
  }
  epa_make_face(F[nf++], V, 0, 1, 2);
  epa_make_face(F[nf++], V, 0, 3, 1);
  epa_make_face(F[nf++], V, 0, 2, 3);
  epa_make_face(F[nf++], V, 1, 3, 2);
  int best = 0;
#pragma unroll 1
  for (int it = 0; it < EPA_MAXIT; it++) {
#ifdef COL_TIMING
    out.dbg_epa = it + 1;
#endif
    best = -1;
    for (int f = 0; f < nf; f++) if (F[f].alive && (best < 0 || F[f].d < F[best].d)) best = f;
    if (best < 0) return false;
    v3 w, wa;
    mink_support(A, B, F[best].n, w, wa);
    const float gain = dot(w, F[best].n) - F[best].d;
    const float tol = CVX_TOL * (1.f + fabsf(F[best].d));
    if (gain <= tol || nv >= EPA_MAXV) break;
    int ne = 0, nvis = 0;
    for (int f = 0; f < nf; f++) {
      if (!F[f].alive) continue;
      if (dot(F[f].n, w - V[F[f].v[0]]) > 0.f || F[f].d >= 1e30f) {
        F[f].alive = 0; nvis++;
        for (int k = 0; k < 3; k++) { E[ne][0] = F[f].v[k]; E[ne][1] = F[f].v[(k + 1) % 3]; ne++; }
      }
    }
    if (nvis == 0) break;
    V[nv] = w; VA[nv] = wa;
    int added = 0;
    bool full = false;
    for (int e = 0; e < ne; e++) {
      bool twin = false;
      for (int g = 0; g < ne; g++) if (E[g][0] == E[e][1] && E[g][1] == E[e][0]) { twin = true; break; }
      if (twin) continue;
      int slot = -1;
      for (int f = 0; f < nf; f++) if (!F[f].alive) { slot = f; break; }
      if (slot < 0) { if (nf >= EPA_MAXF) { full = true; break; } slot = nf++; }
      epa_make_face(F[slot], V, E[e][0], E[e][1], nv);
      added++;
    }
    nv++;
    if (full || added == 0) break;
  }
  best = -1;
  for (int f = 0; f < nf; f++) if (F[f].alive && (best < 0 || F[f].d < F[best].d)) best = f;
  if (best < 0) return false;
  Simplex t;
  for (int i = 0; i < 3; i++) { t.Y[i] = V[F[best].v[i]]; t.A[i] = VA[F[best].v[i]]; }
  t.n = 3;
  closest_triangle(t);
  const v3 p = sx_point(t.Y, t);
  out.pa = sx_point(t.A, t);
  out.n = F[best].n;
  out.pb = out.pa - p;
  out.dist = -fmaxf(F[best].d, 0.f);
  return true;
}

// The same algorithm for warp-level callers: all 32 lanes are here with identical arguments and ONE polytope in shared memory, so the
// per-face and per-edge loops are split over the lanes instead of being run 32 times over -- closest face = warp arg-min of (d,
// index); visibility = one lane per face, the removed faces' edges appended in face order through a ballot; horizon = one lane
// per edge looking for its twin; new faces = the j-th horizon edge (ascending) takes the j-th dead slot (ascending), then fresh
// slots, which is what the serial loop does one edge at a time.  Same polytope after every round, hence the same result.
// (Measured with -DCOL_TIMING before: 15 us per round, one 25-round pair = 388 us = most of k_convex_tasks' duration.)
__device__ __noinline__ bool epa_penetration_warp(const Cvx& A, const Cvx& B, const Simplex& s, CvxRes& out, EpaWork* wk) {
  const int lane = threadIdx.x & 31;
  const unsigned lt = (1u << lane) - 1u;
  v3* const V = wk->V; v3* const VA = wk->VA;
  EpaFace* const F = wk->F;
  unsigned char (*const E)[2] = wk->E;
  int nv = 4, nf = 0;
  __syncwarp();
  if (lane == 0) {
    for (int i = 0; i < 4; i++) { V[i] = s.Y[i]; VA[i] = s.A[i]; }
    if (dot(cross(V[1] - V[0], V[2] - V[0]), V[3] - V[0]) > 0.f) {      // make (0,1,2) face away from vertex 3
      v3 t = V[1]; V[1] = V[2]; V[2] = t; t = VA[1]; VA[1] = VA[2]; VA[2] = t;
    }
    epa_make_face(F[0], V, 0, 1, 2);
    epa_make_face(F[1], V, 0, 3, 1);
    epa_make_face(F[2], V, 0, 2, 3);
    epa_make_face(F[3], V, 1, 3, 2);
  }
  nf = 4;
  __syncwarp();
  auto closest_face = [&]() {
    float bd = 0.f; int best = -1;
    for (int f = lane; f < nf; f += 32) if (F[f].alive && (best < 0 || F[f].d < bd)) { bd = F[f].d; best = f; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ob = __shfl_xor_sync(0xffffffffu, bd, o);
      const int of = __shfl_xor_sync(0xffffffffu, best, o);
      if (of >= 0 && (best < 0 || ob < bd || (ob == bd && of < best))) { bd = ob; best = of; }
    }
    return best;
  };
  int best = 0;
#pragma unroll 1
  for (int it = 0; it < EPA_MAXIT; it++) {
#ifdef COL_TIMING
    out.dbg_epa = it + 1;
#endif
    best = closest_face();
    if (best < 0) return false;
    const v3 bn = F[best].n; const float bdist = F[best].d;
    v3 w, wa;
    mink_support(A, B, bn, w, wa);
    const float gain = dot(w, bn) - bdist;
    const float tol = CVX_TOL * (1.f + fabsf(bdist));
    if (gain <= tol || nv >= EPA_MAXV) break;
    int ne = 0, nvis = 0;
    for (int f0 = 0; f0 < nf; f0 += 32) {
      const int f = f0 + lane;
      bool vis = false;
      if (f < nf && F[f].alive) vis = dot(F[f].n, w - V[F[f].v[0]]) > 0.f || F[f].d >= 1e30f;
      const unsigned vm = __ballot_sync(0xffffffffu, vis);
      if (vis) {
        F[f].alive = 0;
        const int r = ne + 3 * __popc(vm & lt);
        for (int k = 0; k < 3; k++) { E[r + k][0] = F[f].v[k]; E[r + k][1] = F[f].v[(k + 1) % 3]; }
      }
      ne += 3 * __popc(vm); nvis += __popc(vm);
    }
    if (nvis == 0) break;
    if (lane == 0) { V[nv] = w; VA[nv] = wa; }
    __syncwarp();
    // dead slots in ascending order (EPA_MAXF <= 64: two ballots)
    const unsigned dead0 = __ballot_sync(0xffffffffu, lane < nf && !F[lane].alive);
    const unsigned dead1 = __ballot_sync(0xffffffffu, lane + 32 < nf && !F[lane + 32].alive);
    const int nd0 = __popc(dead0), ndead = nd0 + __popc(dead1);
    int nedge = 0;        // horizon edges so far
    for (int e0 = 0; e0 < ne; e0 += 32) {
      const int e = e0 + lane;
      bool boundary = false;
      if (e < ne) {
        boundary = true;
        const unsigned char a = E[e][0], b = E[e][1];
        for (int g = 0; g < ne; g++) if (E[g][0] == b && E[g][1] == a) { boundary = false; break; }
      }
      const unsigned bm = __ballot_sync(0xffffffffu, boundary);
      if (boundary) {
        const int j = nedge + __popc(bm & lt);
        int slot;
        if (j < nd0) slot = (int)__fns(dead0, 0, j + 1);
        else if (j < ndead) slot = 32 + (int)__fns(dead1, 0, j - nd0 + 1);
        else slot = nf + (j - ndead);
        if (slot < EPA_MAXF) epa_make_face(F[slot], V, E[e][0], E[e][1], nv);
      }
      nedge += __popc(bm);
    }
    // the serial loop stops at the first edge that finds no slot: faces exist for the edges before it
    const int fresh = max(nedge - ndead, 0);
    const bool full = nf + fresh > EPA_MAXF;
    const int added = full ? min(nedge, ndead + (EPA_MAXF - nf)) : nedge;
    nf = full ? EPA_MAXF : nf + fresh;
    nv++;
    __syncwarp();
    if (full || added == 0) break;
  }
  best = closest_face();
  if (best < 0) return false;
  Simplex t;
  for (int i = 0; i < 3; i++) { t.Y[i] = V[F[best].v[i]]; t.A[i] = VA[F[best].v[i]]; }
  t.n = 3;
  closest_triangle(t);
  const v3 p = sx_point(t.Y, t);
  out.pa = sx_point(t.A, t);
  out.n = F[best].n;
  out.pb = out.pa - p;
  out.dist = -fmaxf(F[best].d, 0.f);
  return true;
}

// GJK on the cores.  false: farther apart than `cutoff`.  Separated: dist > 0, n from A to B, pa / pb closest core
// points.  Overlapping cores: the simplex is grown to a tetrahedron around the origin and handed to EPA.
// `axis` (optional, 3 floats of per-pair storage that lives across steps; warp-level callers only): TEMPORAL COHERENCE.  A pair that
// was separated in the last step is almost always separated by nearly the same plane now, so the direction GJK ended with is
// tried first: one support evaluation, and if that plane still separates the cores by more than `cutoff` there is no contact --
// the same bound (valid for ANY direction) that ends the GJK loop early, so the outcome is the one the full search would reach.
// Otherwise the search runs exactly as without the cache (same start direction, same iterates).  The cache is refreshed
// whenever the search proves separation and cleared when it finds a contact.
// `work` (optional): GJK iterations of this call, + 64 if EPA ran (k_convex_tasks keeps it per pair to queue expensive pairs first).
__device__ __noinline__ bool gjk_epa(const Cvx& A, const Cvx& B, float cutoff, CvxRes& out, float* axis = nullptr, EpaWork* wk = nullptr, int* work = nullptr) {
  Simplex s;
  if (axis) {
    const v3 vc = ld3(axis);
    const float vv = dot(vc, vc);
    if (vv > 0.f) {
      v3 w, wa;
      mink_support(A, B, vc * -1.f, w, wa);
      const float vw = dot(vc, w);
      if (vw > 0.f && vw * vw > cutoff * cutoff * vv) { if (work) *work = 0; return false; }
    }
    __syncwarp();
  }
  v3 dir = B.pos - A.pos;
  if (dot(dir, dir) < 1e-20f) dir = mk3(1.f, 0.f, 0.f);
  mink_support(A, B, dir, s.Y[0], s.A[0]);
  s.n = 1; s.lam[0] = 1.f;
  v3 v = s.Y[0];
  float scale2 = dot(v, v);
  bool inside = false;
  int it = 0;
#ifdef COL_TIMING
  out.dbg_gjk = 0; out.dbg_epa = 0; out.dbg_epa_cyc = 0;
#endif
#pragma unroll 1
  for (; it < GJK_MAXIT; it++) {
    const float vv = dot(v, v);
    if (vv <= CVX_TOL * CVX_TOL * scale2 || vv < 1e-30f) { inside = true; break; }
    v3 w, wa;
    mink_support(A, B, v * -1.f, w, wa);
    const float vw = dot(v, w);
    if (vw > 0.f && vw * vw > cutoff * cutoff * vv) {
#ifdef COL_TIMING
      out.dbg_gjk = it + 1;
#endif
      if (work) *work = it + 1;
      if (axis && (threadIdx.x & 31) == 0) st3(axis, v); return false; }
    const float ww = dot(w, w);
    if (ww > scale2) scale2 = ww;
    if (vv - vw <= CVX_TOL * vv + 1e-12f * scale2) break;
    bool dup = false;
    for (int i = 0; i < s.n; i++) { const v3 e = w - s.Y[i]; if (dot(e, e) <= 1e-14f * scale2) dup = true; }
    if (dup) break;
    s.Y[s.n] = w; s.A[s.n] = wa; s.n++;
    if (s.n == 2) closest_segment(s);
    else if (s.n == 3) closest_triangle(s);
    else if (wk ? closest_tetrahedron_warp(s, scale2) : closest_tetrahedron(s, scale2)) { inside = true; break; }
    v = sx_point(s.Y, s);
  }
#ifdef COL_TIMING
  out.dbg_gjk = it;
#endif
  if (work) *work = it + (inside ? 64 : 0);
  if (axis && (threadIdx.x & 31) == 0) {      // separated beyond the cutoff: remember the direction; anything else: no cached plane
    const bool sep = !inside && dot(v, v) > cutoff * cutoff;
    st3(axis, sep ? v : mk3(0.f, 0.f, 0.f));
  }
  if (!inside) {
    const float vv = dot(v, v), dist = sqrtf(vv);
    if (dist > cutoff) return false;
    out.pa = sx_point(s.A, s);
    out.pb = out.pa - v;
    out.n = mk3(-v.x / dist, -v.y / dist, -v.z / dist);
    out.dist = dist;
    return true;
  }
  if (s.n < 4) {
    if (s.n == 1) {
      for (int a = 0; a < 6 && s.n == 1; a++) {
        const v3 ax = mk3(a == 0 ? 1.f : (a == 1 ? -1.f : 0.f), a == 2 ? 1.f : (a == 3 ? -1.f : 0.f), a == 4 ? 1.f : (a == 5 ? -1.f : 0.f));
        v3 w, wa; mink_support(A, B, ax, w, wa);
        const v3 e = w - s.Y[0];
        if (dot(e, e) > 1e-12f * scale2 + 1e-30f) { s.Y[1] = w; s.A[1] = wa; s.n = 2; }
      }
    }
    if (s.n == 2) {
      const v3 ab = s.Y[1] - s.Y[0];
      for (int a = 0; a < 6 && s.n == 2; a++) {
        const v3 ax = mk3(a == 0 ? 1.f : (a == 1 ? -1.f : 0.f), a == 2 ? 1.f : (a == 3 ? -1.f : 0.f), a == 4 ? 1.f : (a == 5 ? -1.f : 0.f));
        const v3 dd = cross(ab, ax);
        if (dot(dd, dd) < 1e-12f * dot(ab, ab)) continue;
        v3 w, wa; mink_support(A, B, dd, w, wa);
        const v3 cr = cross(ab, w - s.Y[0]);
        if (dot(cr, cr) > 1e-12f * scale2 * dot(ab, ab)) { s.Y[2] = w; s.A[2] = wa; s.n = 3; }
      }
    }
    if (s.n == 3) {
      const v3 nn = cross(s.Y[1] - s.Y[0], s.Y[2] - s.Y[0]);
      for (int sgn = 0; sgn < 2 && s.n == 3; sgn++) {
        v3 w, wa; mink_support(A, B, sgn ? nn * -1.f : nn, w, wa);
        const float vol = dot(w - s.Y[0], nn);
        if (vol * vol > 1e-12f * dot(nn, nn) * scale2) { s.Y[3] = w; s.A[3] = wa; s.n = 4; }
      }
    }
    if (s.n < 4) {       // flat Minkowski difference: zero-depth contact
      v3 nn = B.pos - A.pos;
      const float l = norm(nn);
      nn = l < 1e-12f ? mk3(0.f, 0.f, 1.f) : mk3(nn.x / l, nn.y / l, nn.z / l);
      out.pa = s.A[0]; out.pb = s.A[0] - s.Y[0]; out.n = nn; out.dist = 0.f;
      return true;
    }
  }
#ifdef COL_TIMING
  const long long t_epa = clock64();
  const bool okk = wk ? epa_penetration_warp(A, B, s, out, wk) : epa_penetration(A, B, s, out, wk);
  out.dbg_epa_cyc = clock64() - t_epa;
  return okk;
#else
  return wk ? epa_penetration_warp(A, B, s, out, wk) : epa_penetration(A, B, s, out, wk);
#endif
}

// one contact between two convex geoms (cores + radii); normal from A to B.  Rc is the kernel's contact scratch type.
template <class Rc>
DEV int convex_pair(Rc* c, float margin, const Cvx& A, const Cvx& B) {
  CvxRes r;
  if (!gjk_epa(A, B, margin + A.radius + B.radius, r)) return 0;
  const float dist = r.dist - A.radius - B.radius;
  if (dist > margin) return 0;
  const v3 sa = r.pa + r.n * A.radius, sb = r.pb - r.n * B.radius;
  c->dist = dist; c->pos = (sa + sb) * 0.5f; c->n = r.n; c->t = mk3(0.f, 0.f, 0.f);
  return 1;
}

// ---- warp-level drivers.  All 32 lanes call these with identical arguments; results land in `out` (shared memory).
struct CvxStage { float dist; v3 pos, n; int idx; };

// (max value, lowest index) over the warp of a per-lane candidate
DEV void warp_argmax(float& v, int& i) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, v, o);
    const int oi = __shfl_xor_sync(0xffffffffu, i, o);
    if (ov > v || (ov == v && oi < i)) { v = ov; i = oi; }
  }
}

// plane - mesh: hull vertices within 1 mm of the deepest one are candidates; up to 4 span the manifold
// (A deepest, B farthest from A, C farthest from line AB, D farthest from line BC or AC among the rest).
// Four lane-strided scans of the hull with warp arg-max (ties: lowest vertex index, as a sequential scan would pick).
DEV int plane_mesh_warp(CvxStage* out, float margin, v3 ppos, const m33& pmat, v3 mpos, const m33& mmat, const float* vert, int nvert, int lane) {
  const v3 n = col(pmat, 2), nl = mulTv(mmat, n), pl = mulTv(mmat, ppos - mpos);
  const float pd = dot(pl, nl);
  float smax = -1e30f; int ia = 0x7fffffff;
  {
    const float4* v4 = reinterpret_cast<const float4*>(vert);       // 4 loads in flight per lane, ascending index order per lane
#pragma unroll 1
    for (int i0 = lane; i0 < nvert; i0 += 128) {
      float4 p[4];
#pragma unroll
      for (int u = 0; u < 4; u++) p[u] = __ldg(v4 + min(i0 + 32 * u, nvert - 1));
#pragma unroll
      for (int u = 0; u < 4; u++) {
        const int i = i0 + 32 * u;
        const float sup = pd - (p[u].x * nl.x + p[u].y * nl.y + p[u].z * nl.z);
        if (i < nvert && sup > smax) { smax = sup; ia = i; }
      }
    }
  }
  warp_argmax(smax, ia);
  if (ia == 0x7fffffff || -smax > margin) return 0;
  float thr = smax - 1e-3f;
  if (thr < 0.f && smax >= 0.f) thr = 0.f;
  if (-thr > margin) thr = -margin;
  int idx[4] = {ia, -1, -1, -1};
  const v3 a = ld3(vert + 4 * ia);
  float best = 0.f; int ib = 0x7fffffff;
  for (int i = lane; i < nvert; i += 32) {
    const v3 v = ld3(vert + 4 * i);
    if (pd - dot(v, nl) < thr) continue;
    const v3 e = v - a; const float q = dot(e, e);
    if (q > best) { best = q; ib = i; }
  }
  warp_argmax(best, ib);
  if (ib != 0x7fffffff && best > 1e-10f) {
    idx[1] = ib;
    const v3 b = ld3(vert + 4 * ib), ab = b - a, side = cross(nl, ab);
    best = 0.f; int ic = 0x7fffffff;
    for (int i = lane; i < nvert; i += 32) {
      const v3 v = ld3(vert + 4 * i);
      if (pd - dot(v, nl) < thr) continue;
      const float q = fabsf(dot(v - a, side));
      if (q > best) { best = q; ic = i; }
    }
    warp_argmax(best, ic);
    if (ic != 0x7fffffff && best > 1e-7f) {
      idx[2] = ic;
      const v3 cc = ld3(vert + 4 * ic), sac = cross(nl, cc - a), sbc = cross(nl, cc - b);
      float bestb = 0.f, besta = 0.f; int jb = 0x7fffffff, ja = 0x7fffffff;
      for (int i = lane; i < nvert; i += 32) {
        const v3 v = ld3(vert + 4 * i);
        if (pd - dot(v, nl) < thr) continue;
        if (i == idx[0] || i == idx[1] || i == idx[2]) continue;
        const float qb = fabsf(dot(v - b, sbc)), qa = fabsf(dot(v - a, sac));
        if (qb > bestb) { bestb = qb; jb = i; }
        if (qa > besta) { besta = qa; ja = i; }
      }
      warp_argmax(bestb, jb); warp_argmax(besta, ja);
      float bq; int id;
      if (bestb >= besta) { bq = bestb; id = jb; } else { bq = besta; id = ja; }
      if (id != 0x7fffffff && bq > 1e-7f) idx[3] = id;
    }
  }
  int nc = 0;
#pragma unroll
  for (int k = 0; k < 4; k++) {
    const int i = idx[k];
    if (i < 0) continue;
    bool dup = false;
    for (int q = 0; q < k; q++) if (idx[q] == i) dup = true;
    if (dup) continue;
    const v3 v = ld3(vert + 4 * i);
    const float sup = pd - dot(v, nl);
    if (-sup > margin) continue;
    if (lane == 0) { out[nc].dist = -sup; out[nc].pos = mulv(mmat, v) + mpos + n * (sup * 0.5f); out[nc].n = n; out[nc].idx = nc; }
    nc++;
  }
  return nc;
}

// one convex pair, every lane in the call: the hull scans are shared (Cvx::coop), the simplex / polytope logic runs
// redundantly and identically on all lanes
DEV int convex_pair_warp(CvxStage* out, float margin, const Cvx& A, const Cvx& B, int lane, float* axis = nullptr, EpaWork* wk = nullptr, int* dbg = nullptr, int* work = nullptr) {
  CvxRes r;
#ifdef COL_TIMING
  r.dbg_gjk = 0; r.dbg_epa = 0; r.dbg_epa_cyc = 0;
  const bool hit = gjk_epa(A, B, margin + A.radius + B.radius, r, axis, wk, work);
  if (dbg && lane == 0) { dbg[0] = r.dbg_gjk | (r.dbg_epa << 8) | (min(A.nvert + B.nvert, 4095) << 16); dbg[1] = (int)min(r.dbg_epa_cyc >> 6, 65535LL) << 16; }
  if (!hit) return 0;
#else
  if (!gjk_epa(A, B, margin + A.radius + B.radius, r, axis, wk, work)) return 0;
#endif
  const float dist = r.dist - A.radius - B.radius;
  if (dist > margin) return 0;
  const v3 sa = r.pa + r.n * A.radius, sb = r.pb - r.n * B.radius;
  if (lane == 0) { out[0].dist = dist; out[0].pos = (sa + sb) * 0.5f; out[0].n = r.n; out[0].idx = 0; }
  return 1;
}

// height field (geom 1) against convex B.  B's exact bounding box in the field's frame (six support calls) selects the
// cells; their triangular prisms, numbered t = (row * ncol + col) * 2 + (upper triangle), are handed out one per lane,
// 32 per round, and each lane runs GJK/EPA on its prism unless the prism's top lies below B's lowest point.  One
// contact per prism; the MAXC deepest are kept, ordered by t (ties: lower t stays).  `out` holds MAXC + 32 entries.
// hs = (x half, y half, z top, base).
template <int MAXC>
DEV int hfield_convex_warp(CvxStage* out, float margin, const float* hs, int nrow, int ncol, const float* data, v3 hpos, const m33& hmat,
                           const Cvx& B, int lane) {
  const float sx = hs[0], sy = hs[1], sz = hs[2], sb = hs[3];
  float lo[3], hi[3];
  {
    Cvx Bc = B;
    Bc.coop = (B.type == CVX_VERTS);        // all lanes are here with the same B: share the hull scans
#pragma unroll
    for (int a = 0; a < 3; a++) {
      const v3 ax = col(hmat, a);
      hi[a] = dot(cvx_support(Bc, ax) - hpos, ax) + B.radius + margin;
      lo[a] = dot(cvx_support(Bc, ax * -1.f) - hpos, ax) - B.radius - margin;
    }
  }
  if (hi[0] < -sx || lo[0] > sx || hi[1] < -sy || lo[1] > sy || lo[2] > sz || hi[2] < -sb) return 0;
  const float dx = 2.f * sx / (float)(ncol - 1), dy = 2.f * sy / (float)(nrow - 1);
  const int cmin = max((int)floorf((lo[0] + sx) / dx), 0), cmax = min((int)ceilf((hi[0] + sx) / dx), ncol - 1);
  const int rmin = max((int)floorf((lo[1] + sy) / dy), 0), rmax = min((int)ceilf((hi[1] + sy) / dy), nrow - 1);
  const int ncell = cmax - cmin, T = (rmax - rmin) * ncell * 2;
  int nk = 0;
#pragma unroll 1
  for (int t0 = 0; t0 < T; t0 += 32) {
    const int t = t0 + lane;
    bool has = false;
    CvxStage one;
    if (t < T) {
      const int tri = t & 1, cell = t >> 1, rr = rmin + cell / ncell, cc = cmin + cell % ncell;
      const int r0 = tri ? rr + 1 : rr, k0 = cc, r1 = rr + 1, k1 = tri ? cc + 1 : cc, r2 = rr, k2 = cc + 1;
      float wv[18];
      float zmax = -sb;
      const int cr[3] = {r0, r1, r2}, ck[3] = {k0, k1, k2};
#pragma unroll
      for (int i = 0; i < 3; i++) {
        const float z = data[cr[i] * ncol + ck[i]] * sz;
        const v3 top = mk3((float)ck[i] * dx - sx, (float)cr[i] * dy - sy, z), bot = mk3(top.x, top.y, -sb);
        st3(wv + 3 * i, mulv(hmat, top) + hpos);
        st3(wv + 9 + 3 * i, mulv(hmat, bot) + hpos);
        zmax = fmaxf(zmax, z);
      }
      if (zmax >= lo[2]) {
        Cvx P;
        P.type = CVX_VERTS; P.vert = wv; P.nvert = 6; P.world = true; P.coop = false; P.radius = 0.f;
        P.pos = mk3((wv[0] + wv[3] + wv[6] + wv[9] + wv[12] + wv[15]) / 6.f, (wv[1] + wv[4] + wv[7] + wv[10] + wv[13] + wv[16]) / 6.f,
                    (wv[2] + wv[5] + wv[8] + wv[11] + wv[14] + wv[17]) / 6.f);
        CvxRes res;
        if (gjk_epa(P, B, margin + B.radius, res)) {
          const float dist = res.dist - B.radius;
          if (dist <= margin) {
            has = true;
            const v3 sbp = res.pb - res.n * B.radius;
            one.dist = dist; one.pos = (res.pa + sbp) * 0.5f; one.n = res.n; one.idx = (rr * ncol + cc) * 2 + tri;
          }
        }
      }
    }
    const unsigned mask = __ballot_sync(0xffffffffu, has);
    if (has) out[nk + __popc(mask & ((1u << lane) - 1u))] = one;
    const int ntot = nk + __popc(mask);
    __syncwarp();
    if (ntot > MAXC) {
      // keep the MAXC smallest by (dist, t); entries are already in t order, so compaction preserves it
      CvxStage e0, e1;
      bool k0 = false, k1 = false;
      if (lane < ntot) {
        e0 = out[lane]; int rank = 0;
        for (int j = 0; j < ntot; j++) rank += (out[j].dist < e0.dist) || (out[j].dist == e0.dist && out[j].idx < e0.idx);
        k0 = rank < MAXC;
      }
      if (lane + 32 < ntot) {
        e1 = out[lane + 32]; int rank = 0;
        for (int j = 0; j < ntot; j++) rank += (out[j].dist < e1.dist) || (out[j].dist == e1.dist && out[j].idx < e1.idx);
        k1 = rank < MAXC;
      }
      const unsigned m0 = __ballot_sync(0xffffffffu, k0), m1 = __ballot_sync(0xffffffffu, k1);
      __syncwarp();
      if (k0) out[__popc(m0 & ((1u << lane) - 1u))] = e0;
      if (k1) out[__popc(m0) + __popc(m1 & ((1u << lane) - 1u))] = e1;
      __syncwarp();
      nk = MAXC;
    } else nk = ntot;
  }
  return nk;
}

// ---- flattened height-field tasks (used by k_collision): per-lane set-up and one prism task
#define CVX_STAGE_CAP 96
#define HF_LIST_CAP 128      // live prism tasks of one pass (32 survivors) that k_collision compacts before the GJK rounds
struct HfRange { int cmin, ncell, rmin, T; float loz; };

// B's bounding box in the field's frame (six support calls, this lane only) -> cell range and prism count
DEV HfRange hfield_range(float margin, const float* hs, int nrow, int ncol, v3 hpos, const m33& hmat, const Cvx& B) {
  HfRange R; R.cmin = 0; R.ncell = 0; R.rmin = 0; R.T = 0; R.loz = 0.f;
  const float sx = hs[0], sy = hs[1], sz = hs[2], sb = hs[3];
  float lo[3], hi[3];
#pragma unroll
  for (int a = 0; a < 3; a++) {
    const v3 ax = col(hmat, a);
    hi[a] = dot(cvx_support(B, ax) - hpos, ax) + B.radius + margin;
    lo[a] = dot(cvx_support(B, ax * -1.f) - hpos, ax) - B.radius - margin;
  }
  if (hi[0] < -sx || lo[0] > sx || hi[1] < -sy || lo[1] > sy || lo[2] > sz || hi[2] < -sb) return R;
  const float dx = 2.f * sx / (float)(ncol - 1), dy = 2.f * sy / (float)(nrow - 1);
  const int cmin = max((int)floorf((lo[0] + sx) / dx), 0), cmax = min((int)ceilf((hi[0] + sx) / dx), ncol - 1);
  const int rmin = max((int)floorf((lo[1] + sy) / dy), 0), rmax = min((int)ceilf((hi[1] + sy) / dy), nrow - 1);
  R.cmin = cmin; R.ncell = max(cmax - cmin, 0); R.rmin = rmin; R.T = max(rmax - rmin, 0) * R.ncell * 2; R.loz = lo[2];
  return R;
}
// cheap part of a prism task: false when the prism's top lies below B's lowest point (hfield_prism_task would return false
// before running GJK); k_collision compacts the live prisms of a pass before the GJK rounds
DEV bool hfield_prism_live(int t, const HfRange& R, const float* hs, int ncol, const float* data) {
  const float sz = hs[2], sb = hs[3];
  const int tri = t & 1, cell = t >> 1, rr = R.rmin + cell / R.ncell, cc = R.cmin + cell % R.ncell;
  const int cr[3] = {tri ? rr + 1 : rr, rr + 1, rr}, ck[3] = {cc, tri ? cc + 1 : cc, cc + 1};
  float zmax = -sb;
#pragma unroll
  for (int i = 0; i < 3; i++) zmax = fmaxf(zmax, data[cr[i] * ncol + ck[i]] * sz);
  return !(zmax < R.loz);
}
// prism number t (local to the range) against B; true + `one` when they touch within the margin
DEV bool hfield_prism_task(CvxStage& one, int t, const HfRange& R, float margin, const float* hs, int nrow, int ncol, const float* data,
                           v3 hpos, const m33& hmat, const Cvx& B) {
  const float sx = hs[0], sy = hs[1], sz = hs[2], sb = hs[3];
  const float dx = 2.f * sx / (float)(ncol - 1), dy = 2.f * sy / (float)(nrow - 1);
  const int tri = t & 1, cell = t >> 1, rr = R.rmin + cell / R.ncell, cc = R.cmin + cell % R.ncell;
  const int cr[3] = {tri ? rr + 1 : rr, rr + 1, rr}, ck[3] = {cc, tri ? cc + 1 : cc, cc + 1};
  float wv[18];
  float zmax = -sb;
#pragma unroll
  for (int i = 0; i < 3; i++) {
    const float z = data[cr[i] * ncol + ck[i]] * sz;
    const v3 top = mk3((float)ck[i] * dx - sx, (float)cr[i] * dy - sy, z), bot = mk3(top.x, top.y, -sb);
    st3(wv + 3 * i, mulv(hmat, top) + hpos);
    st3(wv + 9 + 3 * i, mulv(hmat, bot) + hpos);
    zmax = fmaxf(zmax, z);
  }
  if (zmax < R.loz) return false;
  Cvx P;
  P.type = CVX_VERTS; P.vert = wv; P.nvert = 6; P.world = true; P.coop = false; P.radius = 0.f;
  P.pos = mk3((wv[0] + wv[3] + wv[6] + wv[9] + wv[12] + wv[15]) / 6.f, (wv[1] + wv[4] + wv[7] + wv[10] + wv[13] + wv[16]) / 6.f,
              (wv[2] + wv[5] + wv[8] + wv[11] + wv[14] + wv[17]) / 6.f);
  CvxRes res;
  if (!gjk_epa(P, B, margin + B.radius, res)) return false;
  const float dist = res.dist - B.radius;
  if (dist > margin) return false;
  const v3 sbp = res.pb - res.n * B.radius;
  one.dist = dist; one.pos = (res.pa + sbp) * 0.5f; one.n = res.n; one.idx = (rr * ncol + cc) * 2 + tri;
  return true;
}
