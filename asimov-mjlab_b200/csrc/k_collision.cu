// k_collision.cu -- broadphase + primitive narrowphase, one warp per world.
//
// The candidate list is the static geom-pair table built at put_model (contype/conaffinity,
// same-weld-body, parent-child and <exclude> filters applied once on the host).  Per world a warp
//   A) tests every pair's bounding spheres (planes: signed distance), 32 pairs per pass, and compacts
//      the survivors IN PAIR ORDER with ballot/popc into shared memory;
//   B) runs the exact primitive test for 32 survivors at a time and appends the contacts with a warp
//      exclusive scan, again in pair order.
// Contacts of a world are therefore stored in a fixed per-world slice, in a deterministic order
// (pair index, then contact index) with no atomics -- unlike the reference backend's global pool
// with an atomic counter (src/mjlab/utils/random.py:12-14 notes its non-determinism).
#include "b2s_math.cuh"
#include "b2s_api.h"
#include "k_convex.cuh"

extern __shared__ float b2s_smem[];

struct RawCon { float dist; v3 pos, n, t; };

DEV void make_frame(v3 n, v3 t, float* f) {
  n = normalized(n);
  if (norm(t) < 0.5f) t = (n.y < 0.5f && n.y > -0.5f) ? mk3(0.f, 1.f, 0.f) : mk3(0.f, 0.f, 1.f);
  t = normalized(t - n * dot(n, t));
  v3 b = cross(n, t);
  st3(f, n); st3(f + 3, t); st3(f + 6, b);
}
DEV int plane_sphere(RawCon* c, float margin, v3 ppos, v3 n, v3 spos, float r) {
  float cdist = dot(spos - ppos, n);
  if (cdist > margin + r) return 0;
  c->dist = cdist - r;
  c->pos = spos - n * (c->dist * 0.5f + r);
  c->n = n; c->t = mk3(0.f, 0.f, 0.f);
  return 1;
}
DEV int sphere_sphere(RawCon* c, float margin, v3 p1, float r1, v3 p2, float r2) {
  v3 dif = p2 - p1;
  float cd = norm(dif);
  if (cd > margin + r1 + r2) return 0;
  c->dist = cd - r1 - r2;
  c->n = (cd < B2S_MINVAL) ? mk3(1.f, 0.f, 0.f) : dif * (1.0f / cd);
  c->pos = p1 + c->n * (r1 + c->dist * 0.5f);
  c->t = mk3(0.f, 0.f, 0.f);
  return 1;
}
DEV int plane_capsule(RawCon* c, float margin, v3 p1, const m33& m1, v3 p2, const m33& m2, v3 s2) {
  v3 n = col(m1, 2), axis = col(m2, 2), seg = axis * s2.y;
  int cnt = plane_sphere(c, margin, p1, n, p2 + seg, s2.x);
  cnt += plane_sphere(c + cnt, margin, p1, n, p2 - seg, s2.x);
  v3 y = axis - n * dot(axis, n);
  y = (norm(y) >= 0.5f) ? normalized(y) : mk3(0.f, 0.f, 0.f);
  for (int i = 0; i < cnt; i++) c[i].t = y;
  return cnt;
}
DEV int plane_box(RawCon* c, float margin, v3 p1, const m33& m1, v3 p2, const m33& m2, v3 s2) {
  v3 n = col(m1, 2);
  float dist = dot(p2 - p1, n);
  int cnt = 0;
  for (int i = 0; i < 8; i++) {
    v3 v = mk3((i & 1) ? s2.x : -s2.x, (i & 2) ? s2.y : -s2.y, (i & 4) ? s2.z : -s2.z);
    v3 corner = mulv(m2, v);
    float ldist = dist + dot(n, corner);
    if (ldist > margin) continue;
    c[cnt].dist = ldist; c[cnt].pos = p2 + corner - n * (ldist * 0.5f); c[cnt].n = n; c[cnt].t = mk3(0.f, 0.f, 0.f);
    if (++cnt >= 4) break;
  }
  return cnt;
}
DEV int sphere_capsule(RawCon* c, float margin, v3 p1, v3 s1, v3 p2, const m33& m2, v3 s2) {
  v3 axis = col(m2, 2);
  float x = clampf(dot(axis, p1 - p2), -s2.y, s2.y);
  return sphere_sphere(c, margin, p1, s1.x, p2 + axis * x, s2.x);
}
DEV int capsule_capsule(RawCon* c, float margin, v3 p1, const m33& m1, v3 s1, v3 p2, const m33& m2, v3 s2) {
  v3 a1 = col(m1, 2), a2 = col(m2, 2), dif = p1 - p2;
  float ma = dot(a1, a1), mb = -dot(a1, a2), mc = dot(a2, a2), u = -dot(a1, dif), v = dot(a2, dif);
  float det = ma * mc - mb * mb;
  if (fabsf(det) >= B2S_MINVAL) {
    float x1 = (mc * u - mb * v) / det, x2 = (ma * v - mb * u) / det;
    if (x1 > s1.y) { x1 = s1.y; x2 = (v - mb * s1.y) / mc; }
    else if (x1 < -s1.y) { x1 = -s1.y; x2 = (v + mb * s1.y) / mc; }
    if (x2 > s2.y) { x2 = s2.y; x1 = clampf((u - mb * s2.y) / ma, -s1.y, s1.y); }
    else if (x2 < -s2.y) { x2 = -s2.y; x1 = clampf((u + mb * s2.y) / ma, -s1.y, s1.y); }
    return sphere_sphere(c, margin, p1 + a1 * x1, s1.x, p2 + a2 * x2, s2.x);
  }
  int n = 0;
  float x2 = clampf((v - mb * s1.y) / mc, -s2.y, s2.y);
  n += sphere_sphere(c + n, margin, p1 + a1 * s1.y, s1.x, p2 + a2 * x2, s2.x);
  x2 = clampf((v + mb * s1.y) / mc, -s2.y, s2.y);
  n += sphere_sphere(c + n, margin, p1 - a1 * s1.y, s1.x, p2 + a2 * x2, s2.x);
  if (n >= 2) return n;
  float x1 = clampf((u - mb * s2.y) / ma, -s1.y, s1.y);
  n += sphere_sphere(c + n, margin, p1 + a1 * x1, s1.x, p2 + a2 * s2.y, s2.x);
  if (n >= 2) return n;
  x1 = clampf((u + mb * s2.y) / ma, -s1.y, s1.y);
  n += sphere_sphere(c + n, margin, p1 + a1 * x1, s1.x, p2 - a2 * s2.y, s2.x);
  return n;
}
DEV int sphere_box(RawCon* c, float margin, v3 p1, v3 s1, v3 p2, const m33& m2, v3 s2) {
  v3 loc = mulTv(m2, p1 - p2);
  v3 cl = mk3(clampf(loc.x, -s2.x, s2.x), clampf(loc.y, -s2.y, s2.y), clampf(loc.z, -s2.z, s2.z));
  bool inside = (cl.x == loc.x) && (cl.y == loc.y) && (cl.z == loc.z);
  v3 nl; float dist; const float r = s1.x;
  if (!inside) {
    v3 dv = loc - cl; float dn = norm(dv);
    if (dn - r > margin) return 0;
    dist = dn - r; nl = dv * (-1.0f / dn);
  } else {
    float bx = s2.x - fabsf(loc.x), by = s2.y - fabsf(loc.y), bz = s2.z - fabsf(loc.z);
    nl = mk3(0.f, 0.f, 0.f); float best = bx; int kb = 0;
    if (by < best) { best = by; kb = 1; }
    if (bz < best) { best = bz; kb = 2; }
    if (kb == 0) nl.x = loc.x >= 0.f ? -1.f : 1.f; else if (kb == 1) nl.y = loc.y >= 0.f ? -1.f : 1.f; else nl.z = loc.z >= 0.f ? -1.f : 1.f;
    dist = -best - r;
  }
  c->n = mulv(m2, nl); c->dist = dist; c->pos = p1 + c->n * (r + dist * 0.5f); c->t = mk3(0.f, 0.f, 0.f);
  return 1;
}

// capsule-box: each contact is the sphere-box contact of one point of the capsule segment -- the point closest to
// the box (or the deepest sampled point if the segment enters it), then a segment end whose own contact faces the
// same way (oracle: col_capsule_box).
DEV float seg_box_closest(v3 a, v3 dd, v3 h, bool* inside_out) {
  const float av[3] = {a.x, a.y, a.z}, dv[3] = {dd.x, dd.y, dd.z}, hv[3] = {h.x, h.y, h.z};
  float ts[8]; int nt = 0;
  ts[nt++] = 0.f; ts[nt++] = 1.f;
  for (int i = 0; i < 3; i++) if (dv[i] != 0.f) for (int sgn = -1; sgn <= 1; sgn += 2) {
    const float t = ((float)sgn * hv[i] - av[i]) / dv[i];
    if (t > 0.f && t < 1.f) ts[nt++] = t;
  }
  for (int i = 1; i < nt; i++) { const float v = ts[i]; int j = i - 1; while (j >= 0 && ts[j] > v) { ts[j + 1] = ts[j]; j--; } ts[j + 1] = v; }
  float best = 1e30f, bt = 0.f, tin = 0.f; bool inside = false;
  for (int k = 0; k + 1 < nt; k++) {
    const float t0 = ts[k], t1 = ts[k + 1], tm = 0.5f * (t0 + t1);
    float A = 0.f, B = 0.f; int nout = 0;
    for (int i = 0; i < 3; i++) {
      const float pm = av[i] + tm * dv[i];
      const float r = pm > hv[i] ? hv[i] : (pm < -hv[i] ? -hv[i] : 0.f);
      if (pm > hv[i] || pm < -hv[i]) { A += dv[i] * dv[i]; B += dv[i] * (av[i] - r); nout++; }
    }
    float t = t0;
    if (nout == 0) { inside = true; tin = tm; }   // the inside part of the segment is ONE interval: take its middle
    else if (A > B2S_MINVAL) t = clampf(-B / A, t0, t1);
    float d2 = 0.f;
    for (int i = 0; i < 3; i++) { const float pv = av[i] + t * dv[i]; const float c = clampf(pv, -hv[i], hv[i]); d2 += (pv - c) * (pv - c); }
    if (d2 < best) { best = d2; bt = t; }
  }
  *inside_out = inside;
  if (inside) return tin;
  return bt;
}
DEV int capsule_box(RawCon* c, float margin, v3 p1, const m33& m1, v3 s1, v3 p2, const m33& m2, v3 s2) {
  const v3 axis = col(m1, 2), e0 = p1 - axis * s1.y, e1 = p1 + axis * s1.y;
  const v3 a = mulTv(m2, e0 - p2), b = mulTv(m2, e1 - p2);
  bool inside;
  const float tb = seg_box_closest(a, b - a, s2, &inside);
  const v3 rad = mk3(s1.x, 0.f, 0.f);
  const v3 pt = e0 + (e1 - e0) * tb;
  int n;
  if (!inside) n = sphere_box(c, margin, pt, rad, p2, m2, s2);
  else {
    // segment through the box: contact at the middle of the inside part, pushed out along the box axis of least
    // depth towards the side the capsule CENTRE is on (robust for thin boxes)
    const v3 ps = a + (b - a) * tb, ctr = a + (b - a) * 0.5f;
    const float psv[3] = {ps.x, ps.y, ps.z}, cv[3] = {ctr.x, ctr.y, ctr.z}, hv[3] = {s2.x, s2.y, s2.z};
    int kb = 0; float bestd = hv[0] - fabsf(psv[0]);
    for (int k = 1; k < 3; k++) { const float e = hv[k] - fabsf(psv[k]); if (e < bestd) { bestd = e; kb = k; } }
    const float sg = cv[kb] >= 0.f ? 1.f : -1.f;
    const v3 nl = mk3(kb == 0 ? -sg : 0.f, kb == 1 ? -sg : 0.f, kb == 2 ? -sg : 0.f);
    const float depth = hv[kb] - sg * psv[kb];
    c->n = mulv(m2, nl); c->dist = -depth - s1.x; c->pos = pt + c->n * (s1.x + c->dist * 0.5f); c->t = mk3(0.f, 0.f, 0.f);
    n = 1;
  }
  if (n == 0) return 0;
  for (int e = 0; e < 2 && n < 2; e++) {
    const float te = (float)e;
    if (fabsf(te - tb) * 2.f * s1.y < 1e-3f * s1.x + 1e-6f) continue;
    if (sphere_box(c + n, margin, e == 0 ? e0 : e1, rad, p2, m2, s2) && dot(c[n].n, c[0].n) > 0.9f) n++;
  }
  return n;
}
// box-box: separating-axis test (15 axes), then a face manifold (incident face clipped against the reference face's
// side planes, up to 8 points) or one edge-edge contact (oracle: col_box_box).
// plane - ellipsoid: the support point against the plane normal
DEV int plane_ellipsoid(RawCon* c, float margin, v3 ppos, const m33& pmat, v3 epos, const m33& emat, v3 size) {
  const v3 n = col(pmat, 2), dl = mulTv(emat, n * -1.f);
  const v3 e = mk3(size.x * dl.x, size.y * dl.y, size.z * dl.z);
  const float h = sqrtf(dot(e, e));
  const v3 loc = h > B2S_MINVAL ? mk3(size.x * e.x / h, size.y * e.y / h, size.z * e.z / h) : mk3(size.x, 0.f, 0.f);
  const v3 sp = mulv(emat, loc) + epos;
  const float dist = dot(sp - ppos, n);
  if (dist > margin) return 0;
  c[0].dist = dist; c[0].pos = sp - n * (0.5f * dist); c[0].n = n; c[0].t = mk3(0.f, 0.f, 0.f);
  return 1;
}
// plane - cylinder, up to 4 contacts in a fixed order (oracle col_plane_cylinder): rim point along u (the radial direction
// into the plane) of the nearer cap, the same on the other cap, rim points of the nearer cap at +-120 degrees from u
DEV int plane_cylinder(RawCon* c, float margin, v3 ppos, const m33& pmat, v3 cpos, const m33& cmat, v3 size) {
  const v3 n = col(pmat, 2), ax = col(cmat, 2);
  const float r = size.x, h = size.y, an = dot(ax, n);
  v3 u = (n - ax * an) * -1.f;
  const float ul = sqrtf(dot(u, u));
  u = ul > 1e-6f ? u * (1.0f / ul) : col(cmat, 0);
  const float sgn = an > 0.f ? -1.f : 1.f;
  const v3 v = cross(ax, u);
  int cnt = 0;
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const float cap = i == 1 ? -sgn : sgn, cs = i < 2 ? 1.f : -0.5f, sn = i < 2 ? 0.f : (i == 2 ? 0.8660254037844386f : -0.8660254037844386f);
    const v3 p = cpos + ax * (cap * h) + (u * cs + v * sn) * r;
    const float dist = dot(p - ppos, n);
    if (dist > margin) continue;
#pragma unroll
    for (int k = 0; k < 4; k++) if (k == cnt) { c[k].dist = dist; c[k].pos = p - n * (0.5f * dist); c[k].n = n; c[k].t = mk3(0.f, 0.f, 0.f); }
    cnt++;
  }
  return cnt;
}
DEV int box_box(RawCon* c, float margin, v3 p1, const m33& M1, v3 S1, v3 p2, const m33& M2, v3 S2) {
  const float* m1 = M1.m; const float* m2 = M2.m;
  const float s1[3] = {S1.x, S1.y, S1.z}, s2[3] = {S2.x, S2.y, S2.z};
  const v3 tv = mulTv(M1, p2 - p1);
  const float t[3] = {tv.x, tv.y, tv.z};
  float R[9], AR[9];
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) {
    float v = 0.f;
    for (int k = 0; k < 3; k++) v += m1[3 * k + i] * m2[3 * k + j];
    R[3 * i + j] = v; AR[3 * i + j] = fabsf(v) + 1e-6f;
  }
  float best = -1e30f; int bestax = -1; float bestsign = 1.f;
  for (int i = 0; i < 3; i++) {
    const float rb = s2[0] * AR[3 * i] + s2[1] * AR[3 * i + 1] + s2[2] * AR[3 * i + 2];
    const float sep = fabsf(t[i]) - (s1[i] + rb);
    if (sep > margin) return 0;
    if (sep > best) { best = sep; bestax = i; bestsign = t[i] >= 0.f ? 1.f : -1.f; }
  }
  for (int j = 0; j < 3; j++) {
    const float tj = t[0] * R[j] + t[1] * R[3 + j] + t[2] * R[6 + j];
    const float ra = s1[0] * AR[j] + s1[1] * AR[3 + j] + s1[2] * AR[6 + j];
    const float sep = fabsf(tj) - (ra + s2[j]);
    if (sep > margin) return 0;
    if (sep > best) { best = sep; bestax = 3 + j; bestsign = tj >= 0.f ? 1.f : -1.f; }
  }
  float ebest = -1e30f; int ei = -1, ej = -1; float esign = 1.f, eaxis[3] = {0.f, 0.f, 0.f};
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) {
    const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
    float L[3] = {0.f, 0.f, 0.f};
    L[i1] = -R[3 * i2 + j]; L[i2] = R[3 * i1 + j];
    const float ln = sqrtf(L[0] * L[0] + L[1] * L[1] + L[2] * L[2]);
    if (ln < 1e-6f) continue;
    const float ra = s1[i1] * fabsf(R[3 * i2 + j]) + s1[i2] * fabsf(R[3 * i1 + j]);
    const float rb = s2[j1] * fabsf(R[3 * i + j2]) + s2[j2] * fabsf(R[3 * i + j1]);
    const float tl = t[i2] * R[3 * i1 + j] - t[i1] * R[3 * i2 + j];
    const float sep = (fabsf(tl) - (ra + rb)) / ln;
    if (sep > margin) return 0;
    if (sep > ebest) { ebest = sep; ei = i; ej = j; esign = tl >= 0.f ? 1.f : -1.f; for (int k = 0; k < 3; k++) eaxis[k] = L[k] / ln; }
  }
  if (ei >= 0 && ebest > best + 1e-3f * fabsf(best) + 1e-5f) {
    const float nl[3] = {eaxis[0] * esign, eaxis[1] * esign, eaxis[2] * esign};
    const v3 nw = mulv(M1, mk3(nl[0], nl[1], nl[2]));
    float pa[3] = {p1.x, p1.y, p1.z}, pb[3] = {p2.x, p2.y, p2.z};
    for (int k = 0; k < 3; k++) if (k != ei) {
      const float sg = nl[k] > 0.f ? 1.f : -1.f;
      for (int q = 0; q < 3; q++) pa[q] += sg * s1[k] * m1[3 * q + k];
    }
    for (int k = 0; k < 3; k++) if (k != ej) {
      const float nb = nl[0] * R[k] + nl[1] * R[3 + k] + nl[2] * R[6 + k];
      const float sg = nb > 0.f ? -1.f : 1.f;
      for (int q = 0; q < 3; q++) pb[q] += sg * s2[k] * m2[3 * q + k];
    }
    const v3 ua = col(M1, ei), ub = col(M2, ej);
    const v3 w = mk3(pb[0] - pa[0], pb[1] - pa[1], pb[2] - pa[2]);
    const float uaub = dot(ua, ub), q1 = dot(ua, w), q2 = -dot(ub, w), den = 1.f - uaub * uaub;
    float al = 0.f, be = 0.f;
    if (den > 1e-8f) { al = (q1 + uaub * q2) / den; be = (uaub * q1 + q2) / den; }
    al = clampf(al, -s1[ei], s1[ei]); be = clampf(be, -s2[ej], s2[ej]);
    const v3 A = mk3(pa[0], pa[1], pa[2]) + ua * al, B = mk3(pb[0], pb[1], pb[2]) + ub * be;
    c->dist = ebest; c->pos = (A + B) * 0.5f; c->n = nw; c->t = mk3(0.f, 0.f, 0.f);
    return 1;
  }
  const bool ref1 = bestax < 3;
  const v3 pr = ref1 ? p1 : p2, pi = ref1 ? p2 : p1;
  const float* mr = ref1 ? m1 : m2; const float* mi = ref1 ? m2 : m1;
  const float* sr = ref1 ? s1 : s2; const float* si = ref1 ? s2 : s1;
  const int ax = ref1 ? bestax : bestax - 3;
  const float flip = ref1 ? 1.f : -1.f, sgn = bestsign * flip;
  const v3 nref = mk3(sgn * mr[ax], sgn * mr[3 + ax], sgn * mr[6 + ax]);
  int iax = 0; float imin = 2.f, isg = 1.f;
  for (int k = 0; k < 3; k++) {
    const float dcol = nref.x * mi[k] + nref.y * mi[3 + k] + nref.z * mi[6 + k];
    if (-fabsf(dcol) < imin) { imin = -fabsf(dcol); iax = k; isg = dcol > 0.f ? -1.f : 1.f; }
  }
  const int u = (iax + 1) % 3, v = (iax + 2) % 3;
  v3 poly[16], tmp[16]; int np = 4;
  const float su[4] = {1.f, -1.f, -1.f, 1.f}, sv[4] = {1.f, 1.f, -1.f, -1.f};
  for (int q = 0; q < 4; q++)
    poly[q] = pi + mk3(mi[iax], mi[3 + iax], mi[6 + iax]) * (isg * si[iax]) + mk3(mi[u], mi[3 + u], mi[6 + u]) * (su[q] * si[u]) +
              mk3(mi[v], mi[3 + v], mi[6 + v]) * (sv[q] * si[v]);
  const int ru = (ax + 1) % 3, rv = (ax + 2) % 3;
  for (int side = 0; side < 4 && np > 0; side++) {
    const int ka = side < 2 ? ru : rv; const float sg = (side & 1) ? -1.f : 1.f;
    const v3 pn = mk3(sg * mr[ka], sg * mr[3 + ka], sg * mr[6 + ka]); const float lim = sr[ka];
    int nn = 0;
    for (int q = 0; q < np; q++) {
      const v3 A = poly[q], B = poly[(q + 1) % np];
      const float da = dot(A - pr, pn) - lim, db = dot(B - pr, pn) - lim;
      if (da <= 0.f) tmp[nn++] = A;
      if ((da < 0.f && db > 0.f) || (da > 0.f && db < 0.f)) { const float f = da / (da - db); tmp[nn++] = A + (B - A) * f; }
    }
    np = nn > 16 ? 16 : nn;
    for (int q = 0; q < np; q++) poly[q] = tmp[q];
  }
  int n = 0;
  for (int q = 0; q < np && n < 8; q++) {
    const float dist = dot(poly[q] - pr, nref) - sr[ax];
    if (dist > margin) continue;
    c[n].dist = dist; c[n].pos = poly[q] - nref * (0.5f * dist); c[n].n = nref * flip; c[n].t = mk3(0.f, 0.f, 0.f);
    n++;
  }
  return n;
}

// separating-axis test of two oriented boxes (Gottschalk et al.): t = centre of box 2 in box 1's frame; true when a
// separating axis leaves more than `margin` between them
DEV bool obb_separated(const m33& R1, v3 a, const m33& R2, v3 b, v3 t, float margin) {
  float R[3][3], AR[3][3];
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) {
      R[i][j] = R1.m[i] * R2.m[j] + R1.m[3 + i] * R2.m[3 + j] + R1.m[6 + i] * R2.m[6 + j];     // column i of R1 . column j of R2
      AR[i][j] = fabsf(R[i][j]) + 1e-6f;
    }
  const float A[3] = {a.x, a.y, a.z}, B[3] = {b.x, b.y, b.z}, T[3] = {t.x, t.y, t.z};
#pragma unroll
  for (int i = 0; i < 3; i++) if (fabsf(T[i]) > A[i] + B[0] * AR[i][0] + B[1] * AR[i][1] + B[2] * AR[i][2] + margin) return true;
#pragma unroll
  for (int j = 0; j < 3; j++)
    if (fabsf(T[0] * R[0][j] + T[1] * R[1][j] + T[2] * R[2][j]) > A[0] * AR[0][j] + A[1] * AR[1][j] + A[2] * AR[2][j] + B[j] + margin) return true;
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) {
      const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
      const float ra = A[i1] * AR[i2][j] + A[i2] * AR[i1][j], rb = B[j1] * AR[i][j2] + B[j2] * AR[i][j1];
      // the cross-product axis is not unit length: scale the margin by its length bound (<= 1)
      if (fabsf(T[i2] * R[i1][j] - T[i1] * R[i2][j]) > ra + rb + margin) return true;
    }
  return false;
}

// convex core of geom g (k_convex.cuh): sphere / capsule = point / segment + radius, box, mesh hull
DEV Cvx geom_core(const ModelDev& m, const float* gx, const float* gm, int g) {
  Cvx B;
  const int t = m.geom_type[g];
  B.pos = ld3(gx + 3 * g); B.mat = ldm(gm + 9 * g); B.size = ld3(m.geom_size + 3 * g);
  B.vert = nullptr; B.nvert = 0; B.world = false; B.coop = false; B.radius = 0.f;
  if (t == B2S_GEOM_SPHERE) { B.type = CVX_POINT; B.radius = B.size.x; }
  else if (t == B2S_GEOM_CAPSULE) { B.type = CVX_SEGMENT; B.radius = B.size.x; }
  else if (t == B2S_GEOM_BOX) B.type = CVX_BOX;
  else if (t == B2S_GEOM_CYLINDER) B.type = CVX_CYLINDER;
  else if (t == B2S_GEOM_ELLIPSOID) B.type = CVX_ELLIPSOID;
  else { const int id = m.geom_dataid[g]; B.type = CVX_VERTS; B.vert = m.mesh_vert + 4 * m.mesh_vertadr[id]; B.nvert = m.mesh_vertnum[id]; }
  return B;
}

// MAXC = most contacts one pair of this model can produce (put_model picks 2 / 4 / 8): the contact scratch stays in
// registers for sphere / capsule robots and the box colliders are compiled out where no box pair exists
// PHASE (models with GJK / EPA pairs between two convex cores, ModelDev::col_defer): such a pair costs a warp 2 - 100 us and a
// crumpled robot has a hundred of them, so one world's warp used to hold the whole kernel (measured: mean 154 us per world,
// slowest world 6.6 ms).  The stage is split: PHASE 1 = broadphase, survivor list to global memory, one (world, survivor) TASK
// per deferred pair; k_convex_tasks = one warp per task from the global list, results staged per survivor; PHASE 2 = the
// narrowphase as before, reading the staged result where a pair was deferred.  PHASE 0 = everything in one kernel.
// PHASE 3 = PHASE 2 of a model without height-field pairs: with every core-core pair deferred, only the plane-mesh manifolds are left
// of the convex code, so GJK / EPA drop out of this kernel's call graph (232 -> ~100 registers, 2 -> 5 CTAs per SM).
// One task of k_convex_tasks = 32 bytes that carry everything the task needs from tables: world, survivor index, the two geoms (lower
// type first), pair-table index, the geom types and the hull addresses / sizes.  The task kernel used to find these through
// cvx_task -> col_surv -> geom_type -> geom_dataid -> mesh_vertadr, five dependent loads (~3.5 us, as long as the median task).
DEV void task_record(const ModelDev& m, const DataDev& d, size_t k, int w, int i, unsigned key, int pidx) {
  int g1 = (int)(key >> 16), g2 = (int)(key & 0xffffu);
  if (m.geom_type[g1] > m.geom_type[g2]) { const int t = g1; g1 = g2; g2 = t; }
  const int t1 = m.geom_type[g1], t2 = m.geom_type[g2];
  int va1 = 0, vn1 = 0, va2 = 0, vn2 = 0;
  if (t1 == B2S_GEOM_MESH) { const int id = m.geom_dataid[g1]; va1 = m.mesh_vertadr[id]; vn1 = m.mesh_vertnum[id]; }
  if (t2 == B2S_GEOM_MESH) { const int id = m.geom_dataid[g2]; va2 = m.mesh_vertadr[id]; vn2 = m.mesh_vertnum[id]; }
  int4* r = reinterpret_cast<int4*>(d.cvx_task) + 2 * k;
  r[0] = make_int4(w, i, g1 | (g2 << 16), pidx);
  r[1] = make_int4(t1 | (t2 << 8), va1, va2, vn1 | (vn2 << 16));
}
DEV Cvx task_core(const ModelDev& m, const float* gx, const float* gm, int g, int t, int vadr, int vnum) {
  Cvx B;
  B.pos = ld3(gx + 3 * g); B.mat = ldm(gm + 9 * g); B.size = ld3(m.geom_size + 3 * g);
  B.vert = nullptr; B.nvert = 0; B.world = false; B.coop = false; B.radius = 0.f;
  if (t == B2S_GEOM_SPHERE) { B.type = CVX_POINT; B.radius = B.size.x; }
  else if (t == B2S_GEOM_CAPSULE) { B.type = CVX_SEGMENT; B.radius = B.size.x; }
  else if (t == B2S_GEOM_BOX) B.type = CVX_BOX;
  else if (t == B2S_GEOM_CYLINDER) B.type = CVX_CYLINDER;
  else if (t == B2S_GEOM_ELLIPSOID) B.type = CVX_ELLIPSOID;
  else { B.type = CVX_VERTS; B.vert = m.mesh_vert + 4 * vadr; B.nvert = vnum; B.coop = true; }
  return B;
}
#ifndef CVX_TASK_MINB
#define CVX_TASK_MINB 4
#endif
#define CVX_COST_HEAVY 12      // GJK iterations (+ 64 for EPA) above which a pair's next task is queued with the expensive ones
// PHASE 4 = PHASE 0 (everything in one kernel) of a convex model without height fields and without deferred pairs: its convex pairs
// are all against a plane, so it is instantiated without GJK / EPA as well (Asimov 12-dof: 228 -> ~80 registers).
template <int MAXC, bool CONVEX, int PHASE>
__global__ void __launch_bounds__(32 * B2S_WARPS_PER_CTA) k_collision(const ModelDev m, const DataDev d) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int w = blockIdx.x * B2S_WARPS_PER_CTA + warp;
  if (w >= d.nworld) return;
  // survivors are held as pair keys (min geom id << 16 | max geom id); ascending key order == the order in which
  // the reference enumerates geom pairs, so contacts come out in a fixed order without atomics in the data path
  unsigned* s_list = (unsigned*)b2s_smem + warp * (2 * B2S_MAX_SURVIVORS + 1);
  unsigned* s_sort = s_list + B2S_MAX_SURVIVORS;
  int* s_cnt = (int*)(s_sort + B2S_MAX_SURVIVORS);
  CvxStage* s_stage = (CvxStage*)((unsigned*)b2s_smem + B2S_WARPS_PER_CTA * (2 * B2S_MAX_SURVIVORS + 1)) + warp * CVX_STAGE_CAP;
  // pair-table index of each survivor (convex models without a terrain grid): addresses the pair's cached separating direction
  const bool use_axis = CONVEX && m.ngrid == 0 && d.col_axis != nullptr;
  int* s_pidx = (int*)(s_stage + (size_t)(B2S_WARPS_PER_CTA - warp) * CVX_STAGE_CAP) + warp * B2S_MAX_SURVIVORS;
  int* s_hft = s_pidx + (B2S_WARPS_PER_CTA - warp) * B2S_MAX_SURVIVORS + warp * HF_LIST_CAP;      // live height-field prism tasks of the current pass
  (void)s_hft;
  if (m.disableflags & (B2S_DSBL_CONSTRAINT | B2S_DSBL_CONTACT)) { if (lane == 0) d.ncon[w] = 0; return; }
  const float* gx = d.geom_xpos + (size_t)w * m.ngeom * 3;
  const float* gm = d.geom_xmat + (size_t)w * m.ngeom * 9;
  const float* gfri = m.geom_friction.at(w);
#ifdef COL_TIMING      // development builds only (tools/exp_col_time.py): phase cycles and pair counts into the tail of efc_force
  long long t_ph[4] = {clock64(), 0, 0, 0}; int n_cvx = 0, n_hit = 0;
#endif
  // A1) static pair table (dynamic-dynamic and plane-dynamic pairs, already in key order): bounding spheres
  int nsurv = 0;
  constexpr bool P2 = PHASE == 2 || PHASE == 3;      // second phase
  constexpr bool HF = PHASE != 3 && PHASE != 4;      // height-field (and in-kernel GJK / EPA) code compiled in
  if constexpr (P2) {       // the survivors phase 1 found
    nsurv = d.col_nsurv[w];
    for (int i = lane; i < nsurv; i += 32) s_list[i] = d.col_surv[(size_t)w * B2S_MAX_SURVIVORS + i];
    __syncwarp();
  }
  for (int base = 0; !P2 && base < m.npair; base += 32) {
    const int p = base + lane;
    bool pass = false;
    unsigned key = 0;
    if (p < m.npair) {
      const int g1 = m.pair_geom[2 * p], g2 = m.pair_geom[2 * p + 1];
      const float margin = fmaxf(m.geom_margin[g1], m.geom_margin[g2]);
      const int ty1 = m.geom_type[g1];
      // broadphase volume of geom 2: box (centre c2, axes R2, half sizes b2) inside a sphere of radius brad
      const m33 R2 = ldm(gm + 9 * g2);
      const v3 b2 = ld3(m.geom_bsize + 3 * g2);
      const v3 c2 = ld3(gx + 3 * g2) + (CONVEX ? mulv(R2, ld3(m.geom_bcen + 3 * g2)) : mk3(0.f, 0.f, 0.f));
      if (ty1 == B2S_GEOM_PLANE) {
        const v3 n = mk3(gm[9 * g1 + 2], gm[9 * g1 + 5], gm[9 * g1 + 8]);
        const v3 nl = mulTv(R2, n);
        const float ext = fabsf(nl.x) * b2.x + fabsf(nl.y) * b2.y + fabsf(nl.z) * b2.z;     // box extent along the normal
        pass = dot(c2 - ld3(gx + 3 * g1), n) <= fminf(ext, m.geom_brad[g2]) + margin;
      } else if (CONVEX && ty1 == B2S_GEOM_HFIELD) {
        // bounding sphere of geom 2 against the field's box, in the field's frame
        const float* hs = m.hfield_size + 4 * m.geom_dataid[g1];
        const v3 pl = mulTv(ldm(gm + 9 * g1), c2 - ld3(gx + 3 * g1));
        const float r = m.geom_brad[g2] + margin;
        pass = !(pl.x + r < -hs[0] || pl.x - r > hs[0] || pl.y + r < -hs[1] || pl.y - r > hs[1] || pl.z - r > hs[2] || pl.z + r < -hs[3]);
      } else {
        const m33 R1 = ldm(gm + 9 * g1);
        const v3 c1 = ld3(gx + 3 * g1) + (CONVEX ? mulv(R1, ld3(m.geom_bcen + 3 * g1)) : mk3(0.f, 0.f, 0.f));
        const v3 dif = c2 - c1;
        const float r1 = m.geom_brad[g1] + margin, r2 = m.geom_brad[g2];
        pass = dot(dif, dif) <= (r1 + r2) * (r1 + r2);
        if (pass) {           // sphere 1 against box 2, then sphere 2 against box 1
          const v3 l = mulTv(R2, dif);
          const float ex = fmaxf(fabsf(l.x) - b2.x, 0.f), ey = fmaxf(fabsf(l.y) - b2.y, 0.f), ez = fmaxf(fabsf(l.z) - b2.z, 0.f);
          pass = ex * ex + ey * ey + ez * ez <= r1 * r1;
        }
        if (pass) {
          const v3 l = mulTv(R1, dif), b1 = ld3(m.geom_bsize + 3 * g1);
          const float ex = fmaxf(fabsf(l.x) - b1.x, 0.f), ey = fmaxf(fabsf(l.y) - b1.y, 0.f), ez = fmaxf(fabsf(l.z) - b1.z, 0.f);
          const float rr = r2 + margin;
          pass = ex * ex + ey * ey + ez * ez <= rr * rr;
          // a mesh pair costs a warp-wide GJK: spend a 15-axis box-box separation test (lane-parallel) on it first
          if (CONVEX && pass && (ty1 == B2S_GEOM_MESH || m.geom_type[g2] == B2S_GEOM_MESH)) pass = !obb_separated(R1, b1, R2, b2, l, margin);
        }
      }
      key = ((unsigned)min(g1, g2) << 16) | (unsigned)max(g1, g2);
    }
    const unsigned mask = __ballot_sync(FULL, pass);
    if (pass) { const int pos = nsurv + __popc(mask & ((1u << lane) - 1u)); if (pos < B2S_MAX_SURVIVORS) { s_list[pos] = key; if (CONVEX && use_axis) s_pidx[pos] = p; } }
    nsurv += __popc(mask);
  }
  nsurv = min(nsurv, B2S_MAX_SURVIVORS);
  // A2) static terrain geoms through the uniform xy grid built at put_model: one lane per dynamic geom walks the
  // cells under its bounding sphere; a static geom that spans several cells is taken in its first common cell only
  if (!P2 && m.ngrid > 0) {
    if (lane == 0) *s_cnt = nsurv;
    __syncwarp();
    for (int k = lane; k < m.ndyngeom; k += 32) {
      const int g = m.dyn_geom[k];
      const v3 c = ld3(gx + 3 * g) + (CONVEX ? mulv(ldm(gm + 9 * g), ld3(m.geom_bcen + 3 * g)) : mk3(0.f, 0.f, 0.f));
      const float r = m.geom_brad[g] + fmaxf(m.geom_margin[g], m.grid_margin);
      const int ix0 = max((int)floorf((c.x - r - m.grid_x0) * m.grid_inv), 0), ix1 = min((int)floorf((c.x + r - m.grid_x0) * m.grid_inv), m.grid_nx - 1);
      const int iy0 = max((int)floorf((c.y - r - m.grid_y0) * m.grid_inv), 0), iy1 = min((int)floorf((c.y + r - m.grid_y0) * m.grid_inv), m.grid_ny - 1);
      for (int ix = ix0; ix <= ix1; ix++) for (int iy = iy0; iy <= iy1; iy++) {
        const int cell = ix * m.grid_ny + iy;
        for (int e = m.grid_adr[cell]; e < m.grid_adr[cell + 1]; e++) {
          const int sg = m.grid_geom[e];
          const float* bb = m.grid_aabb + 6 * sg;
          const int sx0 = max((int)floorf((bb[0] - m.grid_x0) * m.grid_inv), 0), sy0 = max((int)floorf((bb[1] - m.grid_y0) * m.grid_inv), 0);
          if (ix != max(ix0, sx0) || iy != max(iy0, sy0)) continue;           // not the first common cell
          if (!((m.geom_contype[sg] & m.geom_conaffinity[g]) || (m.geom_contype[g] & m.geom_conaffinity[sg]))) continue;
          const float ex = fmaxf(fmaxf(bb[0] - c.x, c.x - bb[3]), 0.f), ey = fmaxf(fmaxf(bb[1] - c.y, c.y - bb[4]), 0.f), ez = fmaxf(fmaxf(bb[2] - c.z, c.z - bb[5]), 0.f);
          if (ex * ex + ey * ey + ez * ez > r * r) continue;
          const int pos = atomicAdd(s_cnt, 1);
          if (pos < B2S_MAX_SURVIVORS) s_list[pos] = ((unsigned)min(sg, g) << 16) | (unsigned)max(sg, g);
        }
      }
    }
    __syncwarp();
    const int ntot = min(*s_cnt, B2S_MAX_SURVIVORS);
    if (ntot > nsurv) {
      // rank sort (keys are unique): deterministic order whatever order the lanes appended in
      for (int i = lane; i < ntot; i += 32) {
        const unsigned ki = s_list[i];
        int rank = 0;
        for (int j = 0; j < ntot; j++) rank += s_list[j] < ki;
        s_sort[rank] = ki;
      }
      __syncwarp();
      for (int i = lane; i < ntot; i += 32) s_list[i] = s_sort[i];
    }
    nsurv = ntot;
  }
  __syncwarp();
#ifdef COL_TIMING
  t_ph[1] = clock64();
#endif
  if constexpr (PHASE == 1) {
    // survivors to global memory; one task per pair that needs GJK / EPA between two convex cores
    __syncwarp();
    for (int base = 0; base < nsurv; base += 32) {
      const int i = base + lane;
      bool defer = false;
      if (i < nsurv) {
        const unsigned key = s_list[i];
        d.col_surv[(size_t)w * B2S_MAX_SURVIVORS + i] = key;
        if (use_axis) d.col_pidx[(size_t)w * B2S_MAX_SURVIVORS + i] = s_pidx[i];
        int t1 = m.geom_type[key >> 16], t2 = m.geom_type[key & 0xffffu];
        if (t1 > t2) { const int tt = t1; t1 = t2; t2 = tt; }
        defer = t1 >= B2S_GEOM_SPHERE && (t2 == B2S_GEOM_MESH || t1 == B2S_GEOM_CYLINDER || t1 == B2S_GEOM_ELLIPSOID || t2 == B2S_GEOM_CYLINDER || t2 == B2S_GEOM_ELLIPSOID);
      }
      // pairs whose last task was expensive (a GJK run of many iterations, or EPA: col_cost) are queued from the END of the list and
      // taken first by k_convex_tasks -- a handful of pairs cycle to the GJK iteration cap for 250 us each, and the kernel used to
      // end when the last of them, started at a random time, was done
      const bool heavy = defer && use_axis && d.col_cost[(size_t)w * m.npair + s_pidx[i]] > CVX_COST_HEAVY;
      const unsigned hm = __ballot_sync(FULL, heavy);
      const unsigned dm = __ballot_sync(FULL, defer && !heavy);
      if (dm) {
        int k0 = 0;
        if (lane == 0) k0 = atomicAdd(d.cvx_ntask, __popc(dm));
        k0 = __shfl_sync(FULL, k0, 0);
        if (defer && !heavy) task_record(m, d, (size_t)(k0 + __popc(dm & ((1u << lane) - 1u))), w, i, s_list[i], use_axis ? s_pidx[i] : 0);
      }
      if (hm) {
        int k0 = 0;
        if (lane == 0) k0 = atomicAdd(d.cvx_ntask + 2, __popc(hm));
        k0 = __shfl_sync(FULL, k0, 0);
        if (heavy) task_record(m, d, (size_t)d.nworld * B2S_MAX_SURVIVORS - 1 - (k0 + __popc(hm & ((1u << lane) - 1u))), w, i, s_list[i], s_pidx[i]);
      }
    }
    if (lane == 0) d.col_nsurv[w] = nsurv;
    return;
  }
  // B) narrowphase, 32 survivors per pass, ordered append
  int ncon = 0;
  const int cmax = d.nconmax;
  for (int base = 0; base < nsurv; base += 32) {
    const int i = base + lane;
    RawCon rc[MAXC];
    int n = 0, g1 = 0, g2 = 0;
    float margin = 0.f;
    bool cvx = false;
    if (i < nsurv) {
      const unsigned key = s_list[i];
      g1 = (int)(key >> 16); g2 = (int)(key & 0xffffu);
      if (m.geom_type[g1] > m.geom_type[g2]) { const int tswap = g1; g1 = g2; g2 = tswap; }
      margin = fmaxf(m.geom_margin[g1], m.geom_margin[g2]);
      const int t1 = m.geom_type[g1], t2 = m.geom_type[g2];
      v3 p1 = ld3(gx + 3 * g1), p2 = ld3(gx + 3 * g2), s1 = ld3(m.geom_size + 3 * g1), s2 = ld3(m.geom_size + 3 * g2);
      m33 m1 = ldm(gm + 9 * g1), m2 = ldm(gm + 9 * g2);
      if (t1 == B2S_GEOM_PLANE) {
        if (t2 == B2S_GEOM_SPHERE) n = plane_sphere(rc, margin, p1, col(m1, 2), p2, s2.x);
        else if (t2 == B2S_GEOM_CAPSULE) n = plane_capsule(rc, margin, p1, m1, p2, m2, s2);
        else if (t2 == B2S_GEOM_BOX) { if constexpr (MAXC >= 4) n = plane_box(rc, margin, p1, m1, p2, m2, s2); }
        else if (t2 == B2S_GEOM_CYLINDER) { if constexpr (CONVEX) n = plane_cylinder(rc, margin, p1, m1, p2, m2, s2); }
        else if (t2 == B2S_GEOM_ELLIPSOID) { if constexpr (CONVEX) n = plane_ellipsoid(rc, margin, p1, m1, p2, m2, s2); }
      } else if (t1 == B2S_GEOM_SPHERE) {
        if (t2 == B2S_GEOM_SPHERE) n = sphere_sphere(rc, margin, p1, s1.x, p2, s2.x);
        else if (t2 == B2S_GEOM_CAPSULE) n = sphere_capsule(rc, margin, p1, s1, p2, m2, s2);
        else if (t2 == B2S_GEOM_BOX) n = sphere_box(rc, margin, p1, s1, p2, m2, s2);
      } else if (t1 == B2S_GEOM_CAPSULE) {
        if (t2 == B2S_GEOM_CAPSULE) n = capsule_capsule(rc, margin, p1, m1, s1, p2, m2, s2);
        else if (t2 == B2S_GEOM_BOX) n = capsule_box(rc, margin, p1, m1, s1, p2, m2, s2);
      } else if (t1 == B2S_GEOM_BOX && t2 == B2S_GEOM_BOX) { if constexpr (MAXC >= 8) n = box_box(rc, margin, p1, m1, s1, p2, m2, s2); }
      // every pair with a mesh, a height field, a cylinder or an ellipsoid (other than plane - cylinder / ellipsoid) is convex work
      if constexpr (CONVEX) cvx = t2 == B2S_GEOM_MESH || t1 == B2S_GEOM_HFIELD ||
                                  (t1 >= B2S_GEOM_SPHERE && (t1 == B2S_GEOM_CYLINDER || t1 == B2S_GEOM_ELLIPSOID || t2 == B2S_GEOM_CYLINDER || t2 == B2S_GEOM_ELLIPSOID));
      if constexpr (CONVEX && P2) {
        if (cvx && t1 >= B2S_GEOM_SPHERE) {      // deferred pair: k_convex_tasks staged its result
          const CvxStage* sg = reinterpret_cast<const CvxStage*>(d.cvx_stage + ((size_t)w * B2S_MAX_SURVIVORS + i) * 8);
          n = sg->idx;
          if (n) { rc[0].dist = sg->dist; rc[0].pos = sg->pos; rc[0].n = sg->n; rc[0].t = mk3(0.f, 0.f, 0.f); }
          cvx = false;
        }
      }
    }
    if constexpr (CONVEX) {
      // mesh / height-field pairs of this pass, one at a time with the whole warp: hull scans are lane-strided, a
      // height field's prisms are handed out one per lane; the results go back to the pair's lane through smem
      // height-field pairs first, flattened: every (pair, prism) of this pass is one task, 32 tasks per round, so a
      // capsule over 12 prisms and a sphere over 4 share rounds.  Hits are staged in task order = (pair, prism) order;
      // each pair's lane then takes its segment.  If the stage overflows (robot buried in the field) the pairs fall
      // back to the one-pair-at-a-time driver below, which has no such limit.
      bool hf = false, hf_done = false;
      if constexpr (HF) { if (cvx) hf = m.geom_type[g1] == B2S_GEOM_HFIELD; }
      if constexpr (HF) if (__any_sync(FULL, hf)) {
        HfRange R; R.cmin = 0; R.ncell = 0; R.rmin = 0; R.T = 0; R.loz = 0.f;
        if (hf) {
          Cvx B = geom_core(m, gx, gm, g2);
          const int h = m.geom_dataid[g1];
          R = hfield_range(margin, m.hfield_size + 4 * h, m.hfield_nrow[h], m.hfield_ncol[h], ld3(gx + 3 * g1), ldm(gm + 9 * g1), B);
        }
        int Ttot;
        const int off = warp_excl_scan(R.T, lane, &Ttot);
        // pass A: the cheap test of every (pair, prism) of this pass -- prism top above B's lowest point -- 32 per round; the live
        // ones are compacted, in (pair, prism) order, into a task list.  Pass B then runs GJK / EPA on full rounds of live tasks
        // (a Go1 pass has ~120 prisms under its geoms' bounding boxes and ~30 live ones: 1 - 2 GJK rounds instead of 4).
        int nlive = 0;
#pragma unroll 1
        for (int q0 = 0; q0 < Ttot; q0 += 32) {
          const int q = min(q0 + lane, Ttot - 1);
          int own = 0;                                  // largest lane whose offset is <= q
#pragma unroll
          for (int b = 16; b > 0; b >>= 1) { const int cand = own + b; const int o = __shfl_sync(FULL, off, cand & 31); if (cand < 32 && o <= q) own = cand; }
          const int og1 = __shfl_sync(FULL, g1, own), ooff = __shfl_sync(FULL, off, own);
          HfRange OR;
          OR.cmin = __shfl_sync(FULL, R.cmin, own); OR.ncell = __shfl_sync(FULL, R.ncell, own); OR.rmin = __shfl_sync(FULL, R.rmin, own);
          OR.T = 0; OR.loz = __shfl_sync(FULL, R.loz, own);
          bool live = false;
          if (q0 + lane < Ttot) {
            const int h = m.geom_dataid[og1];
            live = hfield_prism_live(q - ooff, OR, m.hfield_size + 4 * h, m.hfield_ncol[h], m.hfield_data + m.hfield_adr[h]);
          }
          const unsigned lm = __ballot_sync(FULL, live);
          const int pos = nlive + __popc(lm & ((1u << lane) - 1u));
          if (live && pos < HF_LIST_CAP) s_hft[pos] = (own << 26) | (q - ooff);
          nlive += __popc(lm);
        }
        __syncwarp();
        int nst = 0;
        if (nlive > HF_LIST_CAP) nst = CVX_STAGE_CAP + 1;      // too many for the list: the one-pair-at-a-time driver below takes these pairs
#pragma unroll 1
        for (int e0 = 0; e0 < nlive && nlive <= HF_LIST_CAP; e0 += 32) {
          const int e = min(e0 + lane, nlive - 1);
          const int ent = s_hft[e], own = (ent >> 26) & 31, t = ent & 0x3ffffff;
          const int og1 = __shfl_sync(FULL, g1, own), og2 = __shfl_sync(FULL, g2, own);
          const float om = __shfl_sync(FULL, margin, own);
          HfRange OR;
          OR.cmin = __shfl_sync(FULL, R.cmin, own); OR.ncell = __shfl_sync(FULL, R.ncell, own); OR.rmin = __shfl_sync(FULL, R.rmin, own);
          OR.T = 0; OR.loz = __shfl_sync(FULL, R.loz, own);
          bool has = false;
          CvxStage one;
          if (e0 + lane < nlive) {
            const Cvx B = geom_core(m, gx, gm, og2);
            const int h = m.geom_dataid[og1];
            has = hfield_prism_task(one, t, OR, om, m.hfield_size + 4 * h, m.hfield_nrow[h], m.hfield_ncol[h], m.hfield_data + m.hfield_adr[h],
                                    ld3(gx + 3 * og1), ldm(gm + 9 * og1), B);
            one.idx |= own << 26;
          }
          const unsigned hm = __ballot_sync(FULL, has);
          const int pos = nst + __popc(hm & ((1u << lane) - 1u));
          if (has && pos < CVX_STAGE_CAP) s_stage[pos] = one;
          nst += __popc(hm);
        }
        __syncwarp();
        hf_done = nst <= CVX_STAGE_CAP;
        if (hf_done && hf) {
          int cnt = 0;
          for (int e = 0; e < nst; e++) cnt += (s_stage[e].idx >> 26) == lane;
          n = 0;
          for (int e = 0; e < nst; e++) {
            if ((s_stage[e].idx >> 26) != lane) continue;
            bool keep = true;
            if (cnt > MAXC) {        // the MAXC deepest by (dist, prism id), still in prism order
              int rank = 0;
              const float de = s_stage[e].dist; const int ie = s_stage[e].idx;
              for (int j = 0; j < nst; j++)
                if ((s_stage[j].idx >> 26) == lane) rank += (s_stage[j].dist < de) || (s_stage[j].dist == de && s_stage[j].idx < ie);
              keep = rank < MAXC;
            }
            if (keep) {
#pragma unroll
              for (int k = 0; k < MAXC; k++) if (k == n) { rc[k].dist = s_stage[e].dist; rc[k].pos = s_stage[e].pos; rc[k].n = s_stage[e].n; rc[k].t = mk3(0.f, 0.f, 0.f); }
              n++;
            }
          }
        }
        __syncwarp();
      }
      unsigned todo = __ballot_sync(FULL, cvx && !(hf && hf_done));
#ifdef COL_TIMING
      n_cvx += __popc(todo);
#endif
      while (todo) {
#ifdef COL_TIMING
        const long long t_c0 = clock64();
#endif
        const int L = __ffs(todo) - 1;
        todo &= todo - 1;
        const int bg1 = __shfl_sync(FULL, g1, L), bg2 = __shfl_sync(FULL, g2, L);
        const float bm = __shfl_sync(FULL, margin, L);
        const int t1 = m.geom_type[bg1], t2 = m.geom_type[bg2];
        const v3 p1 = ld3(gx + 3 * bg1), p2 = ld3(gx + 3 * bg2), s1 = ld3(m.geom_size + 3 * bg1), s2 = ld3(m.geom_size + 3 * bg2);
        const m33 m1 = ldm(gm + 9 * bg1), m2 = ldm(gm + 9 * bg2);
        Cvx A, B;
        B.pos = p2; B.mat = m2; B.size = s2; B.vert = nullptr; B.nvert = 0; B.world = false; B.coop = false; B.radius = 0.f;
        if (t2 == B2S_GEOM_SPHERE) { B.type = CVX_POINT; B.radius = s2.x; }
        else if (t2 == B2S_GEOM_CAPSULE) { B.type = CVX_SEGMENT; B.radius = s2.x; }
        else if (t2 == B2S_GEOM_BOX) B.type = CVX_BOX;
        else if (t2 == B2S_GEOM_CYLINDER) B.type = CVX_CYLINDER;
        else if (t2 == B2S_GEOM_ELLIPSOID) B.type = CVX_ELLIPSOID;
        else { const int id = m.geom_dataid[bg2]; B.type = CVX_VERTS; B.vert = m.mesh_vert + 4 * m.mesh_vertadr[id]; B.nvert = m.mesh_vertnum[id]; }
        int nres;
        if (t1 == B2S_GEOM_PLANE) nres = plane_mesh_warp(s_stage, bm, p1, m1, p2, m2, B.vert, B.nvert, lane);
        else if constexpr (HF) {
        if (t1 == B2S_GEOM_HFIELD) {
          const int h = m.geom_dataid[bg1];
          nres = hfield_convex_warp<MAXC>(s_stage, bm, m.hfield_size + 4 * h, m.hfield_nrow[h], m.hfield_ncol[h], m.hfield_data + m.hfield_adr[h],
                                          p1, m1, B, lane);
        } else {
          A.pos = p1; A.mat = m1; A.size = s1; A.vert = nullptr; A.nvert = 0; A.world = false; A.coop = false; A.radius = 0.f;
          if (t1 == B2S_GEOM_SPHERE) { A.type = CVX_POINT; A.radius = s1.x; }
          else if (t1 == B2S_GEOM_CAPSULE) { A.type = CVX_SEGMENT; A.radius = s1.x; }
          else if (t1 == B2S_GEOM_BOX) A.type = CVX_BOX;
          else if (t1 == B2S_GEOM_CYLINDER) A.type = CVX_CYLINDER;
          else if (t1 == B2S_GEOM_ELLIPSOID) A.type = CVX_ELLIPSOID;
          else { const int id = m.geom_dataid[bg1]; A.type = CVX_VERTS; A.vert = m.mesh_vert + 4 * m.mesh_vertadr[id]; A.nvert = m.mesh_vertnum[id]; A.coop = true; }
          B.coop = (B.type == CVX_VERTS);
          float* axis = use_axis ? d.col_axis + ((size_t)w * m.npair + s_pidx[base + L]) * 3 : nullptr;
          nres = convex_pair_warp(s_stage, bm, A, B, lane, axis);
        }
        } else nres = 0;      // PHASE 3: every core-core pair was a task, no height fields
        __syncwarp();
#ifdef COL_TIMING
        n_hit += nres > 0; t_ph[2] += clock64() - t_c0;
#endif
        if (lane == L) {
          n = nres;
#pragma unroll
          for (int k = 0; k < MAXC; k++) if (k < nres) { rc[k].dist = s_stage[k].dist; rc[k].pos = s_stage[k].pos; rc[k].n = s_stage[k].n; rc[k].t = mk3(0.f, 0.f, 0.f); }
        }
        __syncwarp();
      }
    }
    int total;
    const int excl = warp_excl_scan(n, lane, &total);
    if (n > 0) {
      // contact parameters (mj_contactParam): priority, else max condim / max friction / solmix-weighted solref, solimp
      int condim; float f0, f1, f2, sr0, sr1, si[5];
      const int pr1 = m.geom_priority[g1], pr2 = m.geom_priority[g2];
      if (pr1 != pr2) {
        const int g = pr1 > pr2 ? g1 : g2;
        condim = m.geom_condim[g]; f0 = gfri[3 * g]; f1 = gfri[3 * g + 1]; f2 = gfri[3 * g + 2];
        sr0 = m.geom_solref[2 * g]; sr1 = m.geom_solref[2 * g + 1];
        for (int k = 0; k < 5; k++) si[k] = m.geom_solimp[5 * g + k];
      } else {
        condim = max(m.geom_condim[g1], m.geom_condim[g2]);
        f0 = fmaxf(gfri[3 * g1], gfri[3 * g2]); f1 = fmaxf(gfri[3 * g1 + 1], gfri[3 * g2 + 1]); f2 = fmaxf(gfri[3 * g1 + 2], gfri[3 * g2 + 2]);
        const float a = m.geom_solmix[g1], b = m.geom_solmix[g2];
        float mix;
        if (a >= B2S_MINVAL && b >= B2S_MINVAL) mix = a / (a + b);
        else if (a < B2S_MINVAL && b < B2S_MINVAL) mix = 0.5f;
        else mix = (a < B2S_MINVAL) ? 0.f : 1.f;
        if (m.geom_solref[2 * g1] > 0.f && m.geom_solref[2 * g2] > 0.f) {
          sr0 = mix * m.geom_solref[2 * g1] + (1.f - mix) * m.geom_solref[2 * g2];
          sr1 = mix * m.geom_solref[2 * g1 + 1] + (1.f - mix) * m.geom_solref[2 * g2 + 1];
        } else {
          sr0 = fminf(m.geom_solref[2 * g1], m.geom_solref[2 * g2]); sr1 = fminf(m.geom_solref[2 * g1 + 1], m.geom_solref[2 * g2 + 1]);
        }
        for (int k = 0; k < 5; k++) si[k] = mix * m.geom_solimp[5 * g1 + k] + (1.f - mix) * m.geom_solimp[5 * g2 + k];
      }
      f0 = fmaxf(f0, B2S_MINMU); f1 = fmaxf(f1, B2S_MINMU); f2 = fmaxf(f2, B2S_MINMU);
      const float gap = fmaxf(m.geom_gap[g1], m.geom_gap[g2]);
#pragma unroll
      for (int k = 0; k < MAXC; k++) {
        if (k >= n) break;
        const int c = ncon + excl + k;
        if (c >= cmax) break;
        const size_t o = (size_t)w * cmax + c;
        B2S_ASSERT(c >= 0 && g1 >= 0 && g1 < m.ngeom && g2 >= 0 && g2 < m.ngeom && condim >= 1 && condim <= 6);
        d.con_dist[o] = rc[k].dist;
        st3(d.con_pos + 3 * o, rc[k].pos);
        make_frame(rc[k].n, rc[k].t, d.con_frame + 9 * o);
        d.con_includemargin[o] = margin - gap;
        float* fr = d.con_friction + 5 * o;
        fr[0] = f0; fr[1] = f0; fr[2] = f1; fr[3] = f2; fr[4] = f2;
        d.con_solref[2 * o] = sr0; d.con_solref[2 * o + 1] = sr1;
        for (int q = 0; q < 5; q++) d.con_solimp[5 * o + q] = si[q];
        d.con_dim[o] = condim; d.con_geom[2 * o] = g1; d.con_geom[2 * o + 1] = g2; d.con_efc_address[o] = -1;
      }
    }
    ncon += total;
  }
  B2S_ASSERT(nsurv <= B2S_MAX_SURVIVORS);
#ifdef COL_TIMING
  if (lane == 0) {
    float* o = d.efc_force + (size_t)(w + 1) * d.njmax - 8;
    o[0] = (float)(t_ph[1] - t_ph[0]); o[1] = (float)(clock64() - t_ph[1]); o[2] = (float)t_ph[2]; o[3] = (float)nsurv; o[4] = (float)n_cvx; o[5] = (float)n_hit;
  }
#endif
  if (lane == 0) d.ncon[w] = min(ncon, cmax);
}

// One warp per (world, survivor) task of the global list phase 1 filled: GJK / EPA between the two convex cores with the hull
// scans shared by the 32 lanes, result staged for phase 2.  A persistent grid strides over the list, so a world with a hundred
// such pairs is spread over the whole machine instead of serialising on its own warp.
__global__ void __launch_bounds__(32 * B2S_WARPS_PER_CTA, CVX_TASK_MINB) k_convex_tasks(const ModelDev m, const DataDev d) {
  const int lane = threadIdx.x & 31;
  const int gw = blockIdx.x * B2S_WARPS_PER_CTA + (threadIdx.x >> 5), nwarp = gridDim.x * B2S_WARPS_PER_CTA;
  const int cap = d.nworld * B2S_MAX_SURVIVORS, nheavy = min(d.cvx_ntask[2], cap), ntask = min(d.cvx_ntask[0] + nheavy, cap);
  const bool use_axis = m.ngrid == 0;
  EpaWork* wk = reinterpret_cast<EpaWork*>(b2s_smem) + (threadIdx.x >> 5);       // this warp's EPA polytope
  // tasks differ by two orders of magnitude (a cached separating plane: one support evaluation; a deep overlap: GJK + 24 EPA
  // rounds): warps pull the next task from a cursor instead of striding statically
  (void)gw; (void)nwarp;
#pragma unroll 1
  while (true) {
    // one task per cursor increment: chunks of 4 / 8 consecutive cheap tasks per increment were measured slower (5.24 -> 4.88 / 4.45 M:
    // consecutive tasks belong to one world and their costs are correlated)
    int k = 0;
    if (lane == 0) k = atomicAdd(d.cvx_ntask + 1, 1);
    k = __shfl_sync(FULL, k, 0);
    if (k >= ntask) break;
    k = k < nheavy ? cap - 1 - k : k - nheavy;      // the expensive tasks first (queued from the end of the list), then the others
    const int4 r0 = reinterpret_cast<const int4*>(d.cvx_task)[2 * (size_t)k], r1 = reinterpret_cast<const int4*>(d.cvx_task)[2 * (size_t)k + 1];
    const int w = r0.x, i = r0.y, g1 = r0.z & 0xffff, g2 = r0.z >> 16;
    const float* gx = d.geom_xpos + (size_t)w * m.ngeom * 3;
    const float* gm = d.geom_xmat + (size_t)w * m.ngeom * 9;
    const float margin = fmaxf(m.geom_margin[g1], m.geom_margin[g2]);
    const Cvx A = task_core(m, gx, gm, g1, r1.x & 0xff, r1.y, r1.w & 0xffff), B = task_core(m, gx, gm, g2, r1.x >> 8, r1.z, r1.w >> 16);
    CvxStage* out = reinterpret_cast<CvxStage*>(d.cvx_stage + ((size_t)w * B2S_MAX_SURVIVORS + i) * 8);
    const size_t pslot = use_axis ? (size_t)w * m.npair + r0.w : 0;
    float* axis = use_axis ? d.col_axis + pslot * 3 : nullptr;
    int work = 0;
#ifdef COL_TIMING      // development builds only (tools/exp_col_time.py): the consumed task record becomes (iterations | EPA rounds | hull vertices, task cycles / 64 | EPA cycles / 64 << 16)
    const long long t_task = clock64();
    const int nres = convex_pair_warp(out, margin, A, B, lane, axis, wk, d.cvx_task + 8 * (size_t)k, &work);
    if (lane == 0) d.cvx_task[8 * (size_t)k + 1] |= (int)min((clock64() - t_task) >> 6, 65535LL);
#else
    const int nres = convex_pair_warp(out, margin, A, B, lane, axis, wk, nullptr, &work);
#endif
    __syncwarp();
    if (lane == 0) { out->idx = nres; if (use_axis) d.col_cost[pslot] = work; }
    __syncwarp();
  }
}

#define COL_DISPATCH(M, EXPR)                                                                   \
  if ((M).col_convex && !(M).col_hf && !(M).col_defer) { constexpr int MAXC = 8; constexpr bool CONVEX = true; constexpr int PH = 4; EXPR; } \
  else if ((M).col_convex) { constexpr int MAXC = 8; constexpr bool CONVEX = true; constexpr int PH = 0; EXPR; }           \
  else switch ((M).col_maxc) {                                                                  \
    case 2: { constexpr int MAXC = 2; constexpr bool CONVEX = false; constexpr int PH = 0; EXPR; } break;             \
    case 4: { constexpr int MAXC = 4; constexpr bool CONVEX = false; constexpr int PH = 0; EXPR; } break;             \
    default: { constexpr int MAXC = 8; constexpr bool CONVEX = false; constexpr int PH = 0; EXPR; } break;            \
  }
void launch_collision(const ModelDev& m, const DataDev& d, cudaStream_t st) {
  const int grid = (d.nworld + B2S_WARPS_PER_CTA - 1) / B2S_WARPS_PER_CTA;
  if (m.col_convex && m.col_defer) {      // broadphase + task emission, the GJK / EPA tasks over the whole machine, narrowphase + ordered append
    k_collision<8, true, 1><<<grid, 32 * B2S_WARPS_PER_CTA, smem_collision(m), st>>>(m, d);
    k_convex_tasks<<<148 * CVX_TASK_MINB, 32 * B2S_WARPS_PER_CTA, B2S_WARPS_PER_CTA * sizeof(EpaWork), st>>>(m, d);
    if (m.col_hf) k_collision<8, true, 2><<<grid, 32 * B2S_WARPS_PER_CTA, smem_collision(m), st>>>(m, d);
    else k_collision<8, true, 3><<<grid, 32 * B2S_WARPS_PER_CTA, smem_collision(m), st>>>(m, d);
    b2s_count_launch(); b2s_count_launch(); b2s_count_launch();
    return;
  }
  COL_DISPATCH(m, (k_collision<MAXC, CONVEX, PH><<<grid, 32 * B2S_WARPS_PER_CTA, smem_collision(m), st>>>(m, d)));
  b2s_count_launch();
}
size_t smem_collision(const ModelDev& m) {
  return (size_t)B2S_WARPS_PER_CTA * ((2 * B2S_MAX_SURVIVORS + 1) * sizeof(int) + (m.col_convex ? CVX_STAGE_CAP * sizeof(CvxStage) + (B2S_MAX_SURVIVORS + HF_LIST_CAP) * sizeof(int) : 0));
}
const void* fn_collision(const ModelDev& m) {
  const void* f = nullptr;
  COL_DISPATCH(m, (f = (const void*)k_collision<MAXC, CONVEX, PH>));
  return f;
}
