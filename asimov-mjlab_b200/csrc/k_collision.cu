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

extern __shared__ float b2s_smem[];
#define B2S_MAX_SURVIVORS 160

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

__global__ void __launch_bounds__(32 * B2S_WARPS_PER_CTA) k_collision(const ModelDev m, const DataDev d) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int w = blockIdx.x * B2S_WARPS_PER_CTA + warp;
  if (w >= d.nworld) return;
  int* s_list = (int*)b2s_smem + warp * B2S_MAX_SURVIVORS;
  if (m.disableflags & (B2S_DSBL_CONSTRAINT | B2S_DSBL_CONTACT)) { if (lane == 0) d.ncon[w] = 0; return; }
  const float* gx = d.geom_xpos + (size_t)w * m.ngeom * 3;
  const float* gm = d.geom_xmat + (size_t)w * m.ngeom * 9;
  const float* gfri = m.geom_friction.at(w);
  // A) broadphase over the static pair table
  int nsurv = 0;
  for (int base = 0; base < m.npair; base += 32) {
    const int p = base + lane;
    bool pass = false;
    if (p < m.npair) {
      const int g1 = m.pair_geom[2 * p], g2 = m.pair_geom[2 * p + 1];
      const float margin = fmaxf(m.geom_margin[g1], m.geom_margin[g2]);
      v3 dif = ld3(gx + 3 * g2) - ld3(gx + 3 * g1);
      if (m.geom_type[g1] == B2S_GEOM_PLANE) {
        v3 n = mk3(gm[9 * g1 + 2], gm[9 * g1 + 5], gm[9 * g1 + 8]);
        pass = dot(dif, n) <= m.geom_rbound[g2] + margin;
      } else {
        const float bound = m.geom_rbound[g1] + m.geom_rbound[g2] + margin;
        pass = dot(dif, dif) <= bound * bound;
      }
    }
    const unsigned mask = __ballot_sync(FULL, pass);
    if (pass) { const int pos = nsurv + __popc(mask & ((1u << lane) - 1u)); if (pos < B2S_MAX_SURVIVORS) s_list[pos] = p; }
    nsurv += __popc(mask);
  }
  nsurv = min(nsurv, B2S_MAX_SURVIVORS);
  __syncwarp();
  // B) narrowphase, 32 survivors per pass, ordered append
  int ncon = 0;
  const int cmax = d.nconmax;
  for (int base = 0; base < nsurv; base += 32) {
    const int i = base + lane;
    RawCon rc[4];
    int n = 0, g1 = 0, g2 = 0;
    float margin = 0.f;
    if (i < nsurv) {
      const int p = s_list[i];
      g1 = m.pair_geom[2 * p]; g2 = m.pair_geom[2 * p + 1];
      margin = fmaxf(m.geom_margin[g1], m.geom_margin[g2]);
      const int t1 = m.geom_type[g1], t2 = m.geom_type[g2];
      v3 p1 = ld3(gx + 3 * g1), p2 = ld3(gx + 3 * g2), s1 = ld3(m.geom_size + 3 * g1), s2 = ld3(m.geom_size + 3 * g2);
      m33 m1 = ldm(gm + 9 * g1), m2 = ldm(gm + 9 * g2);
      if (t1 == B2S_GEOM_PLANE) {
        if (t2 == B2S_GEOM_SPHERE) n = plane_sphere(rc, margin, p1, col(m1, 2), p2, s2.x);
        else if (t2 == B2S_GEOM_CAPSULE) n = plane_capsule(rc, margin, p1, m1, p2, m2, s2);
        else if (t2 == B2S_GEOM_BOX) n = plane_box(rc, margin, p1, m1, p2, m2, s2);
      } else if (t1 == B2S_GEOM_SPHERE) {
        if (t2 == B2S_GEOM_SPHERE) n = sphere_sphere(rc, margin, p1, s1.x, p2, s2.x);
        else if (t2 == B2S_GEOM_CAPSULE) n = sphere_capsule(rc, margin, p1, s1, p2, m2, s2);
        else if (t2 == B2S_GEOM_BOX) n = sphere_box(rc, margin, p1, s1, p2, m2, s2);
      } else if (t1 == B2S_GEOM_CAPSULE && t2 == B2S_GEOM_CAPSULE) n = capsule_capsule(rc, margin, p1, m1, s1, p2, m2, s2);
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
      for (int k = 0; k < n; k++) {
        const int c = ncon + excl + k;
        if (c >= cmax) break;
        const size_t o = (size_t)w * cmax + c;
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
  if (lane == 0) d.ncon[w] = min(ncon, cmax);
}

void launch_collision(const ModelDev& m, const DataDev& d, cudaStream_t st) {
  const int grid = (d.nworld + B2S_WARPS_PER_CTA - 1) / B2S_WARPS_PER_CTA;
  k_collision<<<grid, 32 * B2S_WARPS_PER_CTA, smem_collision(m), st>>>(m, d);
  b2s_count_launch();
}
size_t smem_collision(const ModelDev&) { return (size_t)B2S_WARPS_PER_CTA * B2S_MAX_SURVIVORS * sizeof(int); }
const void* fn_collision() { return (const void*)k_collision; }
