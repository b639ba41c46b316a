/* convex_oracle.h -- TEST INFRASTRUCTURE (CPU oracle), included by mjstep_oracle.c.
 *
 * convex / height-field narrowphase -- GJK distance + EPA penetration between convex
 * cores (point, segment, box, vertex cloud, height-field prism), spheres and capsules handled as a core plus a radius;
 * plane-mesh by vertex scan with a 4-point manifold; height fields as triangular prisms over the sub-grid under the
 * geom's bounding sphere.  mujoco_warp (collision_convex.py, collision_gjk.py @ e605c406) is NOT vendored in
 * /root/reference, so this is a restatement of the published algorithms (Gilbert-Johnson-Keerthi 1988 with the
 * Voronoi-region sub-simplex solve of Ericson, "Real-Time Collision Detection" ch. 5/9; van den Bergen's EPA) wired the
 * way MuJoCo's documentation describes its convex pipeline (one contact per convex pair, one per prism for height
 * fields, geom1 = the lower geom type, normal from geom1 to geom2).  Reference call sites that need it:
 * src/mjlab/asset_zoo/robots/asimov/xmls/asimov_toe.xml:12,49 (mesh collision geoms) and
 * src/mjlab/terrains/heightfield_terrains.py:213-242 (hfield terrain).  PARITY UNPINNED.
 */
#ifndef CONVEX_ORACLE_H
#define CONVEX_ORACLE_H

enum { CVX_POINT = 0, CVX_SEGMENT = 1, CVX_BOX = 2, CVX_VERTS = 3, CVX_CYLINDER = 4, CVX_ELLIPSOID = 5 };
typedef struct {
  int type;
  real pos[3];
  real mat[9];
  real size[3];        /* SEGMENT: size[1] = half length along local z; BOX: half sizes; CYLINDER: radius, half height (z);
                          ELLIPSOID: the three radii */
  const double* vert;  /* VERTS: model vertices (double, local frame) ... */
  const real* vertr;   /* ... or `real` vertices already in the world frame when mat is unused (prisms) */
  int nvert;
  real radius;         /* inflation (sphere / capsule) */
} cvx;

#define GJK_MAXIT 48
#define EPA_MAXIT 48
#define EPA_MAXV 28
#define EPA_MAXF 56       /* a closed triangle mesh with V vertices has 2V - 4 faces */
#define CVX_TOL ((real)1e-6)

/* support point of the CORE of `s` in world direction `dir` (ties: lowest vertex index) */
static void cvx_support(const cvx* s, const real* dir, real* out) {
  if (s->type == CVX_POINT) { out[0] = s->pos[0]; out[1] = s->pos[1]; out[2] = s->pos[2]; return; }
  if (s->type == CVX_VERTS && s->vertr) {
    int best = 0; real bd = dot3(s->vertr, dir);
    for (int i = 1; i < s->nvert; i++) { real t = dot3(s->vertr + 3*i, dir); if (t > bd) { bd = t; best = i; } }
    out[0] = s->vertr[3*best]; out[1] = s->vertr[3*best + 1]; out[2] = s->vertr[3*best + 2];
    return;
  }
  real dl[3], loc[3];
  matT_vec(dl, s->mat, dir);
  if (s->type == CVX_SEGMENT) { loc[0] = 0; loc[1] = 0; loc[2] = dl[2] >= 0 ? s->size[1] : -s->size[1]; }
  else if (s->type == CVX_BOX) { for (int k = 0; k < 3; k++) loc[k] = dl[k] >= 0 ? s->size[k] : -s->size[k]; }
  else if (s->type == CVX_CYLINDER) {      /* rim point of the cap facing dir */
    real h = (real)sqrt((double)(dl[0]*dl[0] + dl[1]*dl[1]));
    loc[0] = h > MINVAL ? s->size[0]*dl[0]/h : 0; loc[1] = h > MINVAL ? s->size[0]*dl[1]/h : 0;
    loc[2] = dl[2] >= 0 ? s->size[1] : -s->size[1];
  } else if (s->type == CVX_ELLIPSOID) {   /* S^2 d / |S d| */
    real e[3] = {s->size[0]*dl[0], s->size[1]*dl[1], s->size[2]*dl[2]};
    real h = (real)sqrt((double)dot3(e, e));
    for (int k = 0; k < 3; k++) loc[k] = h > MINVAL ? s->size[k]*e[k]/h : (k == 0 ? s->size[0] : 0);
  } else {
    int best = 0; real bd = (real)s->vert[0]*dl[0] + (real)s->vert[1]*dl[1] + (real)s->vert[2]*dl[2];
    for (int i = 1; i < s->nvert; i++) {
      real t = (real)s->vert[3*i]*dl[0] + (real)s->vert[3*i + 1]*dl[1] + (real)s->vert[3*i + 2]*dl[2];
      if (t > bd) { bd = t; best = i; }
    }
    loc[0] = (real)s->vert[3*best]; loc[1] = (real)s->vert[3*best + 1]; loc[2] = (real)s->vert[3*best + 2];
  }
  mat_vec(out, s->mat, loc);
  out[0] += s->pos[0]; out[1] += s->pos[1]; out[2] += s->pos[2];
}

/* Minkowski-difference support: y = supA(dir) - supB(-dir); also returns supA */
static void mink_support(const cvx* A, const cvx* B, const real* dir, real* y, real* a) {
  real b[3], nd[3] = {-dir[0], -dir[1], -dir[2]};
  cvx_support(A, dir, a);
  cvx_support(B, nd, b);
  y[0] = a[0] - b[0]; y[1] = a[1] - b[1]; y[2] = a[2] - b[2];
}

/* ---- closest point to the origin on a simplex (Ericson 5.1); reduces the simplex in place.
 * Y: up to 4 Minkowski points, Pa: matching supports on A, *n: count.  Writes barycentric weights lam[0..*n-1] of the
 * reduced simplex and the closest point v.  Returns 1 when the origin is enclosed by the tetrahedron. */
static void simplex_keep(real Y[4][3], real Pa[4][3], int* n, const int* idx, int cnt) {
  real ty[4][3], ta[4][3];
  for (int i = 0; i < cnt; i++) for (int k = 0; k < 3; k++) { ty[i][k] = Y[idx[i]][k]; ta[i][k] = Pa[idx[i]][k]; }
  for (int i = 0; i < cnt; i++) for (int k = 0; k < 3; k++) { Y[i][k] = ty[i][k]; Pa[i][k] = ta[i][k]; }
  *n = cnt;
}
static void closest_segment(real Y[4][3], real Pa[4][3], int* n, real* lam) {
  real ab[3] = {Y[1][0] - Y[0][0], Y[1][1] - Y[0][1], Y[1][2] - Y[0][2]};
  real t = -dot3(Y[0], ab), den = dot3(ab, ab);
  if (t <= 0 || den <= 0) { int id[1] = {0}; simplex_keep(Y, Pa, n, id, 1); lam[0] = 1; return; }
  if (t >= den) { int id[1] = {1}; simplex_keep(Y, Pa, n, id, 1); lam[0] = 1; return; }
  t /= den; lam[0] = 1 - t; lam[1] = t;
}
static void closest_triangle(real Y[4][3], real Pa[4][3], int* n, real* lam) {
  const real *a = Y[0], *b = Y[1], *c = Y[2];
  real ab[3] = {b[0]-a[0], b[1]-a[1], b[2]-a[2]}, ac[3] = {c[0]-a[0], c[1]-a[1], c[2]-a[2]};
  real d1 = -dot3(ab, a), d2 = -dot3(ac, a);
  if (d1 <= 0 && d2 <= 0) { int id[1] = {0}; simplex_keep(Y, Pa, n, id, 1); lam[0] = 1; return; }
  real d3 = -dot3(ab, b), d4 = -dot3(ac, b);
  if (d3 >= 0 && d4 <= d3) { int id[1] = {1}; simplex_keep(Y, Pa, n, id, 1); lam[0] = 1; return; }
  real vc = d1*d4 - d3*d2;
  if (vc <= 0 && d1 >= 0 && d3 <= 0) { real t = d1/(d1 - d3); int id[2] = {0, 1}; simplex_keep(Y, Pa, n, id, 2); lam[0] = 1 - t; lam[1] = t; return; }
  real d5 = -dot3(ab, c), d6 = -dot3(ac, c);
  if (d6 >= 0 && d5 <= d6) { int id[1] = {2}; simplex_keep(Y, Pa, n, id, 1); lam[0] = 1; return; }
  real vb = d5*d2 - d1*d6;
  if (vb <= 0 && d2 >= 0 && d6 <= 0) { real t = d2/(d2 - d6); int id[2] = {0, 2}; simplex_keep(Y, Pa, n, id, 2); lam[0] = 1 - t; lam[1] = t; return; }
  real va = d3*d6 - d5*d4;
  if (va <= 0 && (d4 - d3) >= 0 && (d5 - d6) >= 0) {
    real t = (d4 - d3)/((d4 - d3) + (d5 - d6)); int id[2] = {1, 2}; simplex_keep(Y, Pa, n, id, 2); lam[0] = 1 - t; lam[1] = t; return;
  }
  real den = va + vb + vc;
  if (den <= 0) { int id[2] = {0, 1}; simplex_keep(Y, Pa, n, id, 2); closest_segment(Y, Pa, n, lam); return; }  /* collinear */
  lam[1] = vb/den; lam[2] = vc/den; lam[0] = 1 - lam[1] - lam[2];
}
static void bary_point(real P[4][3], int n, const real* lam, real* out) {
  out[0] = out[1] = out[2] = 0;
  for (int i = 0; i < n; i++) for (int k = 0; k < 3; k++) out[k] += lam[i]*P[i][k];
}
/* returns 1 if the origin is inside the tetrahedron (simplex untouched), else reduces to the closest face feature */
static int closest_tetrahedron(real Y[4][3], real Pa[4][3], int* n, real* lam, real scale2) {
  static const int F[4][4] = {{0, 1, 2, 3}, {0, 2, 3, 1}, {0, 3, 1, 2}, {1, 3, 2, 0}};
  real bestd = -1, bY[4][3], bA[4][3], bl[4]; int bn = 0;
  for (int f = 0; f < 4; f++) {
    const real *a = Y[F[f][0]], *b = Y[F[f][1]], *c = Y[F[f][2]], *dd = Y[F[f][3]];
    real ab[3] = {b[0]-a[0], b[1]-a[1], b[2]-a[2]}, ac[3] = {c[0]-a[0], c[1]-a[1], c[2]-a[2]}, ad[3] = {dd[0]-a[0], dd[1]-a[1], dd[2]-a[2]}, nn[3];
    cross3(nn, ab, ac);
    real sp = -dot3(a, nn), sd = dot3(ad, nn);
    /* origin on the outer side of this face, or the 4th vertex (nearly) in the face's plane: a flat tetrahedron
     * cannot certify enclosure in floating point, so it is always resolved through its faces */
    if (sp*sd < 0 || sd*sd <= (real)1e-10*dot3(nn, nn)*scale2) {
      real tY[4][3], tA[4][3], tl[4], v[3]; int tn = 3;
      for (int i = 0; i < 3; i++) for (int k = 0; k < 3; k++) { tY[i][k] = Y[F[f][i]][k]; tA[i][k] = Pa[F[f][i]][k]; }
      closest_triangle(tY, tA, &tn, tl);
      bary_point(tY, tn, tl, v);
      real d2 = dot3(v, v);
      if (bestd < 0 || d2 < bestd) {
        bestd = d2; bn = tn;
        for (int i = 0; i < tn; i++) { bl[i] = tl[i]; for (int k = 0; k < 3; k++) { bY[i][k] = tY[i][k]; bA[i][k] = tA[i][k]; } }
      }
    }
  }
  if (bestd < 0) return 1;
  for (int i = 0; i < bn; i++) { lam[i] = bl[i]; for (int k = 0; k < 3; k++) { Y[i][k] = bY[i][k]; Pa[i][k] = bA[i][k]; } }
  *n = bn;
  return 0;
}

typedef struct { real dist; real n[3]; real pa[3]; real pb[3]; } cvxres;

static int epa_penetration(const cvx* A, const cvx* B, real Y[4][3], real Pa[4][3], cvxres* out);

/* GJK on the cores.  Returns 0 = farther apart than `cutoff` (core distance), 1 = result valid.
 * Separated: dist > 0, n = direction from A to B, pa / pb = closest core points.  Overlapping cores: EPA. */
static int gjk_epa(const cvx* A, const cvx* B, real cutoff, cvxres* out) {
  real Y[4][3], Pa[4][3], lam[4], v[3], dir[3] = {B->pos[0] - A->pos[0], B->pos[1] - A->pos[1], B->pos[2] - A->pos[2]};
  int n = 0;
  if (dot3(dir, dir) < (real)1e-20) { dir[0] = 1; dir[1] = 0; dir[2] = 0; }
  mink_support(A, B, dir, Y[0], Pa[0]);
  n = 1; lam[0] = 1;
  v[0] = Y[0][0]; v[1] = Y[0][1]; v[2] = Y[0][2];
  real scale2 = dot3(v, v);
  int inside = 0;
  for (int it = 0; it < GJK_MAXIT; it++) {
    real vv = dot3(v, v);
    if (vv <= CVX_TOL*CVX_TOL*scale2 || vv < (real)1e-30) { inside = 1; break; }   /* touching: treat as overlap */
    real w[3], wa[3], nd[3] = {-v[0], -v[1], -v[2]};
    mink_support(A, B, nd, w, wa);
    real vw = dot3(v, w);
    if (vw > 0 && vw*vw > cutoff*cutoff*vv) return 0;          /* separating plane farther than the cutoff */
    real ww = dot3(w, w); if (ww > scale2) scale2 = ww;
    if (vv - vw <= CVX_TOL*vv + (real)1e-12*scale2) break;     /* no progress possible: v is the closest point */
    int dup = 0;
    for (int i = 0; i < n; i++) { real e[3] = {w[0]-Y[i][0], w[1]-Y[i][1], w[2]-Y[i][2]}; if (dot3(e, e) <= (real)1e-14*scale2) dup = 1; }
    if (dup) break;
    for (int k = 0; k < 3; k++) { Y[n][k] = w[k]; Pa[n][k] = wa[k]; }
    n++;
    if (n == 2) closest_segment(Y, Pa, &n, lam);
    else if (n == 3) closest_triangle(Y, Pa, &n, lam);
    else if (closest_tetrahedron(Y, Pa, &n, lam, scale2)) { inside = 1; break; }
    bary_point(Y, n, lam, v);
  }
  if (!inside) {
    real vv = dot3(v, v), dist = (real)sqrt((double)vv);
    if (dist > cutoff) return 0;
    bary_point(Pa, n, lam, out->pa);
    for (int k = 0; k < 3; k++) { out->pb[k] = out->pa[k] - v[k]; out->n[k] = -v[k]/dist; }
    out->dist = dist;
    return 1;
  }
  /* overlap: grow the simplex to a tetrahedron that encloses the origin, then EPA */
  if (n < 4) {
    static const real AX[6][3] = {{1,0,0},{-1,0,0},{0,1,0},{0,-1,0},{0,0,1},{0,0,-1}};
    if (n == 1) {
      for (int a = 0; a < 6 && n == 1; a++) {
        real w[3], wa[3]; mink_support(A, B, AX[a], w, wa);
        real e[3] = {w[0]-Y[0][0], w[1]-Y[0][1], w[2]-Y[0][2]};
        if (dot3(e, e) > (real)1e-12*scale2 + (real)1e-30) { for (int k = 0; k < 3; k++) { Y[1][k] = w[k]; Pa[1][k] = wa[k]; } n = 2; }
      }
    }
    if (n == 2) {
      real ab[3] = {Y[1][0]-Y[0][0], Y[1][1]-Y[0][1], Y[1][2]-Y[0][2]};
      for (int a = 0; a < 6 && n == 2; a++) {
        real dd[3]; cross3(dd, ab, AX[a]);
        if (dot3(dd, dd) < (real)1e-12*dot3(ab, ab)) continue;
        real w[3], wa[3]; mink_support(A, B, dd, w, wa);
        real e[3] = {w[0]-Y[0][0], w[1]-Y[0][1], w[2]-Y[0][2]}, cr[3]; cross3(cr, ab, e);
        if (dot3(cr, cr) > (real)1e-12*scale2*dot3(ab, ab)) { for (int k = 0; k < 3; k++) { Y[2][k] = w[k]; Pa[2][k] = wa[k]; } n = 3; }
      }
    }
    if (n == 3) {
      real ab[3] = {Y[1][0]-Y[0][0], Y[1][1]-Y[0][1], Y[1][2]-Y[0][2]}, ac[3] = {Y[2][0]-Y[0][0], Y[2][1]-Y[0][1], Y[2][2]-Y[0][2]}, nn[3];
      cross3(nn, ab, ac);
      for (int sgn = 0; sgn < 2 && n == 3; sgn++) {
        real dd[3] = {sgn ? -nn[0] : nn[0], sgn ? -nn[1] : nn[1], sgn ? -nn[2] : nn[2]}, w[3], wa[3];
        mink_support(A, B, dd, w, wa);
        real e[3] = {w[0]-Y[0][0], w[1]-Y[0][1], w[2]-Y[0][2]};
        real vol = dot3(e, nn);
        if (vol*vol > (real)1e-12*dot3(nn, nn)*scale2) { for (int k = 0; k < 3; k++) { Y[3][k] = w[k]; Pa[3][k] = wa[k]; } n = 4; }
      }
    }
    if (n < 4) {   /* flat Minkowski difference (e.g. point against a face): zero-depth contact along the last direction */
      real nn[3] = {B->pos[0] - A->pos[0], B->pos[1] - A->pos[1], B->pos[2] - A->pos[2]};
      if (normalize3(nn) < (real)1e-12) { nn[0] = 0; nn[1] = 0; nn[2] = 1; }
      bary_point(Pa, 1, lam, out->pa);
      for (int k = 0; k < 3; k++) { out->pa[k] = Pa[0][k]; out->pb[k] = Pa[0][k] - Y[0][k]; out->n[k] = nn[k]; }
      out->dist = 0;
      return 1;
    }
  }
  return epa_penetration(A, B, Y, Pa, out);
}

/* ---- EPA (expanding polytope): faces are triangles with outward normals; the closest face to the origin is expanded
 * along its normal until the support point adds less than the tolerance.  Deterministic: ties -> lowest face index. */
typedef struct { int v[3]; real n[3]; real d; int alive; } epaface;
static int epa_make_face(epaface* f, real V[][3], int a, int b, int c) {
  real ab[3] = {V[b][0]-V[a][0], V[b][1]-V[a][1], V[b][2]-V[a][2]}, ac[3] = {V[c][0]-V[a][0], V[c][1]-V[a][1], V[c][2]-V[a][2]};
  f->v[0] = a; f->v[1] = b; f->v[2] = c; f->alive = 1;
  cross3(f->n, ab, ac);
  real len = norm3(f->n);
  if (len < (real)1e-20) { f->n[0] = f->n[1] = f->n[2] = 0; f->d = (real)1e30; return 0; }   /* sliver: never chosen */
  f->n[0] /= len; f->n[1] /= len; f->n[2] /= len;
  f->d = dot3(f->n, V[a]);
  return 1;
}
static int epa_penetration(const cvx* A, const cvx* B, real Y[4][3], real Pa[4][3], cvxres* out) {
  real V[EPA_MAXV][3], VA[EPA_MAXV][3];
  epaface F[EPA_MAXF];
  int nv = 4, nf = 0;
  for (int i = 0; i < 4; i++) for (int k = 0; k < 3; k++) { V[i][k] = Y[i][k]; VA[i][k] = Pa[i][k]; }
  /* orient the tetrahedron: faces wound so that normals point away from the 4th vertex */
  {
    real ab[3] = {V[1][0]-V[0][0], V[1][1]-V[0][1], V[1][2]-V[0][2]}, ac[3] = {V[2][0]-V[0][0], V[2][1]-V[0][1], V[2][2]-V[0][2]};
    real ad[3] = {V[3][0]-V[0][0], V[3][1]-V[0][1], V[3][2]-V[0][2]}, nn[3];
    cross3(nn, ab, ac);
    if (dot3(nn, ad) > 0) {   /* swap 1 and 2 so that (0,1,2) faces away from 3 */
      for (int k = 0; k < 3; k++) { real t = V[1][k]; V[1][k] = V[2][k]; V[2][k] = t; t = VA[1][k]; VA[1][k] = VA[2][k]; VA[2][k] = t; }
    }
  }
  epa_make_face(&F[nf++], V, 0, 1, 2);
  epa_make_face(&F[nf++], V, 0, 3, 1);
  epa_make_face(&F[nf++], V, 0, 2, 3);
  epa_make_face(&F[nf++], V, 1, 3, 2);
  int best = 0;
  for (int it = 0; it < EPA_MAXIT; it++) {
    best = -1;
    for (int f = 0; f < nf; f++) if (F[f].alive && (best < 0 || F[f].d < F[best].d)) best = f;
    if (best < 0) return 0;
    real w[3], wa[3];
    mink_support(A, B, F[best].n, w, wa);
    real gain = dot3(w, F[best].n) - F[best].d;
    real tol = CVX_TOL*(1 + (real)fabs((double)F[best].d));
    if (gain <= tol || nv >= EPA_MAXV) break;
    /* faces visible from w die; their boundary is the horizon */
    int E[EPA_MAXF*3][2], ne = 0, nvis = 0;
    for (int f = 0; f < nf; f++) {
      if (!F[f].alive) continue;
      real e[3] = {w[0]-V[F[f].v[0]][0], w[1]-V[F[f].v[0]][1], w[2]-V[F[f].v[0]][2]};
      if (dot3(F[f].n, e) > 0 || F[f].d >= (real)1e30) {
        F[f].alive = 0; nvis++;
        for (int k = 0; k < 3; k++) { E[ne][0] = F[f].v[k]; E[ne][1] = F[f].v[(k + 1)%3]; ne++; }
      }
    }
    if (nvis == 0) break;
    for (int k = 0; k < 3; k++) { V[nv][k] = w[k]; VA[nv][k] = wa[k]; }
    int added = 0, full = 0;
    for (int e = 0; e < ne; e++) {
      int twin = 0;
      for (int g = 0; g < ne; g++) if (E[g][0] == E[e][1] && E[g][1] == E[e][0]) { twin = 1; break; }
      if (twin) continue;
      int slot = -1;
      for (int f = 0; f < nf; f++) if (!F[f].alive) { slot = f; break; }   /* reuse dead slots first */
      if (slot < 0) { if (nf >= EPA_MAXF) { full = 1; break; } slot = nf++; }
      epa_make_face(&F[slot], V, E[e][0], E[e][1], nv);
      added++;
    }
    nv++;
    if (full || added == 0) break;
  }
  /* closest face -> contact: project the origin on it, barycentric weights give the witness point on A */
  best = -1;
  for (int f = 0; f < nf; f++) if (F[f].alive && (best < 0 || F[f].d < F[best].d)) best = f;
  if (best < 0) return 0;
  {
    real T[4][3], TA[4][3], lam[4]; int tn = 3;
    for (int i = 0; i < 3; i++) for (int k = 0; k < 3; k++) { T[i][k] = V[F[best].v[i]][k]; TA[i][k] = VA[F[best].v[i]][k]; }
    closest_triangle(T, TA, &tn, lam);
    real p[3];
    bary_point(T, tn, lam, p);
    bary_point(TA, tn, lam, out->pa);
    real depth = F[best].d > 0 ? F[best].d : 0;
    for (int k = 0; k < 3; k++) { out->n[k] = F[best].n[k]; out->pb[k] = out->pa[k] - p[k]; }
    out->dist = -depth;
  }
  return 1;
}

/* ---- geom -> convex core */
static void cvx_from_geom(const b2s_model_desc* m, const orc_data* d, int g, cvx* s) {
  int t = m->geom_type[g];
  for (int k = 0; k < 3; k++) { s->pos[k] = d->geom_xpos[3*g + k]; s->size[k] = (real)m->geom_size[3*g + k]; }
  for (int k = 0; k < 9; k++) s->mat[k] = d->geom_xmat[9*g + k];
  s->vert = 0; s->vertr = 0; s->nvert = 0; s->radius = 0;
  if (t == B2S_GEOM_SPHERE) { s->type = CVX_POINT; s->radius = s->size[0]; }
  else if (t == B2S_GEOM_CAPSULE) { s->type = CVX_SEGMENT; s->radius = s->size[0]; }
  else if (t == B2S_GEOM_BOX) s->type = CVX_BOX;
  else if (t == B2S_GEOM_CYLINDER) s->type = CVX_CYLINDER;
  else if (t == B2S_GEOM_ELLIPSOID) s->type = CVX_ELLIPSOID;
  else {  /* mesh */
    int id = m->geom_dataid[g];
    s->type = CVX_VERTS; s->vert = m->mesh_vert + 3*m->mesh_vertadr[id]; s->nvert = m->mesh_vertnum[id];
  }
}
/* one contact between two convex geoms (cores + radii); contact normal from A to B */
static int col_convex(rawcon* c, real margin, const cvx* A, const cvx* B) {
  cvxres r;
  if (!gjk_epa(A, B, margin + A->radius + B->radius, &r)) return 0;
  real dist = r.dist - A->radius - B->radius;
  if (dist > margin) return 0;
  c->dist = dist;
  for (int k = 0; k < 3; k++) {
    real sa = r.pa[k] + r.n[k]*A->radius, sb = r.pb[k] - r.n[k]*B->radius;
    c->pos[k] = (real)0.5*(sa + sb); c->frame[k] = r.n[k]; c->frame[3 + k] = 0;
  }
  return 1;
}

/* plane - ellipsoid: the support point against the plane normal (one contact) */
static int col_plane_ellipsoid(rawcon* c, real margin, const real* ppos, const real* pmat, const cvx* E) {
  real n[3] = {pmat[2], pmat[5], pmat[8]}, nn[3] = {-n[0], -n[1], -n[2]}, sp[3];
  cvx_support(E, nn, sp);
  real dif[3] = {sp[0] - ppos[0], sp[1] - ppos[1], sp[2] - ppos[2]};
  real dist = dot3(dif, n);
  if (dist > margin) return 0;
  c->dist = dist;
  for (int k = 0; k < 3; k++) { c->pos[k] = sp[k] - n[k]*(real)0.5*dist; c->frame[k] = n[k]; c->frame[3 + k] = 0; }
  return 1;
}
/* plane - cylinder (own formulation, up to 4 contacts in a fixed order): u = the radial direction that points into the
 * plane; 0: rim point along u of the cap nearer to the plane, 1: the same on the other cap (a cylinder on its side),
 * 2, 3: rim points of the nearer cap at +-120 degrees from u (a cylinder standing on a cap).  Each is kept if its own
 * distance is within the margin.  When the axis is normal to the plane u is the local x axis. */
static int col_plane_cylinder(rawcon* c, real margin, const real* ppos, const real* pmat, const real* cpos, const real* cmat, const real* size) {
  real n[3] = {pmat[2], pmat[5], pmat[8]}, ax[3] = {cmat[2], cmat[5], cmat[8]}, ex[3] = {cmat[0], cmat[3], cmat[6]};
  real r = size[0], h = size[1];
  real an = dot3(ax, n), u[3], ul;
  for (int k = 0; k < 3; k++) u[k] = -(n[k] - an*ax[k]);
  ul = (real)sqrt((double)dot3(u, u));
  if (ul > (real)1e-6) { for (int k = 0; k < 3; k++) u[k] /= ul; } else { for (int k = 0; k < 3; k++) u[k] = ex[k]; }
  real sgn = an > 0 ? (real)-1 : (real)1;      /* cap whose centre is nearer to the plane */
  real v[3]; cross3(v, ax, u);
  const real cs[4] = {1, 1, (real)-0.5, (real)-0.5}, sn[4] = {0, 0, (real)0.8660254037844386, (real)-0.8660254037844386};
  int cnt = 0;
  for (int i = 0; i < 4; i++) {
    real cap = i == 1 ? -sgn : sgn, p[3], dif[3];
    for (int k = 0; k < 3; k++) { p[k] = cpos[k] + cap*h*ax[k] + r*(cs[i]*u[k] + sn[i]*v[k]); dif[k] = p[k] - ppos[k]; }
    real dist = dot3(dif, n);
    if (dist > margin) continue;
    c[cnt].dist = dist;
    for (int k = 0; k < 3; k++) { c[cnt].pos[k] = p[k] - n[k]*(real)0.5*dist; c[cnt].frame[k] = n[k]; c[cnt].frame[3 + k] = 0; }
    cnt++;
  }
  return cnt;
}

/* plane - mesh: vertices within 1 mm of the deepest one form the candidate set; up to 4 of them span the manifold
 * (A deepest, B farthest from A, C farthest from line AB, D farthest from the triangle's other two edges) */
static int col_plane_mesh(rawcon* c, real margin, const real* ppos, const real* pmat, const real* mpos, const real* mmat,
                          const double* vert, int nvert) {
  real n[3] = {pmat[2], pmat[5], pmat[8]}, nl[3], pl[3], dif[3] = {ppos[0]-mpos[0], ppos[1]-mpos[1], ppos[2]-mpos[2]};
  matT_vec(nl, mmat, n);
  matT_vec(pl, mmat, dif);
  real pd = dot3(pl, nl), smax = (real)-1e30;
  int ia = -1;
  for (int i = 0; i < nvert; i++) {
    real sup = pd - ((real)vert[3*i]*nl[0] + (real)vert[3*i+1]*nl[1] + (real)vert[3*i+2]*nl[2]);   /* depth below the plane */
    if (sup > smax) { smax = sup; ia = i; }
  }
  if (ia < 0 || -smax > margin) return 0;
  real thr = smax - (real)1e-3; if (thr < 0 && smax >= 0) thr = 0;
  if (-thr > margin) thr = -margin;
  int idx[4] = {ia, -1, -1, -1};
  real a[3] = {(real)vert[3*ia], (real)vert[3*ia+1], (real)vert[3*ia+2]}, best;
  best = 0;
  for (int i = 0; i < nvert; i++) {
    real v[3] = {(real)vert[3*i], (real)vert[3*i+1], (real)vert[3*i+2]};
    if (pd - dot3(v, nl) < thr) continue;
    real e[3] = {v[0]-a[0], v[1]-a[1], v[2]-a[2]}, q = dot3(e, e);
    if (q > best) { best = q; idx[1] = i; }
  }
  if (idx[1] >= 0 && best > (real)1e-10) {
    real b[3] = {(real)vert[3*idx[1]], (real)vert[3*idx[1]+1], (real)vert[3*idx[1]+2]};
    real ab[3] = {b[0]-a[0], b[1]-a[1], b[2]-a[2]}, side[3];
    cross3(side, nl, ab);
    best = 0;
    for (int i = 0; i < nvert; i++) {
      real v[3] = {(real)vert[3*i], (real)vert[3*i+1], (real)vert[3*i+2]};
      if (pd - dot3(v, nl) < thr) continue;
      real e[3] = {v[0]-a[0], v[1]-a[1], v[2]-a[2]}, q = (real)fabs((double)dot3(e, side));
      if (q > best) { best = q; idx[2] = i; }
    }
    if (idx[2] >= 0 && best > (real)1e-7) {
      real cc[3] = {(real)vert[3*idx[2]], (real)vert[3*idx[2]+1], (real)vert[3*idx[2]+2]};
      real ac[3] = {cc[0]-a[0], cc[1]-a[1], cc[2]-a[2]}, bc[3] = {cc[0]-b[0], cc[1]-b[1], cc[2]-b[2]}, sac[3], sbc[3];
      cross3(sac, nl, ac); cross3(sbc, nl, bc);
      real bestb = 0, besta = 0; int ib = -1, iaa = -1;
      for (int i = 0; i < nvert; i++) {
        real v[3] = {(real)vert[3*i], (real)vert[3*i+1], (real)vert[3*i+2]};
        if (pd - dot3(v, nl) < thr) continue;
        real ea[3] = {v[0]-a[0], v[1]-a[1], v[2]-a[2]}, eb[3] = {v[0]-b[0], v[1]-b[1], v[2]-b[2]};
        if (i == idx[0] || i == idx[1] || i == idx[2]) continue;
        real qb = (real)fabs((double)dot3(eb, sbc)), qa = (real)fabs((double)dot3(ea, sac));   /* off line BC, off line AC */
        if (qb > bestb) { bestb = qb; ib = i; }
        if (qa > besta) { besta = qa; iaa = i; }
      }
      if (bestb >= besta) { best = bestb; idx[3] = ib; } else { best = besta; idx[3] = iaa; }
      if (best <= (real)1e-7) idx[3] = -1;
    } else idx[2] = -1;
  } else idx[1] = -1;
  int nc = 0;
  for (int k = 0; k < 4; k++) {
    int i = idx[k];
    if (i < 0) continue;
    int dup = 0;
    for (int q = 0; q < k; q++) if (idx[q] == i) dup = 1;
    if (dup) continue;
    real v[3] = {(real)vert[3*i], (real)vert[3*i+1], (real)vert[3*i+2]}, wv[3];
    real sup = pd - dot3(v, nl);
    if (-sup > margin) continue;
    mat_vec(wv, mmat, v);
    c[nc].dist = -sup;
    for (int q = 0; q < 3; q++) { c[nc].pos[q] = wv[q] + mpos[q] + n[q]*(sup*(real)0.5); c[nc].frame[q] = n[q]; c[nc].frame[3 + q] = 0; }
    nc++;
  }
  return nc;
}

/* height field (geom1) against a convex geom B: triangular prisms of the cells under B's bounding sphere, visited in
 * (row, column, lower/upper triangle) order; each prism is a 6-vertex convex (top = terrain surface, bottom = the
 * field's base).  One contact per prism; when more than `maxc` prisms touch, the deepest `maxc` are kept (in visiting
 * order).  Cell (r, c) is split along the diagonal (r+1, c) - (r, c+1). */
static int col_hfield(rawcon* c, int maxc, real margin, const b2s_model_desc* m, int hid, const real* hpos, const real* hmat,
                      const cvx* B, real rbound) {
  const double* hs = m->hfield_size + 4*hid;
  const int nrow = m->hfield_nrow[hid], ncol = m->hfield_ncol[hid];
  const double* data = m->hfield_data + m->hfield_adr[hid];
  real sx = (real)hs[0], sy = (real)hs[1], sz = (real)hs[2], sb = (real)hs[3];
  real dif[3] = {B->pos[0]-hpos[0], B->pos[1]-hpos[1], B->pos[2]-hpos[2]}, p[3];
  matT_vec(p, hmat, dif);
  real r = rbound + margin;
  if (p[0] + r < -sx || p[0] - r > sx || p[1] + r < -sy || p[1] - r > sy || p[2] - r > sz || p[2] + r < -sb) return 0;
  real dx = 2*sx/(real)(ncol - 1), dy = 2*sy/(real)(nrow - 1);
  int cmin = (int)floor((double)((p[0] - r + sx)/dx)), cmax = (int)ceil((double)((p[0] + r + sx)/dx));
  int rmin = (int)floor((double)((p[1] - r + sy)/dy)), rmax = (int)ceil((double)((p[1] + r + sy)/dy));
  if (cmin < 0) cmin = 0; if (rmin < 0) rmin = 0;
  if (cmax > ncol - 1) cmax = ncol - 1; if (rmax > nrow - 1) rmax = nrow - 1;
  int nc = 0;
  for (int rr = rmin; rr < rmax; rr++) for (int cc = cmin; cc < cmax; cc++) for (int tri = 0; tri < 2; tri++) {
    /* corner (row, col) offsets of the triangle */
    int cr[3], ck[3];
    if (tri == 0) { cr[0] = rr; ck[0] = cc; cr[1] = rr + 1; ck[1] = cc; cr[2] = rr; ck[2] = cc + 1; }
    else { cr[0] = rr + 1; ck[0] = cc; cr[1] = rr + 1; ck[1] = cc + 1; cr[2] = rr; ck[2] = cc + 1; }
    real loc[6][3], zmax = -sb;
    for (int i = 0; i < 3; i++) {
      real z = (real)data[cr[i]*ncol + ck[i]]*sz;
      loc[i][0] = (real)ck[i]*dx - sx; loc[i][1] = (real)cr[i]*dy - sy; loc[i][2] = z;
      loc[3 + i][0] = loc[i][0]; loc[3 + i][1] = loc[i][1]; loc[3 + i][2] = -sb;
      if (z > zmax) zmax = z;
    }
    if (zmax < p[2] - r) continue;                      /* the whole prism lies below the geom */
    real wv[18];
    for (int i = 0; i < 6; i++) { mat_vec(wv + 3*i, hmat, loc[i]); for (int k = 0; k < 3; k++) wv[3*i + k] += hpos[k]; }
    cvx P; P.type = CVX_VERTS; P.vert = 0; P.vertr = wv; P.nvert = 6; P.radius = 0;
    for (int k = 0; k < 3; k++) P.pos[k] = (wv[k] + wv[3 + k] + wv[6 + k] + wv[9 + k] + wv[12 + k] + wv[15 + k])/6;
    rawcon one;
    if (!col_convex(&one, margin, &P, B)) continue;
    if (nc < maxc) c[nc++] = one;
    else {
      int sh = 0;
      for (int i = 1; i < maxc; i++) if (c[i].dist >= c[sh].dist) sh = i;    /* shallowest kept contact, the latest of equals:
                                                                                 the kept set is the maxc smallest by (dist, visit index) */
      if (one.dist < c[sh].dist) { for (int i = sh; i < maxc - 1; i++) c[i] = c[i + 1]; c[maxc - 1] = one; }
    }
  }
  return nc;
}
#endif
