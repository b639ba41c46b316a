// k_post.cu -- sensors (+ the acceleration half of rne_postconstraint, subtree_vel) and the integrator,
// one warp per world.
#include "b2s_math.cuh"
#include "b2s_api.h"

extern __shared__ __align__(16) float b2s_smem[];

// 6D velocity [ang, lin] of a frame at `pos` riding on `body` (mj_objectVelocity), world orientation
DEV sv6 object_velocity(const ModelDev& m, const float* cvel, const float* com, int body, v3 pos) {
  sv6 cv = ld6(cvel + 6 * body);
  v3 ang = mk3(cv.v[0], cv.v[1], cv.v[2]);
  v3 lin = mk3(cv.v[3], cv.v[4], cv.v[5]) + cross(ang, pos - ld3(com + 3 * m.body_rootid[body]));
  sv6 r; r.v[0] = ang.x; r.v[1] = ang.y; r.v[2] = ang.z; r.v[3] = lin.x; r.v[4] = lin.y; r.v[5] = lin.z;
  return r;
}
// Per-world contact cache for the contact sensors, filled once by the whole warp (one lane per contact): geoms, bodies,
// the pyramid-decoded wrench in the contact frame (mju_decodePyramid) and rotated to the world frame, the position.
// The sensors then match and reduce out of shared memory instead of chasing global loads contact by contact.
#define CC_WORDS 19
// per-warp words of the contact cache: [nconmax][4] ints, [nconmax][15] floats, then the match masks [nsensor][R][2]
__host__ __device__ inline int cc_mask_words(int nconmax) { return (nconmax + 31) / 32; }
__host__ __device__ inline int cc_warp_words(int nconmax, int nsensor) { return CC_WORDS * nconmax + 2 * cc_mask_words(nconmax) * nsensor; }
struct ConCache {
  const int* i;     // [ncon][4]: geom1, geom2, body1, body2
  const float* f;   // [ncon][15]: f6 contact frame | force world 3 | torque world 3 | pos 3
  const unsigned* mm;   // [nsensor][R][2]: per contact sensor, bit c of word r = contact 32 r + c matches forward / backward
  int R;                // words per mask = ceil(nconmax / 32)
};
DEV void fill_contact_cache(const ModelDev& m, const DataDev& d, int w, int ncon, int* ci, float* cf, int lane) {
  const size_t con0 = (size_t)w * d.nconmax, row0 = (size_t)w * d.njmax;
  for (int c = lane; c < ncon; c += 32) {
    const int g1 = d.con_geom[2 * (con0 + c)], g2 = d.con_geom[2 * (con0 + c) + 1];
    ci[4 * c] = g1; ci[4 * c + 1] = g2; ci[4 * c + 2] = m.geom_bodyid[g1]; ci[4 * c + 3] = m.geom_bodyid[g2];
    float f6[6];
#pragma unroll
    for (int k = 0; k < 6; k++) f6[k] = 0.f;
    const int adr = d.con_efc_address[con0 + c], dim = d.con_dim[con0 + c];
    if (adr >= 0) {
      if (dim == 1) f6[0] = d.efc_force[row0 + adr];
      else if (d.efc_type[row0 + adr] == 6) { for (int k = 0; k < dim; k++) f6[k] = d.efc_force[row0 + adr + k]; }      // elliptic: the rows are the frame axes
      else for (int k = 0; k < dim - 1; k++) {
        const float a = d.efc_force[row0 + adr + 2 * k], b = d.efc_force[row0 + adr + 2 * k + 1];
        f6[0] += a + b;
        f6[1 + k] = d.con_friction[5 * (con0 + c) + k] * (a - b);
      }
    }
    const m33 fr = ldm(d.con_frame + 9 * (con0 + c));
    float* o = cf + 15 * c;
#pragma unroll
    for (int k = 0; k < 6; k++) o[k] = f6[k];
    st3(o + 6, mulTv(fr, mk3(f6[0], f6[1], f6[2])));
    st3(o + 9, mulTv(fr, mk3(f6[3], f6[4], f6[5])));
    st3(o + 12, ld3(d.con_pos + 3 * (con0 + c)));
  }
}
DEV bool side_match(const ModelDev& m, int objtype, int objid, int geom, int body) {
  if (objtype == B2S_OBJ_GEOM) return geom == objid;
  if (objtype == B2S_OBJ_BODY) return body == objid;
  if (objtype == B2S_OBJ_XBODY) return m.body_ancmask[(size_t)body * m.nbody + objid] != 0;
  return true;
}
struct Frame { v3 pos; m33 mat; int body; };
DEV Frame obj_frame(const ModelDev& m, const DataDev& d, int w, int type, int id) {
  Frame f;
  if (type == B2S_OBJ_SITE) { const size_t si = b2s_site_index(m, d, w, id); f.pos = ld3(d.site_xpos + si * 3); f.mat = ldm(d.site_xmat + si * 9); f.body = m.site_bodyid[id]; }
  else if (type == B2S_OBJ_GEOM) { f.pos = ld3(d.geom_xpos + ((size_t)w * m.ngeom + id) * 3); f.mat = ldm(d.geom_xmat + ((size_t)w * m.ngeom + id) * 9); f.body = m.geom_bodyid[id]; }
  else if (type == B2S_OBJ_XBODY) { f.pos = ld3(d.xpos + ((size_t)w * m.nbody + id) * 3); f.mat = ldm(d.xmat + ((size_t)w * m.nbody + id) * 9); f.body = id; }
  else { f.pos = ld3(d.xipos + ((size_t)w * m.nbody + id) * 3); f.mat = ldm(d.ximat + ((size_t)w * m.nbody + id) * 9); f.body = id; }
  return f;
}

// Which contacts each contact sensor matches, for the whole warp at once: one lane per contact, one ballot per sensor and
// 32 contacts.  (The per-sensor lanes used to re-evaluate side_match for every contact in every pass of their reduction:
// 35 % of the kernel's samples on 3 active lanes for the G1.)
DEV void match_contacts(const ModelDev& m, int ncon, const int* ci, unsigned* mm, int R, int lane) {
  for (int s = 0; s < m.nsensor; s++) {
    if (m.sensor_type[s] != B2S_SENS_CONTACT) continue;
    const int ot = m.sensor_objtype[s], id = m.sensor_objid[s], rt = m.sensor_reftype[s], rid = m.sensor_refid[s];
    for (int r = 0; r < R; r++) {
      const int c = 32 * r + lane;
      bool fwd = false, bwd = false;
      if (c < ncon) {
        const int g1 = ci[4 * c], g2 = ci[4 * c + 1], b1 = ci[4 * c + 2], b2 = ci[4 * c + 3];
        fwd = side_match(m, ot, id, g1, b1) && (rid < 0 || side_match(m, rt, rid, g2, b2));
        bwd = side_match(m, ot, id, g2, b2) && (rid < 0 || side_match(m, rt, rid, g1, b1));
      }
      const unsigned mf = __ballot_sync(FULL, fwd), mb = __ballot_sync(FULL, bwd);
      if (lane == 0) { mm[(s * R + r) * 2] = mf; mm[(s * R + r) * 2 + 1] = mb; }
    }
  }
}
// iterate the matched contacts of sensor s in contact order: c = contact, fwd = matched in the forward direction
#define FOR_MATCHED(cc, s, ...)                                                          \
  for (int r_ = 0; r_ < (cc).R; r_++) {                                                  \
    const unsigned mf_ = (cc).mm[((s) * (cc).R + r_) * 2];                               \
    unsigned any_ = mf_ | (cc).mm[((s) * (cc).R + r_) * 2 + 1];                          \
    while (any_) {                                                                       \
      const int bit_ = __ffs(any_) - 1; any_ &= any_ - 1;                                \
      const int c = 32 * r_ + bit_; const bool fwd = (mf_ >> bit_) & 1u; (void)fwd;      \
      __VA_ARGS__                                                                        \
    }                                                                                    \
  }

DEV void contact_sensor(const ModelDev& m, const DataDev& d, int w, int s, float* out, int ncon, const ConCache cc) {
  const int bits = m.sensor_intprm[3 * s], reduce = m.sensor_intprm[3 * s + 1], nslot = m.sensor_intprm[3 * s + 2];
  const int dim = m.sensor_dim[s], maxm = m.contact_sensor_maxmatch;
  const size_t con0 = (size_t)w * d.nconmax;
  for (int k = 0; k < dim; k++) out[k] = 0.f;
  if (nslot <= 0) return;
  const int slotdim = dim / nslot;
  // pass 1: count matches (capped), and for netforce the force-weighted centroid
  int nmatch = 0; float wsum = 0.f; v3 P = mk3(0.f, 0.f, 0.f), Pm = mk3(0.f, 0.f, 0.f);
  FOR_MATCHED(cc, s, {
    if (nmatch >= maxm) break;
    nmatch++;
    if (reduce == B2S_REDUCE_NETFORCE) {
      const float* f6 = cc.f + 15 * c;
      const float wt = sqrtf(f6[0] * f6[0] + f6[1] * f6[1] + f6[2] * f6[2]);
      v3 cp = ld3(f6 + 12);
      P = P + cp * wt; Pm = Pm + cp; wsum += wt;
    }
  })
  (void)ncon;
  if (nmatch == 0) return;
  if (reduce == B2S_REDUCE_NETFORCE) {
    P = wsum > B2S_MINVAL ? P * (1.f / wsum) : Pm * (1.f / (float)nmatch);
    v3 F = mk3(0.f, 0.f, 0.f), Tq = mk3(0.f, 0.f, 0.f);
    int seen = 0;
    FOR_MATCHED(cc, s, {
      if (seen >= maxm) break;
      seen++;
      const float dir = fwd ? 1.f : -1.f;
      const float* f = cc.f + 15 * c;
      v3 fg = ld3(f + 6) * dir, tg = ld3(f + 9) * dir;
      v3 arm = ld3(f + 12) - P;
      F = F + fg; Tq = Tq + tg + cross(arm, fg);
    })
    float* o = out;
    if (bits & B2S_CON_FOUND) *o++ = (float)nmatch;
    if (bits & B2S_CON_FORCE) { st3(o, F); o += 3; }
    if (bits & B2S_CON_TORQUE) { st3(o, Tq); o += 3; }
    if (bits & B2S_CON_DIST) *o++ = 0.f;
    if (bits & B2S_CON_POS) { st3(o, P); o += 3; }
    if (bits & B2S_CON_NORMAL) { st3(o, mk3(1.f, 0.f, 0.f)); o += 3; }
    if (bits & B2S_CON_TANGENT) { st3(o, mk3(0.f, 1.f, 0.f)); o += 3; }
    return;
  }
  // slots: selection by repeated scan in (key, match-order) order -- no per-lane match list needed
  float pkey = -3.0e38f; int pord = -1;
  for (int sl = 0; sl < nslot && sl < nmatch; sl++) {
    float bkey = 3.0e38f; int bord = -1, bc = -1; float bdir = 1.f;
    int ord = 0;
    FOR_MATCHED(cc, s, {
      if (ord >= maxm) break;
      float key;
      if (reduce == B2S_REDUCE_NONE) key = (float)ord;
      else if (reduce == B2S_REDUCE_MINDIST) key = d.con_dist[con0 + c];
      else { const float* f6 = cc.f + 15 * c; key = -sqrtf(f6[0] * f6[0] + f6[1] * f6[1] + f6[2] * f6[2]); }
      const bool after = key > pkey || (key == pkey && ord > pord);
      const bool better = key < bkey || (key == bkey && ord < bord);
      if (after && (bord < 0 || better)) { bkey = key; bord = ord; bc = c; bdir = fwd ? 1.f : -1.f; }
      ord++;
    })
    if (bc < 0) break;
    pkey = bkey; pord = bord;
    const float* f6 = cc.f + 15 * bc;
    const float* fr = d.con_frame + 9 * (con0 + bc);
    float* o = out + sl * slotdim;
    if (bits & B2S_CON_FOUND) *o++ = (float)nmatch;
    if (bits & B2S_CON_FORCE) { o[0] = f6[0]; o[1] = bdir * f6[1]; o[2] = bdir * f6[2]; o += 3; }
    if (bits & B2S_CON_TORQUE) { o[0] = f6[3]; o[1] = bdir * f6[4]; o[2] = bdir * f6[5]; o += 3; }
    if (bits & B2S_CON_DIST) *o++ = d.con_dist[con0 + bc];
    if (bits & B2S_CON_POS) { st3(o, ld3(f6 + 12)); o += 3; }
    if (bits & B2S_CON_NORMAL) { st3(o, ld3(fr) * bdir); o += 3; }
    if (bits & B2S_CON_TANGENT) { st3(o, ld3(fr + 3) * bdir); o += 3; }
  }
}

// ---------------------------------------------------------------------------------------------- rays (rangefinder)
// Distance along a ray to the surface of one primitive geom, -1 without intersection (MuJoCo engine_ray.c mj_rayGeom [ext]:
// the ray is taken into the geom frame; quadratics return the smaller non-negative root, so a ray that starts inside a solid
// reports its exit point; planes only count from their front side and inside their rendered rectangle).
DEV float ray_quad(float a, float b, float c) {
  const float det = b * b - a * c;
  if (det < B2S_MINVAL) return -1.f;
  const float sq = sqrtf(det);
  const float x0 = (-b - sq) / a, x1 = (-b + sq) / a;
  return x0 >= 0.f ? x0 : (x1 >= 0.f ? x1 : -1.f);
}
DEV float ray_geom(int type, const float* size, v3 gpos, const m33& gmat, v3 pnt, v3 vec) {
  const v3 p = mulTv(gmat, pnt - gpos), v = mulTv(gmat, vec);
  float x = -1.f;
  switch (type) {
    case B2S_GEOM_PLANE: {
      if (v.z > -B2S_MINVAL) return -1.f;
      const float t = -p.z / v.z;
      if (t < 0.f) return -1.f;
      const float px = p.x + t * v.x, py = p.y + t * v.y;
      if ((size[0] <= 0.f || fabsf(px) <= size[0]) && (size[1] <= 0.f || fabsf(py) <= size[1])) x = t;
    } break;
    case B2S_GEOM_SPHERE: x = ray_quad(dot(v, v), dot(v, p), dot(p, p) - size[0] * size[0]); break;
    case B2S_GEOM_CAPSULE: {
      const float r = size[0], h = size[1];
      float t = ray_quad(v.x * v.x + v.y * v.y, v.x * p.x + v.y * p.y, p.x * p.x + p.y * p.y - r * r);
      if (t >= 0.f && fabsf(p.z + t * v.z) <= h) x = t;
#pragma unroll
      for (int side = -1; side <= 1; side += 2) {
        const v3 q = mk3(p.x, p.y, p.z - side * h);
        t = ray_quad(dot(v, v), dot(v, q), dot(q, q) - r * r);
        if (t >= 0.f && (x < 0.f || t < x) && side * (p.z + t * v.z) >= h) x = t;
      }
    } break;
    case B2S_GEOM_ELLIPSOID: {
      const float s0 = 1.f / (size[0] * size[0]), s1 = 1.f / (size[1] * size[1]), s2 = 1.f / (size[2] * size[2]);
      x = ray_quad(s0 * v.x * v.x + s1 * v.y * v.y + s2 * v.z * v.z, s0 * p.x * v.x + s1 * p.y * v.y + s2 * p.z * v.z,
                   s0 * p.x * p.x + s1 * p.y * p.y + s2 * p.z * p.z - 1.f);
    } break;
    case B2S_GEOM_CYLINDER: {
      const float r = size[0], h = size[1];
      if (fabsf(v.z) > B2S_MINVAL) {
#pragma unroll
        for (int side = -1; side <= 1; side += 2) {
          const float t = (side * h - p.z) / v.z;
          if (t >= 0.f && (x < 0.f || t < x)) {
            const float px = p.x + t * v.x, py = p.y + t * v.y;
            if (px * px + py * py <= r * r) x = t;
          }
        }
      }
      const float t = ray_quad(v.x * v.x + v.y * v.y, v.x * p.x + v.y * p.y, p.x * p.x + p.y * p.y - r * r);
      if (t >= 0.f && (x < 0.f || t < x) && fabsf(p.z + t * v.z) <= h) x = t;
    } break;
    case B2S_GEOM_BOX: {
      const float pp[3] = {p.x, p.y, p.z}, vv[3] = {v.x, v.y, v.z};
#pragma unroll
      for (int i = 0; i < 3; i++) {
        if (fabsf(vv[i]) <= B2S_MINVAL) continue;
#pragma unroll
        for (int side = -1; side <= 1; side += 2) {
          const float t = (side * size[i] - pp[i]) / vv[i];
          if (t >= 0.f && (x < 0.f || t < x)) {
            const int j = (i + 1) % 3, k = (i + 2) % 3;
            if (fabsf(pp[j] + t * vv[j]) <= size[j] && fabsf(pp[k] + t * vv[k]) <= size[k]) x = t;
          }
        }
      }
    } break;
    default: break;
  }
  return x;
}

// shared per warp: cacc 6nb | bvel 6nb | lm 3nb | am 3nb | slv 3nb | dof 6nv | (force / torque sensors only) cfrc 6nb
__host__ __device__ inline int sens_smem_floats(int nb, int nv, int need_cfrc) { return 21 * nb + 6 * nv + (need_cfrc ? 6 * nb : 0); }
__global__ void __launch_bounds__(32 * B2S_WARPS_PER_CTA) k_sensors(const ModelDev m, const DataDev d) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int w = blockIdx.x * B2S_WARPS_PER_CTA + warp;
  if (w >= d.nworld) return;
  const int nb = m.nbody, nv = m.nv;
  const int wfl = sens_smem_floats(nb, nv, m.need_cfrc);
  float* s_cacc = b2s_smem + (size_t)warp * wfl;
  float* s_bvel = s_cacc + 6 * nb;
  float* s_lm = s_bvel + 6 * nb;
  float* s_am = s_lm + 3 * nb;
  float* s_slv = s_am + 3 * nb;
  float* s_dof = s_slv + 3 * nb;
  float* s_cfi = s_dof + 6 * nv;
  const float* qvel = d.snap_qvel + (size_t)w * nv;     // the state this forward pass was computed from (see DataDev::snap_*)
  // contact cache (only when the model has contact sensors): per-warp [nconmax][4] ints then [nconmax][15] floats
  int* s_ci = (int*)(b2s_smem + (size_t)B2S_WARPS_PER_CTA * wfl) + (size_t)warp * cc_warp_words(d.nconmax, m.nsensor);
  float* s_cf = (float*)(s_ci + 4 * d.nconmax);
  unsigned* s_mm = (unsigned*)(s_ci + CC_WORDS * d.nconmax);
  const int ncon = m.need_concache ? d.ncon[w] : 0;
  if (m.need_concache) {
    fill_contact_cache(m, d, w, ncon, s_ci, s_cf, lane);
    __syncwarp();
    match_contacts(m, ncon, s_ci, s_mm, cc_mask_words(d.nconmax), lane);
  }
  const float* qacc = d.qacc + (size_t)w * nv;
  const float* cdof = d.cdof + (size_t)w * 6 * nv;
  const float* cdd = d.cdof_dot + (size_t)w * 6 * nv;
  const float* cvel = d.cvel + (size_t)w * 6 * nb;
  const float* com = d.subtree_com + (size_t)w * 3 * nb;
  const float* xipos = d.xipos + (size_t)w * 3 * nb;
  // cacc = cacc_world + sum over ancestor dofs (cdof_dot qvel + cdof qacc)
  const bool grav = !(m.disableflags & B2S_DSBL_GRAVITY);
  // per-dof term cdof_dot qvel + cdof qacc, once per dof (coalesced global reads), summed per body from shared memory
  for (int i = lane; i < 6 * nv; i += 32) { const int j = i / 6; s_dof[i] = cdd[i] * qvel[j] + cdof[i] * qacc[j]; }
  __syncwarp();
  for (int b = lane; b < nb; b += 32) {
    sv6 a = zero6();
    if (grav) { a.v[3] = -m.gravity[0]; a.v[4] = -m.gravity[1]; a.v[5] = -m.gravity[2]; }
    for (int c = m.body_chain_adr[b]; c < m.body_chain_adr[b + 1]; c++) axpy6(a, 1.0f, ld6(s_dof + 6 * m.body_chain[c]));
    st6(s_cacc + 6 * b, a);
    st6(d.cacc + ((size_t)w * nb + b) * 6, a);
  }
  if (m.need_subtreevel) {
    const float* mass = m.body_mass.at(w);
    const float* inertia = m.body_inertia.at(w);
    for (int b = lane; b < nb; b += 32) {
      const v3 xi = ld3(xipos + 3 * b);
      sv6 v = object_velocity(m, cvel, com, b, xi);
      // s_bvel: [xipos | linear velocity] -- the angular part is only needed for s_am below, and the angular-momentum loop wants
      // the inertial-frame positions of the whole subtree from shared memory (it read xipos and mass from global memory per term)
      st3(s_bvel + 6 * b, xi); st3(s_bvel + 6 * b + 3, mk3(v.v[3], v.v[4], v.v[5]));
      st3(s_lm + 3 * b, mk3(v.v[3], v.v[4], v.v[5]) * mass[b]);
      m33 R = ldm(d.ximat + ((size_t)w * nb + b) * 9);
      v3 loc = mulTv(R, mk3(v.v[0], v.v[1], v.v[2]));
      loc = mk3(loc.x * inertia[3 * b], loc.y * inertia[3 * b + 1], loc.z * inertia[3 * b + 2]);
      st3(s_am + 3 * b, mulv(R, loc));
    }
    __syncwarp();
    for (int b = lane; b < nb; b += 32) {
      v3 acc = mk3(0.f, 0.f, 0.f);
      const int e = m.body_subtree_end[b];
      for (int j = b; j < e; j++) acc = acc + ld3(s_lm + 3 * j);
      acc = acc * (1.0f / fmaxf(m.body_subtreemass[b], B2S_MINVAL));
      st3(s_slv + 3 * b, acc);
      st3(d.subtree_linvel + ((size_t)w * nb + b) * 3, acc);
    }
    __syncwarp();
    for (int b = lane; b < nb; b += 32) s_lm[3 * b] = mass[b];      // the linear momenta are consumed: their slots carry the masses now
    __syncwarp();
    for (int b = lane; b < nb; b += 32) {
      // angular momentum of the subtree about its own COM (closed form of mj_subtreeVel's recursion)
      v3 L = mk3(0.f, 0.f, 0.f);
      const v3 cb = ld3(com + 3 * b), Vb = ld3(s_slv + 3 * b);
      const int e = m.body_subtree_end[b];
#pragma unroll 2
      for (int j = b; j < e; j++) {
        v3 dx = ld3(s_bvel + 6 * j) - cb;
        v3 dv = (ld3(s_bvel + 6 * j + 3) - Vb) * s_lm[3 * j];
        L = L + ld3(s_am + 3 * j) + cross(dx, dv);
      }
      st3(d.subtree_angmom + ((size_t)w * nb + b) * 3, L);
    }
  }
  if (m.need_cfrc) {
    // force half of mj_rnePostConstraint: per body, inertial force cinert cacc + cvel x* (cinert cvel) minus the external
    // wrenches on it (xfrc_applied at xipos; contact forces at the contact point, -f on geom1's body and +f on geom2's), as
    // spatial forces [torque, force] about the subtree COM of the body's root.  A sensor sums its body's subtree.
    __syncwarp();
    for (int b = lane; b < nb; b += 32) {
      sv6 f = zero6();
      if (b > 0) {
        const float* I = d.cinert + ((size_t)w * nb + b) * 10;
        const sv6 v = ld6(cvel + 6 * b);
        f = mul_inert(I, ld6(s_cacc + 6 * b));
        axpy6(f, 1.0f, cross_force(v, mul_inert(I, v)));
        const v3 cr = ld3(com + 3 * m.body_rootid[b]);
        const float* x = d.xfrc_applied + ((size_t)w * nb + b) * 6;
        const v3 frc = ld3(x), trq = ld3(x + 3);
        if (frc.x != 0.f || frc.y != 0.f || frc.z != 0.f || trq.x != 0.f || trq.y != 0.f || trq.z != 0.f) {
          const v3 t = trq - cross(cr - ld3(xipos + 3 * b), frc);
          f.v[0] -= t.x; f.v[1] -= t.y; f.v[2] -= t.z; f.v[3] -= frc.x; f.v[4] -= frc.y; f.v[5] -= frc.z;
        }
        for (int c = 0; c < ncon; c++) {
          const int b1 = s_ci[4 * c + 2], b2 = s_ci[4 * c + 3];
          if (b1 != b && b2 != b) continue;
          if (d.con_efc_address[(size_t)w * d.nconmax + c] < 0) continue;
          const float* cf = s_cf + 15 * c;
          const v3 fw = ld3(cf + 6), tw = ld3(cf + 9);
          const v3 t = tw - cross(cr - ld3(cf + 12), fw);
          const float sg = (b2 == b ? 1.f : 0.f) - (b1 == b ? 1.f : 0.f);
          f.v[0] -= sg * t.x; f.v[1] -= sg * t.y; f.v[2] -= sg * t.z; f.v[3] -= sg * fw.x; f.v[4] -= sg * fw.y; f.v[5] -= sg * fw.z;
        }
      }
      st6(s_cfi + 6 * b, f);
    }
  }
  __syncwarp();
  // one lane per sensor
  for (int s = lane; s < m.nsensor; s += 32) {
    const int type = m.sensor_type[s], id = m.sensor_objid[s], ot = m.sensor_objtype[s], dim = m.sensor_dim[s];
    B2S_ASSERT(m.sensor_adr[s] >= 0 && m.sensor_adr[s] + dim <= m.nsensordata && ncon <= d.nconmax);
    float* out = d.sensordata + (size_t)w * m.nsensordata + m.sensor_adr[s];
    switch (type) {
      case B2S_SENS_GYRO: case B2S_SENS_VELOCIMETER: case B2S_SENS_ACCELEROMETER: {
        Frame f = obj_frame(m, d, w, B2S_OBJ_SITE, id);
        sv6 v = object_velocity(m, cvel, com, f.body, f.pos);
        v3 wl = mulTv(f.mat, mk3(v.v[0], v.v[1], v.v[2])), vl = mulTv(f.mat, mk3(v.v[3], v.v[4], v.v[5]));
        if (type == B2S_SENS_GYRO) st3(out, wl);
        else if (type == B2S_SENS_VELOCIMETER) st3(out, vl);
        else {
          const float* ca = s_cacc + 6 * f.body;
          v3 aang = mk3(ca[0], ca[1], ca[2]);
          v3 alin = mk3(ca[3], ca[4], ca[5]) + cross(aang, f.pos - ld3(com + 3 * m.body_rootid[f.body]));
          st3(out, mulTv(f.mat, alin) + cross(wl, vl));
        }
      } break;
      case B2S_SENS_FORCE: case B2S_SENS_TORQUE: {
        // interaction force between the site's body and its parent (cfrc_int = sum over the body's subtree), moved from
        // the root's subtree COM to the site and expressed in the site frame
        Frame f = obj_frame(m, d, w, B2S_OBJ_SITE, id);
        sv6 ci = zero6();
        for (int j = f.body; j < m.body_subtree_end[f.body]; j++) axpy6(ci, 1.0f, ld6(s_cfi + 6 * j));
        const v3 frc = mk3(ci.v[3], ci.v[4], ci.v[5]);
        const v3 trq = mk3(ci.v[0], ci.v[1], ci.v[2]) - cross(f.pos - ld3(com + 3 * m.body_rootid[f.body]), frc);
        st3(out, mulTv(f.mat, type == B2S_SENS_FORCE ? frc : trq));
      } break;
      case B2S_SENS_JOINTPOS: out[0] = d.snap_qpos[(size_t)w * m.nq + m.jnt_qposadr[id]]; break;
      case B2S_SENS_JOINTVEL: out[0] = qvel[m.jnt_dofadr[id]]; break;
      case B2S_SENS_ACTUATORFRC: out[0] = d.actuator_force[(size_t)w * m.nu + id]; break;
      case B2S_SENS_TENDONPOS: out[0] = d.ten_length[(size_t)w * m.ntendon + id]; break;
      case B2S_SENS_TENDONVEL: out[0] = d.ten_velocity[(size_t)w * m.ntendon + id]; break;
      case B2S_SENS_ACTUATORPOS: out[0] = d.actuator_length[(size_t)w * m.nu + id]; break;
      case B2S_SENS_ACTUATORVEL: out[0] = d.actuator_velocity[(size_t)w * m.nu + id]; break;
      case B2S_SENS_JOINTACTFRC: out[0] = d.qfrc_actuator[(size_t)w * nv + m.jnt_dofadr[id]]; break;
      case B2S_SENS_CLOCK: out[0] = d.snap_time[w]; break;
      case B2S_SENS_MAGNETOMETER: {
        const Frame f = obj_frame(m, d, w, B2S_OBJ_SITE, id);
        st3(out, mulTv(f.mat, mk3(m.magnetic[0], m.magnetic[1], m.magnetic[2])));
      } break;
      case B2S_SENS_JOINTLIMITPOS: case B2S_SENS_JOINTLIMITVEL: case B2S_SENS_JOINTLIMITFRC: {
        // the joint's limit row if one is active (rows nf .. nl-1 are the limits)
        const size_t row0 = (size_t)w * d.njmax;
        float v = 0.f;
        for (int r = d.nf[w]; r < d.nl[w]; r++)
          if (d.efc_id[row0 + r] == id) {
            v = type == B2S_SENS_JOINTLIMITPOS ? d.efc_pos[row0 + r] - d.efc_margin[row0 + r] : (type == B2S_SENS_JOINTLIMITVEL ? d.efc_vel[row0 + r] : d.efc_force[row0 + r]);
            break;
          }
        out[0] = v;
      } break;
      case B2S_SENS_TENDONACTFRC: {
        float f = 0.f;
        for (int u = 0; u < m.nu; u++) if (m.actuator_trntype[u] == B2S_TRN_TENDON && m.actuator_trnid[u] == id) f += d.actuator_force[(size_t)w * m.nu + u];
        out[0] = f;
      } break;
      case B2S_SENS_E_POTENTIAL: {
        float e = 0.f;
        const float* mass = m.body_mass.at(w);
        if (grav) for (int i = 1; i < nb; i++) e -= mass[i] * (m.gravity[0] * xipos[3 * i] + m.gravity[1] * xipos[3 * i + 1] + m.gravity[2] * xipos[3 * i + 2]);
        if (!(m.disableflags & B2S_DSBL_PASSIVE)) {
          const float* stiff = m.jnt_stiffness.at(w);
          const float* qspring = m.qpos_spring.at(w);
          for (int j = 0; j < m.njnt; j++) {
            const float k = stiff[j];
            if (k == 0.f || (m.jnt_type[j] != B2S_JNT_HINGE && m.jnt_type[j] != B2S_JNT_SLIDE)) continue;
            const float dq = d.snap_qpos[(size_t)w * m.nq + m.jnt_qposadr[j]] - qspring[m.jnt_qposadr[j]];
            e += 0.5f * k * dq * dq;
          }
          for (int t = 0; t < m.ntendon; t++) {
            const float k = m.tendon_stiffness[t], len = d.ten_length[(size_t)w * m.ntendon + t], lo = m.tendon_lengthspring[2 * t], hi = m.tendon_lengthspring[2 * t + 1];
            const float dl = len > hi ? len - hi : (len < lo ? len - lo : 0.f);
            e += 0.5f * k * dl * dl;
          }
        }
        out[0] = e;
      } break;
      case B2S_SENS_E_KINETIC: {
        const float* M = d.qM + (size_t)w * nv * m.nvp;
        float e = 0.f;
        for (int i = 0; i < nv; i++) { float r = 0.f; for (int j = 0; j < nv; j++) r += M[i * m.nvp + j] * qvel[j]; e += qvel[i] * r; }
        out[0] = 0.5f * e;
      } break;
      case B2S_SENS_FRAMELINACC: case B2S_SENS_FRAMEANGACC: {
        Frame f = obj_frame(m, d, w, ot, id);
        sv6 v = object_velocity(m, cvel, com, f.body, f.pos);
        const float* ca = s_cacc + 6 * f.body;
        const v3 aang = mk3(ca[0], ca[1], ca[2]);
        if (type == B2S_SENS_FRAMEANGACC) st3(out, aang);
        else st3(out, mk3(ca[3], ca[4], ca[5]) + cross(aang, f.pos - ld3(com + 3 * m.body_rootid[f.body])) +
                      cross(mk3(v.v[0], v.v[1], v.v[2]), mk3(v.v[3], v.v[4], v.v[5])));
      } break;
      case B2S_SENS_SUBTREECOM: st3(out, ld3(com + 3 * id)); break;
      case B2S_SENS_SUBTREELINVEL: st3(out, ld3(s_slv + 3 * id)); break;
      case B2S_SENS_SUBTREEANGMOM: st3(out, ld3(d.subtree_angmom + ((size_t)w * nb + id) * 3)); break;
      case B2S_SENS_FRAMEPOS: case B2S_SENS_FRAMEQUAT: case B2S_SENS_FRAMEXAXIS: case B2S_SENS_FRAMEYAXIS: case B2S_SENS_FRAMEZAXIS: {
        Frame f = obj_frame(m, d, w, ot, id);
        const bool hasref = m.sensor_refid[s] >= 0;
        Frame r = f;
        if (hasref) r = obj_frame(m, d, w, m.sensor_reftype[s], m.sensor_refid[s]);
        if (type == B2S_SENS_FRAMEPOS) st3(out, hasref ? mulTv(r.mat, f.pos - r.pos) : f.pos);
        else if (type != B2S_SENS_FRAMEQUAT) { const int ax = type - B2S_SENS_FRAMEXAXIS; st3(out, hasref ? mulTv(r.mat, col(f.mat, ax)) : col(f.mat, ax)); }
        else {
          m33 R = f.mat;
          if (hasref) for (int a = 0; a < 3; a++) for (int b = 0; b < 3; b++) {
            float acc = 0.f;
            for (int k = 0; k < 3; k++) acc += r.mat.m[3 * k + a] * f.mat.m[3 * k + b];
            R.m[3 * a + b] = acc;
          }
          stq(out, m2q(R));
        }
      } break;
      case B2S_SENS_FRAMELINVEL: case B2S_SENS_FRAMEANGVEL: {
        Frame f = obj_frame(m, d, w, ot, id);
        sv6 v = object_velocity(m, cvel, com, f.body, f.pos);
        if (type == B2S_SENS_FRAMELINVEL) st3(out, mk3(v.v[3], v.v[4], v.v[5])); else st3(out, mk3(v.v[0], v.v[1], v.v[2]));
      } break;
      case B2S_SENS_RANGEFINDER: break;      // written by the warp-wide ray pass below
      case B2S_SENS_CONTACT: { ConCache cc; cc.i = s_ci; cc.f = s_cf; cc.mm = s_mm; cc.R = cc_mask_words(d.nconmax); contact_sensor(m, d, w, s, out, ncon, cc); } break;
      default: for (int k = 0; k < dim; k++) out[k] = 0.f;
    }
    const float cut = m.sensor_cutoff[s];
    if (cut > 0.f && type != B2S_SENS_CONTACT && type != B2S_SENS_RANGEFINDER) for (int k = 0; k < dim; k++) out[k] = clampf(out[k], -cut, cut);
  }
  if (m.need_ray) {
    // rangefinders (mj_ray from the site along its z axis; geoms of the site's own body are skipped): the warp scans the geoms
    // together, nearest hit by a min-reduction; -1 without a hit; a cutoff caps hits only (MuJoCo's "positive" data type)
    const float* gx = d.geom_xpos + (size_t)w * m.ngeom * 3;
    const float* gm = d.geom_xmat + (size_t)w * m.ngeom * 9;
    for (int s = 0; s < m.nsensor; s++) {
      if (m.sensor_type[s] != B2S_SENS_RANGEFINDER) continue;
      const Frame f = obj_frame(m, d, w, B2S_OBJ_SITE, m.sensor_objid[s]);
      const v3 dir = mk3(f.mat.m[2], f.mat.m[5], f.mat.m[8]);
      float best = 3.0e38f;
      for (int g = lane; g < m.ngeom; g += 32) {
        if (m.geom_bodyid[g] == f.body) continue;
        const float x = ray_geom(m.geom_type[g], m.geom_size + 3 * g, ld3(gx + 3 * g), ldm(gm + 9 * g), f.pos, dir);
        if (x >= 0.f && x < best) best = x;
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) best = fminf(best, __shfl_xor_sync(FULL, best, o));
      if (lane == 0) {
        float r = best < 1.0e38f ? best : -1.f;
        const float cut = m.sensor_cutoff[s];
        if (cut > 0.f && r > cut) r = cut;
        d.sensordata[(size_t)w * m.nsensordata + m.sensor_adr[s]] = r;
      }
    }
  }
}

// ------------------------------------------------------------------------------------------- k_integrate
// shared per warp: Cholesky scratch line 2*RC_LINE | A hdim*ld | invd 7T | dg nvp | qa nvp
__host__ __device__ inline int integ_smem_floats(int T, int nvp) { return (2 * RC_LINE + b2s_hdim(T) * b2s_hld(T) + 7 * T + 2 * nvp + 3) & ~3; }
template <int T>
__global__ void __launch_bounds__(32 * B2S_WARPS_PER_CTA) k_integrate(const ModelDev m, const DataDev d) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int w = blockIdx.x * B2S_WARPS_PER_CTA + warp;
  if (w >= d.nworld) return;
  const int nv = m.nv, nvp = m.nvp, ld = b2s_hld(T);
  float* s_col = b2s_smem + (size_t)warp * integ_smem_floats(T, nvp);
  float* A = s_col + 2 * RC_LINE;
  float* invd = A + b2s_hdim(T) * ld;
  float* dg = invd + 7 * T;
  float* qa = dg + nvp;
  const float dt = m.timestep;
  float* qvel = d.qvel + (size_t)w * nv;
  float* qpos = d.qpos + (size_t)w * m.nq;
  const float* qacc = d.qacc + (size_t)w * nv;
  // The world's mass matrix is the only large input and comes from L2 / HBM: its loads are issued FIRST, as one coalesced
  // stream (float4 k of the matrix on lane k mod 32: 512 contiguous bytes per warp instruction), and are consumed after the
  // derivative terms below have been computed (ncu: the old one-row-per-lane read stalled 36 % of the kernel's samples).
  constexpr int MQ = (7 * T * ((7 * T + 3) / 4) + 31) / 32;      // float4 per lane
  const int n4 = nvp >> 2, nm4 = nv * n4;
  const bool implicit = m.integrator == B2S_INT_IMPLICITFAST || (m.integrator == B2S_INT_EULER && !(m.disableflags & (B2S_DSBL_EULERDAMP | B2S_DSBL_PASSIVE)));
  float4 mv[MQ];
  if (implicit) {
    const float4* M4 = reinterpret_cast<const float4*>(d.qM + (size_t)w * nv * nvp);
#pragma unroll
    for (int q = 0; q < MQ; q++) mv[q] = M4[min(lane + 32 * q, nm4 - 1)];
  }
  for (int i = lane; i < nvp; i += 32) { dg[i] = 0.f; qa[i] = i < nv ? qacc[i] : 0.f; }
  __syncwarp();
  bool use = false;
  if (m.integrator == B2S_INT_IMPLICITFAST) {
    use = true;
    if (!(m.disableflags & B2S_DSBL_ACTUATION)) {
      const float* gainprm = m.actuator_gainprm.at(w);
      const float* biasprm = m.actuator_biasprm.at(w);
      for (int u = lane; u < m.nu; u += 32) {
        float bv = 0.f;
        if (m.actuator_biastype[u] == B2S_BIAS_AFFINE) bv += biasprm[3 * u + 2];
        if (m.actuator_gaintype[u] == B2S_GAIN_AFFINE) {
          float c = d.ctrl[(size_t)w * m.nu + u];
          if (m.actuator_ctrllimited[u] && !(m.disableflags & B2S_DSBL_CLAMPCTRL)) c = clampf(c, m.actuator_ctrlrange[2 * u], m.actuator_ctrlrange[2 * u + 1]);
          bv += gainprm[3 * u + 2] * c;
        }
        if (bv == 0.f) continue;
        if (m.actuator_forcelimited[u]) {
          const float f = d.actuator_force[(size_t)w * m.nu + u];
          if (f <= m.actuator_forcerange[2 * u] || f >= m.actuator_forcerange[2 * u + 1]) continue;
        }
        const float g = m.actuator_gear[u];
        if (m.actuator_trntype[u] == B2S_TRN_TENDON) continue;     // rank-1 term, added to A below
        atomicAdd(&dg[m.jnt_dofadr[m.actuator_trnid[u]]], g * g * bv);
      }
    }
    __syncwarp();
    if (!(m.disableflags & B2S_DSBL_PASSIVE)) {
      const float* damping = m.dof_damping.at(w);
      for (int i = lane; i < nv; i += 32) dg[i] -= damping[i];
    }
  } else if (m.integrator == B2S_INT_EULER && !(m.disableflags & (B2S_DSBL_EULERDAMP | B2S_DSBL_PASSIVE))) {
    const float* damping = m.dof_damping.at(w);
    bool any = false;
    for (int i = lane; i < nv; i += 32) if (damping[i] > 0.f) { any = true; dg[i] = -damping[i]; }
    use = __any_sync(FULL, any);
  }
  __syncwarp();
  if (use) {
    // A = lower(M) - dt diag(dg), zero strict upper triangle (the substitution needs it)
#pragma unroll
    for (int q = 0; q < MQ; q++) {
      const int idx = lane + 32 * q;
      if (idx < nm4) {
        const int r = idx / n4, c0 = 4 * (idx - r * n4);
        const float e[4] = {mv[q].x, mv[q].y, mv[q].z, mv[q].w};
        const float dd = dt * dg[r];
        float* Ar = A + r * ld;
#pragma unroll
        for (int k = 0; k < 4; k++) {
          const int c = c0 + k;
          if (c < nv) Ar[c] = c < r ? e[k] : (c == r ? e[k] - dd : 0.f);
        }
      }
    }
    __syncwarp();
    if (m.ntendon > 0 && m.integrator == B2S_INT_IMPLICITFAST && lane == 0) {
      // velocity derivatives carried by fixed tendons, moment^T s moment: actuators in order, then tendon dampers
      // (one lane, sequential: the sum order is part of the result)
      if (!(m.disableflags & B2S_DSBL_ACTUATION)) {
        const float* gainprm = m.actuator_gainprm.at(w);
        const float* biasprm = m.actuator_biasprm.at(w);
        for (int u = 0; u < m.nu; u++) {
          if (m.actuator_trntype[u] != B2S_TRN_TENDON) continue;
          float bv = 0.f;
          if (m.actuator_biastype[u] == B2S_BIAS_AFFINE) bv += biasprm[3 * u + 2];
          if (m.actuator_gaintype[u] == B2S_GAIN_AFFINE) {
            float c = d.ctrl[(size_t)w * m.nu + u];
            if (m.actuator_ctrllimited[u] && !(m.disableflags & B2S_DSBL_CLAMPCTRL)) c = clampf(c, m.actuator_ctrlrange[2 * u], m.actuator_ctrlrange[2 * u + 1]);
            bv += gainprm[3 * u + 2] * c;
          }
          if (bv == 0.f) continue;
          if (m.actuator_forcelimited[u]) {
            const float f = d.actuator_force[(size_t)w * m.nu + u];
            if (f <= m.actuator_forcerange[2 * u] || f >= m.actuator_forcerange[2 * u + 1]) continue;
          }
          const float g = m.actuator_gear[u], sc = g * g * bv;
          const int t = m.actuator_trnid[u], a0 = m.tendon_adr[t], a1 = a0 + m.tendon_num[t];
          for (int a = a0; a < a1; a++)
            for (int b = a0; b < a1; b++) {
              const int r = m.wrap_dof[a], c = m.wrap_dof[b];
              if (c <= r) A[r * ld + c] -= dt * (sc * m.wrap_coef[a] * m.wrap_coef[b]);
            }
        }
      }
      if (!(m.disableflags & B2S_DSBL_PASSIVE))
        for (int t = 0; t < m.ntendon; t++) {
          const float bd = m.tendon_damping[t];
          if (bd == 0.f) continue;
          const int a0 = m.tendon_adr[t], a1 = a0 + m.tendon_num[t];
          for (int a = a0; a < a1; a++)
            for (int b = a0; b < a1; b++) {
              const int r = m.wrap_dof[a], c = m.wrap_dof[b];
              if (c <= r) A[r * ld + c] += dt * (bd * m.wrap_coef[a] * m.wrap_coef[b]);
            }
        }
    }
    __syncwarp();
    rowchol_smem<T>(A, nv, ld, lane, invd, s_col);
    float b0 = 0.f, b1 = 0.f;
    if (lane < nv) b0 = d.qfrc_smooth[(size_t)w * nv + lane] + d.qfrc_constraint[(size_t)w * nv + lane];
    if (lane + 32 < nv) b1 = d.qfrc_smooth[(size_t)w * nv + lane + 32] + d.qfrc_constraint[(size_t)w * nv + lane + 32];
    rowchol_solve<T>(A, invd, ld, lane, b0, b1);
    if (lane < nv) qa[lane] = b0;
    if (lane + 32 < nv) qa[lane + 32] = b1;
  }
  __syncwarp();
  // mj_advance
  for (int i = lane; i < nv; i += 32) { const float v = qvel[i] + dt * qa[i]; qvel[i] = v; dg[i] = v; }
  __syncwarp();
  for (int j = lane; j < m.njnt; j += 32) {
    int pa = m.jnt_qposadr[j], da = m.jnt_dofadr[j];
    const int t = m.jnt_type[j];
    if (t == B2S_JNT_FREE) { for (int k = 0; k < 3; k++) qpos[pa + k] += dt * dg[da + k]; pa += 3; da += 3; }
    if (t == B2S_JNT_FREE || t == B2S_JNT_BALL) {
      v3 wv = mk3(dg[da], dg[da + 1], dg[da + 2]);
      const float ang = norm(wv) * dt;
      q4 q = qnormalize(ldq(qpos + pa));
      q4 dq = axis_angle(normalized(wv), ang);
      stq(qpos + pa, qnormalize(qmul(q, dq)));
    } else qpos[pa] += dt * dg[da];
  }
  for (int i = lane; i < nv; i += 32) d.qacc_warmstart[(size_t)w * nv + i] = qacc[i];
  if (lane == 0) d.time[w] += dt;
}

void launch_sensors(const ModelDev& m, const DataDev& d, cudaStream_t st) {
  const int grid = (d.nworld + B2S_WARPS_PER_CTA - 1) / B2S_WARPS_PER_CTA;
  k_sensors<<<grid, 32 * B2S_WARPS_PER_CTA, smem_sensors(m, d), st>>>(m, d);
  b2s_count_launch();
}
void launch_integrate(const ModelDev& m, const DataDev& d, cudaStream_t st) {
  const int grid = (d.nworld + B2S_WARPS_PER_CTA - 1) / B2S_WARPS_PER_CTA;
  const size_t smem = smem_integrate(m);
  switch (b2s_tile_for(m.nv)) {
    case 1: k_integrate<1><<<grid, 32 * B2S_WARPS_PER_CTA, smem, st>>>(m, d); break;
    case 2: k_integrate<2><<<grid, 32 * B2S_WARPS_PER_CTA, smem, st>>>(m, d); break;
    case 3: k_integrate<3><<<grid, 32 * B2S_WARPS_PER_CTA, smem, st>>>(m, d); break;
    case 4: k_integrate<4><<<grid, 32 * B2S_WARPS_PER_CTA, smem, st>>>(m, d); break;
    case 5: k_integrate<5><<<grid, 32 * B2S_WARPS_PER_CTA, smem, st>>>(m, d); break;
    case 6: k_integrate<6><<<grid, 32 * B2S_WARPS_PER_CTA, smem, st>>>(m, d); break;
    default: k_integrate<8><<<grid, 32 * B2S_WARPS_PER_CTA, smem, st>>>(m, d); break;
  }
  b2s_count_launch();
}
size_t smem_sensors(const ModelDev& m, const DataDev& d) {
  return (size_t)B2S_WARPS_PER_CTA * (sens_smem_floats(m.nbody, m.nv, m.need_cfrc) + (m.need_concache ? cc_warp_words(d.nconmax, m.nsensor) : 0)) * sizeof(float);
}
size_t smem_integrate(const ModelDev& m) {
  const int T = b2s_tile_for(m.nv);
  return (size_t)B2S_WARPS_PER_CTA * integ_smem_floats(T, m.nvp) * sizeof(float);
}
const void* fn_sensors() { return (const void*)k_sensors; }
const void* fn_integrate(const ModelDev& m) {
  switch (b2s_tile_for(m.nv)) {
    case 1: return (const void*)k_integrate<1>;
    case 2: return (const void*)k_integrate<2>;
    case 3: return (const void*)k_integrate<3>;
    case 4: return (const void*)k_integrate<4>;
    case 5: return (const void*)k_integrate<5>;
    case 6: return (const void*)k_integrate<6>;
    default: return (const void*)k_integrate<8>;
  }
}
