// b2s_api.cu -- the C ABI declared in include/b200sim.h: model validation and upload (put_model),
// per-world allocation (put_data), zero-copy field registry, and the forward/step launch sequences.
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "b2s_api.h"

#define B2S_VERSION 100

static thread_local std::string g_err;
static std::atomic<long long> g_launches{0};
void b2s_count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

static int fail(const std::string& msg) { g_err = msg; return -1; }

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-function LIMIT: setting it to a smaller value makes launches that were
// legal before fail with "invalid argument".  Several Data objects of different models share the non-template stage kernels,
// so the limit is only ever raised: the largest request per (device, function) is remembered here.
#include <mutex>
#include <nvtx3/nvToolsExt.h>
static std::mutex g_smem_mu;
static std::map<std::pair<int, const void*>, size_t> g_smem_max;
int b2s_raise_smem(int device, const void* fn, size_t smem) {
  std::lock_guard<std::mutex> lk(g_smem_mu);
  size_t& cur = g_smem_max[{device, fn}];
  if (smem <= cur) return 0;
  if (cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) { cudaGetLastError(); return smem > 48 * 1024 ? -1 : 0; }
  cur = smem;
  return 0;
}
#define CU(call)                                                                         \
  do {                                                                                   \
    cudaError_t e_ = (call);                                                             \
    if (e_ != cudaSuccess) return fail(std::string(#call) + ": " + cudaGetErrorString(e_)); \
  } while (0)

struct FieldRec {
  std::string name;
  int dtype, ndim;
  long long shape[4];
  void* ptr;                 // Data fields / fixed Model fields
  WField<float>* wf;         // expandable Model fields (ptr lives in the ModelDev)
  long long elems;           // per-world element count
  long long shared = 0;      // Data fields: elements stored ONCE ahead of the per-world block (ptr points at the shared block)
};

struct b2s_model {
  int device = 0;
  int overlap = 1;           // option "stage_overlap": 0 serial stages; 1 collision beside dynamics, sensors beside the integrator and the solver's second
                             // launch on side streams; 2 only the solver's second launch.  Set per model at creation (see b2s_model_create)
  ModelDev dev;
  std::vector<void*> allocs;
  std::vector<FieldRec> fields;
  std::map<std::string, int> index;
  int nexpanded = 0;
  ~b2s_model() { for (void* p : allocs) cudaFree(p); }
};
struct b2s_data {
  int device = 0;
  DataDev dev;
  void* arena = nullptr;
  size_t bytes = 0;
  std::vector<FieldRec> fields;
  std::map<std::string, int> index;
  cudaEvent_t ev[16];
  bool have_ev = false;
  // fork / join of independent stages (collision beside dynamics, sensors beside the integrator): one internal side stream,
  // forked from and joined back into the caller's stream with these events, so a stream capture records a DAG
  cudaStream_t side = nullptr;
  cudaStream_t side_hi = nullptr;     // highest priority: the solver's launch for worlds with many rows must get its CTAs in before the main launch fills the machine
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_fork2 = nullptr, ev_join2 = nullptr;
  float* efc_J = nullptr;    // dense Jacobian, allocated and filled on demand (b2s_efc_jacobian)
  ~b2s_data() {
    if (arena) cudaFree(arena);
    if (efc_J) cudaFree(efc_J);
    if (have_ev) for (auto& e : ev) cudaEventDestroy(e);
    for (cudaEvent_t e : {ev_fork, ev_join, ev_fork2, ev_join2}) if (e) cudaEventDestroy(e);
    if (side) cudaStreamDestroy(side);
    if (side_hi) cudaStreamDestroy(side_hi);
  }
};

extern "C" const char* b2s_last_error(void) { return g_err.c_str(); }
extern "C" int b2s_version(void) { return B2S_VERSION; }
extern "C" int b2s_built_arch(void) { return 100; }
extern "C" int64_t b2s_launch_count(void) { return (int64_t)g_launches.load(); }

// tile src[n] into dst[nworld][n]
__global__ void k_tile(float* dst, const float* src, long long n, long long total) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) dst[i] = src[i % n];
}
__global__ void k_fill_i(int* dst, int v, long long total) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) dst[i] = v;
}

namespace {
struct Uploader {
  b2s_model* m;
  bool ok = true;
  std::string err;
  template <typename T>
  T* raw(const std::vector<T>& h) {
    T* p = nullptr;
    size_t n = h.size() ? h.size() : 1;
    if (cudaMalloc(&p, n * sizeof(T)) != cudaSuccess) { ok = false; err = "cudaMalloc failed"; return nullptr; }
    m->allocs.push_back(p);
    if (h.size() && cudaMemcpy(p, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice) != cudaSuccess) { ok = false; err = "cudaMemcpy failed"; }
    return p;
  }
  const float* f(const double* src, long long n) {
    std::vector<float> h((size_t)n);
    for (long long i = 0; i < n; i++) h[(size_t)i] = (float)src[i];
    return raw(h);
  }
  const int* i(const int32_t* src, long long n) {
    std::vector<int> h(src, src + n);
    return raw(h);
  }
};
const char* geom_name(int t) {
  static const char* n[] = {"plane", "hfield", "sphere", "capsule", "ellipsoid", "cylinder", "box", "mesh"};
  return (t >= 0 && t < 8) ? n[t] : "?";
}
bool round_geom(int t) { return t == B2S_GEOM_CYLINDER || t == B2S_GEOM_ELLIPSOID; }
bool convex_core(int t) { return t == B2S_GEOM_SPHERE || t == B2S_GEOM_CAPSULE || t == B2S_GEOM_BOX || t == B2S_GEOM_MESH || round_geom(t); }
bool pair_supported(int t1, int t2) {
  if (t1 == B2S_GEOM_PLANE) return t2 == B2S_GEOM_SPHERE || t2 == B2S_GEOM_CAPSULE || t2 == B2S_GEOM_BOX || t2 == B2S_GEOM_MESH || round_geom(t2);
  if (t1 == B2S_GEOM_HFIELD) return convex_core(t2);
  if (t2 == B2S_GEOM_MESH) return convex_core(t1);
  if (round_geom(t1) || round_geom(t2)) return convex_core(t1) && convex_core(t2);
  if (t1 == B2S_GEOM_SPHERE) return t2 == B2S_GEOM_SPHERE || t2 == B2S_GEOM_CAPSULE || t2 == B2S_GEOM_BOX;
  if (t1 == B2S_GEOM_CAPSULE) return t2 == B2S_GEOM_CAPSULE || t2 == B2S_GEOM_BOX;
  if (t1 == B2S_GEOM_BOX) return t2 == B2S_GEOM_BOX;
  return false;
}
// pairs that go through the GJK/EPA, plane-mesh or height-field colliders (k_convex.cuh)
// (the analytic plane - cylinder / ellipsoid colliders are only compiled into that instantiation too)
bool pair_convex(int t1, int t2) { return t1 == B2S_GEOM_HFIELD || t2 == B2S_GEOM_MESH || round_geom(t1) || round_geom(t2); }
// most contacts the narrowphase returns for a (type-ordered) pair
int pair_max_contacts(int t1, int t2) {
  if (t1 == B2S_GEOM_HFIELD) return 8;
  if (t2 == B2S_GEOM_MESH) return t1 == B2S_GEOM_PLANE ? 4 : 1;
  if (t1 == B2S_GEOM_PLANE && t2 == B2S_GEOM_CYLINDER) return 4;
  if (t1 == B2S_GEOM_BOX && t2 == B2S_GEOM_BOX) return 8;
  if (t1 == B2S_GEOM_PLANE && t2 == B2S_GEOM_BOX) return 4;
  if (t2 == B2S_GEOM_CAPSULE && (t1 == B2S_GEOM_PLANE || t1 == B2S_GEOM_CAPSULE)) return 2;
  if (t1 == B2S_GEOM_CAPSULE && t2 == B2S_GEOM_BOX) return 2;
  return 1;
}
// host-side pose algebra for the static-geom grid (double precision)
void hq_mul(double* r, const double* a, const double* b) {
  r[0] = a[0]*b[0] - a[1]*b[1] - a[2]*b[2] - a[3]*b[3]; r[1] = a[0]*b[1] + a[1]*b[0] + a[2]*b[3] - a[3]*b[2];
  r[2] = a[0]*b[2] - a[1]*b[3] + a[2]*b[0] + a[3]*b[1]; r[3] = a[0]*b[3] + a[1]*b[2] - a[2]*b[1] + a[3]*b[0];
}
void hq_mat(double* m, const double* q) {
  const double w = q[0], x = q[1], y = q[2], z = q[3];
  m[0] = w*w + x*x - y*y - z*z; m[1] = 2*(x*y - w*z); m[2] = 2*(x*z + w*y);
  m[3] = 2*(x*y + w*z); m[4] = w*w - x*x + y*y - z*z; m[5] = 2*(y*z - w*x);
  m[6] = 2*(x*z - w*y); m[7] = 2*(y*z + w*x); m[8] = w*w - x*x - y*y + z*z;
}
void add_field(std::vector<FieldRec>& v, std::map<std::string, int>& idx, const char* name, int dtype, void* ptr, WField<float>* wf,
               const std::vector<long long>& shape) {
  FieldRec r;
  r.name = name; r.dtype = dtype; r.ptr = ptr; r.wf = wf; r.ndim = (int)shape.size(); r.elems = 1;
  int k = 0;
  for (long long s : shape) { r.shape[k++] = s; r.elems *= s; }
  for (; k < 4; k++) r.shape[k] = 1;
  idx[name] = (int)v.size();
  v.push_back(r);
}
}  // namespace

extern "C" int b2s_model_create(const b2s_model_desc* s, int device, b2s_model** out) {
  if (!s || !out) return fail("null argument");
  // ---------------------------------------------------------------- validation: unsupported features are errors
  if (s->na > 0) return fail("stateful actuators (na > 0) are not supported");
  if (s->nv < 1 || s->nv > B2S_MAX_NV) return fail("nv must be in [1, " + std::to_string(B2S_MAX_NV) + "]");
  if (s->opt_cone != B2S_CONE_PYRAMIDAL && s->opt_cone != B2S_CONE_ELLIPTIC) return fail("cone must be pyramidal or elliptic");
  if (s->opt_cone == B2S_CONE_ELLIPTIC) for (int g = 0; g < s->ngeom; g++) if (s->geom_condim[g] > 3) return fail("elliptic cones are supported for condim 1 and 3 only");
  if (s->opt_solver != B2S_SOL_NEWTON && s->opt_solver != B2S_SOL_CG) return fail("solver must be newton or cg (PGS is not supported)");
  if (s->opt_integrator != B2S_INT_EULER && s->opt_integrator != B2S_INT_IMPLICITFAST) return fail("integrator must be euler or implicitfast");
  const int nb = s->nbody, nv = s->nv;
  for (int i = 0; i < nv; i++) if (s->dof_frictionloss[i] < 0.0) return fail("dof_frictionloss must be non-negative");
  for (int j = 0; j < s->njnt; j++) {
    if (s->jnt_type[j] < 0 || s->jnt_type[j] > 3) return fail("unknown joint type");
    if (s->jnt_type[j] == B2S_JNT_BALL && s->jnt_limited[j]) return fail("ball joint limits are not supported");
  }
  for (int g = 0; g < s->ngeom; g++) {
    const bool collides = (s->geom_contype[g] | s->geom_conaffinity[g]) != 0;
    if (s->geom_type[g] == B2S_GEOM_MESH && collides) {
      const int id = s->geom_dataid[g];
      if (id < 0 || id >= s->nmesh || s->mesh_vertnum[id] < 1 || s->mesh_vertadr[id] < 0 || s->mesh_vertadr[id] + s->mesh_vertnum[id] > s->nmeshvert)
        return fail("colliding mesh geom " + std::to_string(g) + " has no convex-hull vertices (geom_dataid / mesh_vert)");
    }
    if (s->geom_type[g] == B2S_GEOM_HFIELD && collides) {
      const int id = s->geom_dataid[g];
      if (id < 0 || id >= s->nhfield || s->hfield_nrow[id] < 2 || s->hfield_ncol[id] < 2 ||
          s->hfield_adr[id] < 0 || s->hfield_adr[id] + s->hfield_nrow[id] * s->hfield_ncol[id] > s->nhfielddata)
        return fail("height-field geom " + std::to_string(g) + " has no elevation data");
      if ((long long)s->hfield_nrow[id] * s->hfield_ncol[id] * 2 >= (1LL << 26)) return fail("height field with more than 2^25 cells is not supported");
    }
  }
  if (s->ntendon < 0 || s->nwrap < 0) return fail("negative tendon sizes");
  for (int t = 0; t < s->ntendon; t++) {
    if (s->tendon_limited[t]) return fail("tendon limits are not supported");
    if (s->tendon_adr[t] < 0 || s->tendon_num[t] < 0 || s->tendon_adr[t] + s->tendon_num[t] > s->nwrap) return fail("tendon wrap range out of bounds");
  }
  for (int k = 0; k < s->nwrap; k++) {
    const int j = s->wrap_objid[k];
    if (s->wrap_type[k] != 1) return fail("only fixed tendons (joint wraps) are supported");
    if (j < 0 || j >= s->njnt || (s->jnt_type[j] != B2S_JNT_HINGE && s->jnt_type[j] != B2S_JNT_SLIDE)) return fail("fixed tendons must wrap hinge or slide joints");
  }
  for (int u = 0; u < s->nu; u++) {
    const int j = s->actuator_trnid[u];
    if (s->actuator_trntype[u] == B2S_TRN_TENDON) {
      if (j < 0 || j >= s->ntendon) return fail("actuator tendon id out of range");
    } else if (s->actuator_trntype[u] != B2S_TRN_JOINT) {
      return fail("actuator transmission must be joint or tendon");
    } else if (j < 0 || j >= s->njnt || (s->jnt_type[j] != B2S_JNT_HINGE && s->jnt_type[j] != B2S_JNT_SLIDE)) {
      return fail("actuators must target hinge or slide joints");
    }
  }
  // bodies must be in DFS pre-order (MuJoCo's compiler guarantees it)
  std::vector<int> send(nb), ssize(nb, 1), depth(nb, 0);
  for (int i = 0; i < nb; i++) send[i] = i + 1;
  for (int i = nb - 1; i > 0; i--) {
    const int p = s->body_parentid[i];
    if (p < 0 || p >= i) return fail("bodies are not topologically ordered");
    send[p] = send[p] > send[i] ? send[p] : send[i];
    ssize[p] += ssize[i];
  }
  for (int i = 0; i < nb; i++) if (send[i] - i != ssize[i]) return fail("bodies are not in depth-first order");
  for (int i = 1; i < nb; i++) depth[i] = depth[s->body_parentid[i]] + 1;
  for (int k = 0; k < s->nsensor; k++) {
    switch (s->sensor_type[k]) {
      case B2S_SENS_ACCELEROMETER: case B2S_SENS_VELOCIMETER: case B2S_SENS_GYRO: case B2S_SENS_JOINTPOS: case B2S_SENS_JOINTVEL:
      case B2S_SENS_ACTUATORFRC: case B2S_SENS_FRAMEPOS: case B2S_SENS_FRAMEQUAT: case B2S_SENS_FRAMEZAXIS: case B2S_SENS_FRAMELINVEL:
      case B2S_SENS_FRAMEANGVEL: case B2S_SENS_SUBTREECOM: case B2S_SENS_SUBTREELINVEL: case B2S_SENS_SUBTREEANGMOM: case B2S_SENS_CONTACT: break;
      case B2S_SENS_ACTUATORPOS: case B2S_SENS_ACTUATORVEL: case B2S_SENS_JOINTACTFRC: case B2S_SENS_FRAMEXAXIS: case B2S_SENS_FRAMEYAXIS:
      case B2S_SENS_FRAMELINACC: case B2S_SENS_FRAMEANGACC: case B2S_SENS_CLOCK: break;
      case B2S_SENS_MAGNETOMETER: case B2S_SENS_E_POTENTIAL: case B2S_SENS_E_KINETIC: break;
      case B2S_SENS_RANGEFINDER:
        if (s->sensor_objtype[k] != B2S_OBJ_SITE || s->sensor_objid[k] < 0 || s->sensor_objid[k] >= s->nsite) return fail("rangefinder needs a site");
        for (int g = 0; g < s->ngeom; g++)
          if (s->geom_type[g] == B2S_GEOM_MESH || s->geom_type[g] == B2S_GEOM_HFIELD) return fail("rangefinder: rays against mesh / height-field geoms are not implemented");
        break;
      case B2S_SENS_FORCE: case B2S_SENS_TORQUE:
        if (s->sensor_objtype[k] != B2S_OBJ_SITE || s->sensor_objid[k] < 0 || s->sensor_objid[k] >= s->nsite) return fail("force / torque sensor needs a site");
        break;
      case B2S_SENS_JOINTLIMITPOS: case B2S_SENS_JOINTLIMITVEL: case B2S_SENS_JOINTLIMITFRC:
        if (s->sensor_objid[k] < 0 || s->sensor_objid[k] >= s->njnt) return fail("joint-limit sensor refers to an unknown joint");
        break;
      case B2S_SENS_TENDONACTFRC:
      case B2S_SENS_TENDONPOS: case B2S_SENS_TENDONVEL:
        if (s->sensor_objid[k] < 0 || s->sensor_objid[k] >= s->ntendon) return fail("tendon sensor refers to an unknown tendon");
        break;
      default: return fail("unsupported sensor type " + std::to_string(s->sensor_type[k]));
    }
  }
  // ---------------------------------------------------------------- static-geom grid (terrain)
  // Non-plane geoms on world-welded bodies go into a uniform xy grid when there are many of them; their pairs with
  // dynamic geoms are then found per step through the grid instead of the static pair table.
  std::vector<char> in_grid((size_t)s->ngeom, 0);
  std::vector<int> grid_geoms, dyn_geoms;
  for (int g = 0; g < s->ngeom; g++) {
    const bool collides = (s->geom_contype[g] | s->geom_conaffinity[g]) != 0;
    // mocap bodies are welded to the world but move: their geoms stay in the per-step pair table, never in the grid
    const bool stat = s->body_weldid[s->geom_bodyid[g]] == 0 && s->body_mocapid[s->geom_bodyid[g]] < 0;
    if (collides && stat && s->geom_type[g] != B2S_GEOM_PLANE) grid_geoms.push_back(g);
    if (collides && !stat) dyn_geoms.push_back(g);
  }
  const bool use_grid = (int)grid_geoms.size() > B2S_GRID_MIN_GEOMS;
  std::vector<float> grid_aabb;
  std::vector<int> grid_adr, grid_list;
  int grid_nx = 0, grid_ny = 0; double grid_x0 = 0, grid_y0 = 0, grid_cell = 1.0, grid_margin = 0;
  if (use_grid) {
    if (s->ngeom > 65535) return fail("more than 65535 geoms are not supported with a static-geom grid");
    grid_aabb.assign((size_t)6 * s->ngeom, 0.f);
    double lo[2] = {1e300, 1e300}, hi[2] = {-1e300, -1e300};
    std::vector<int> types_seen;
    for (int g : grid_geoms) {
      in_grid[g] = 1;
      for (int e = 0; e < s->nexclude; e++)
        if (s->exclude_body1[e] == s->geom_bodyid[g] || s->exclude_body2[e] == s->geom_bodyid[g])
          return fail("contact excludes that involve a static terrain body are not supported");
      // world pose of the static geom: compose the (joint-less) body chain
      double q[4] = {1, 0, 0, 0}, p[3] = {0, 0, 0};
      std::vector<int> chainb;
      for (int b = s->geom_bodyid[g]; b > 0; b = s->body_parentid[b]) chainb.push_back(b);
      for (int k = (int)chainb.size() - 1; k >= 0; k--) {
        const int b = chainb[k];
        double R[9], q2[4];
        hq_mat(R, q);
        for (int a = 0; a < 3; a++) p[a] += R[3*a]*s->body_pos[3*b] + R[3*a+1]*s->body_pos[3*b+1] + R[3*a+2]*s->body_pos[3*b+2];
        hq_mul(q2, q, s->body_quat + 4*b);
        for (int a = 0; a < 4; a++) q[a] = q2[a];
      }
      double R[9], qg[4], Rg[9], c[3];
      hq_mat(R, q);
      for (int a = 0; a < 3; a++) c[a] = p[a] + R[3*a]*s->geom_pos[3*g] + R[3*a+1]*s->geom_pos[3*g+1] + R[3*a+2]*s->geom_pos[3*g+2];
      hq_mul(qg, q, s->geom_quat + 4*g);
      hq_mat(Rg, qg);
      const double* sz = s->geom_size + 3*g;
      double h[3];
      double hfs[3] = {0, 0, 0};
      if (s->geom_type[g] == B2S_GEOM_HFIELD) {       // box [-sx,sx] x [-sy,sy] x [-base, top] in the field's frame
        const double* hs = s->hfield_size + 4*s->geom_dataid[g];
        const double zc = 0.5*(hs[2] - hs[3]);
        for (int a = 0; a < 3; a++) c[a] += Rg[3*a+2]*zc;
        hfs[0] = hs[0]; hfs[1] = hs[1]; hfs[2] = 0.5*(hs[2] + hs[3]);
      }
      for (int a = 0; a < 3; a++) {
        if (s->geom_type[g] == B2S_GEOM_HFIELD) h[a] = std::fabs(Rg[3*a])*hfs[0] + std::fabs(Rg[3*a+1])*hfs[1] + std::fabs(Rg[3*a+2])*hfs[2];
        else if (s->geom_type[g] == B2S_GEOM_BOX) h[a] = std::fabs(Rg[3*a])*sz[0] + std::fabs(Rg[3*a+1])*sz[1] + std::fabs(Rg[3*a+2])*sz[2];
        else if (s->geom_type[g] == B2S_GEOM_SPHERE) h[a] = sz[0];
        else if (s->geom_type[g] == B2S_GEOM_CAPSULE) h[a] = std::fabs(Rg[3*a+2])*sz[1] + sz[0];
        else h[a] = s->geom_rbound[g];
        h[a] += s->geom_margin[g];
      }
      grid_margin = s->geom_margin[g] > grid_margin ? s->geom_margin[g] : grid_margin;
      for (int a = 0; a < 3; a++) { grid_aabb[(size_t)6*g + a] = (float)(c[a] - h[a]); grid_aabb[(size_t)6*g + 3 + a] = (float)(c[a] + h[a]); }
      for (int a = 0; a < 2; a++) { lo[a] = std::fmin(lo[a], c[a] - h[a]); hi[a] = std::fmax(hi[a], c[a] + h[a]); }
      // every dynamic collision geom type must have a narrowphase against this static geom type
      for (int dg : dyn_geoms) {
        if (!((s->geom_contype[g] & s->geom_conaffinity[dg]) || (s->geom_contype[dg] & s->geom_conaffinity[g]))) continue;
        int t1 = s->geom_type[g], t2 = s->geom_type[dg];
        if (t1 > t2) { int tt = t1; t1 = t2; t2 = tt; }
        if (!pair_supported(t1, t2)) return fail(std::string("unsupported collision pair ") + geom_name(t1) + "-" + geom_name(t2) + " (terrain geom " + std::to_string(g) + ")");
      }
    }
    while ((hi[0] - lo[0]) / grid_cell * ((hi[1] - lo[1]) / grid_cell) > 1.0e6) grid_cell *= 2.0;
    grid_x0 = lo[0]; grid_y0 = lo[1];
    grid_nx = (int)std::floor((hi[0] - lo[0]) / grid_cell) + 1; grid_ny = (int)std::floor((hi[1] - lo[1]) / grid_cell) + 1;
    std::vector<int> count((size_t)grid_nx * grid_ny + 1, 0);
    auto cells = [&](int g, int& x0, int& x1, int& y0, int& y1) {
      // computed in float exactly as the kernel does, so both sides agree on a geom's first cell
      const float inv = (float)(1.0 / grid_cell), fx0 = (float)grid_x0, fy0 = (float)grid_y0;
      x0 = (int)std::floor((grid_aabb[(size_t)6*g] - fx0) * inv); x1 = (int)std::floor((grid_aabb[(size_t)6*g + 3] - fx0) * inv);
      y0 = (int)std::floor((grid_aabb[(size_t)6*g + 1] - fy0) * inv); y1 = (int)std::floor((grid_aabb[(size_t)6*g + 4] - fy0) * inv);
      x0 = x0 < 0 ? 0 : x0; y0 = y0 < 0 ? 0 : y0; x1 = x1 > grid_nx - 1 ? grid_nx - 1 : x1; y1 = y1 > grid_ny - 1 ? grid_ny - 1 : y1;
    };
    for (int g : grid_geoms) { int x0, x1, y0, y1; cells(g, x0, x1, y0, y1); for (int x = x0; x <= x1; x++) for (int y = y0; y <= y1; y++) count[(size_t)x * grid_ny + y + 1]++; }
    for (size_t k = 1; k < count.size(); k++) count[k] += count[k - 1];
    grid_adr = count;
    grid_list.assign((size_t)grid_adr.back(), 0);
    std::vector<int> fill(grid_adr.begin(), grid_adr.end() - 1);
    for (int g : grid_geoms) { int x0, x1, y0, y1; cells(g, x0, x1, y0, y1); for (int x = x0; x <= x1; x++) for (int y = y0; y <= y1; y++) grid_list[(size_t)fill[(size_t)x * grid_ny + y]++] = g; }
  }
  // ---------------------------------------------------------------- static collision pair table
  std::vector<int> pairs;
  int col_convex = 0, col_defer = 0, col_hf = 0;
  auto pair_deferred = [&](int t1, int t2) { return pair_convex(t1, t2) && t1 != B2S_GEOM_PLANE && t1 != B2S_GEOM_HFIELD; };
  int col_maxc = 1;          // most contacts one geom pair of this model can produce: picks the narrowphase variant
  for (int ga = 0; ga < s->ngeom; ga++) for (int gb = ga + 1; gb < s->ngeom; gb++) {
    if (in_grid[ga] || in_grid[gb]) continue;      // found through the grid at run time
    const int b1 = s->geom_bodyid[ga], b2 = s->geom_bodyid[gb];
    if (!((s->geom_contype[ga] & s->geom_conaffinity[gb]) || (s->geom_contype[gb] & s->geom_conaffinity[ga]))) continue;
    const int w1 = s->body_weldid[b1], w2 = s->body_weldid[b2];
    if (w1 == w2) continue;
    const int wp1 = s->body_weldid[s->body_parentid[w1]], wp2 = s->body_weldid[s->body_parentid[w2]];
    if (!(s->opt_disableflags & B2S_DSBL_FILTERPARENT) && w1 != 0 && w2 != 0 && (w1 == wp2 || w2 == wp1)) continue;
    bool ex = false;
    for (int e = 0; e < s->nexclude; e++) {
      const int e1 = s->exclude_body1[e], e2 = s->exclude_body2[e];
      if ((e1 == b1 && e2 == b2) || (e1 == b2 && e2 == b1)) ex = true;
    }
    if (ex) continue;
    int g1 = ga, g2 = gb;
    if (s->geom_type[g1] > s->geom_type[g2]) { g1 = gb; g2 = ga; }
    if (!pair_supported(s->geom_type[g1], s->geom_type[g2]))
      return fail(std::string("unsupported collision pair ") + geom_name(s->geom_type[g1]) + "-" + geom_name(s->geom_type[g2]) +
                  " (geoms " + std::to_string(g1) + ", " + std::to_string(g2) + ")");
    pairs.push_back(g1); pairs.push_back(g2);
    col_maxc = std::max(col_maxc, pair_max_contacts(s->geom_type[g1], s->geom_type[g2]));
    col_convex |= pair_convex(s->geom_type[g1], s->geom_type[g2]) ? 1 : 0;
    col_defer |= pair_deferred(s->geom_type[g1], s->geom_type[g2]) ? 1 : 0;
    col_hf |= s->geom_type[g1] == B2S_GEOM_HFIELD ? 1 : 0;
  }
  if (use_grid)   // grid pairs: every dynamic geom against every geom type present in the grid
    for (int ga : dyn_geoms) for (int gb : grid_geoms) {
      if (!((s->geom_contype[ga] & s->geom_conaffinity[gb]) || (s->geom_contype[gb] & s->geom_conaffinity[ga]))) continue;
      const int t1 = std::min(s->geom_type[ga], s->geom_type[gb]), t2 = std::max(s->geom_type[ga], s->geom_type[gb]);
      col_maxc = std::max(col_maxc, pair_max_contacts(t1, t2));
      col_convex |= pair_convex(t1, t2) ? 1 : 0;
      col_defer |= pair_deferred(t1, t2) ? 1 : 0;
      col_hf |= t1 == B2S_GEOM_HFIELD ? 1 : 0;
    }
  // ---------------------------------------------------------------- traversal tables
  int nlevel = 0;
  for (int i = 0; i < nb; i++) nlevel = depth[i] + 1 > nlevel ? depth[i] + 1 : nlevel;
  std::vector<int> level_adr(nlevel + 1, 0), level_body(nb);
  for (int i = 0; i < nb; i++) level_adr[depth[i] + 1]++;
  for (int l = 0; l < nlevel; l++) level_adr[l + 1] += level_adr[l];
  { std::vector<int> fill(level_adr.begin(), level_adr.end() - 1); for (int i = 0; i < nb; i++) level_body[fill[depth[i]]++] = i; }
  std::vector<int> chain_adr(nb + 1, 0), chain;
  std::vector<unsigned char> dofmask((size_t)nb * nv, 0), ancmask((size_t)nb * nb, 0);
  int maxchain = 0;
  for (int b = 0; b < nb; b++) {
    int bb = b;
    while (bb > 0 && s->body_dofnum[bb] == 0) bb = s->body_parentid[bb];
    std::vector<int> c;
    if (bb > 0) for (int i = s->body_dofadr[bb] + s->body_dofnum[bb] - 1; i >= 0; i = s->dof_parentid[i]) c.push_back(i);
    for (size_t k = 1; k < c.size(); k++) if (c[k] >= c[k - 1]) return fail("dof chain is not ordered");
    for (int k = (int)c.size() - 1; k >= 0; k--) { chain.push_back(c[k]); dofmask[(size_t)b * nv + c[k]] = 1; }
    chain_adr[b + 1] = (int)chain.size();
    maxchain = (int)c.size() > maxchain ? (int)c.size() : maxchain;
    for (int a = b;; a = s->body_parentid[a]) { ancmask[(size_t)b * nb + a] = 1; if (a == 0) break; }
  }
  // ---------------------------------------------------------------- upload
  CU(cudaSetDevice(device));
  b2s_model* m = new b2s_model();
  m->device = device;
  Uploader up{m};
  ModelDev& d = m->dev;
  memset(&d, 0, sizeof(d));
  d.nq = s->nq; d.nv = nv; d.nvp = (nv + 3) & ~3; d.nu = s->nu; d.nbody = nb; d.njnt = s->njnt; d.ngeom = s->ngeom; d.nsite = s->nsite;
  d.nsensor = s->nsensor; d.nsensordata = s->nsensordata; d.nmocap = s->nmocap;
  d.nlevel = nlevel; d.npair = (int)pairs.size() / 2; d.maxchain = maxchain; d.nchain = (int)chain.size();
  d.ref_root = nv > 0 ? s->body_rootid[s->dof_bodyid[0]] : 0;
  d.timestep = (float)s->opt_timestep;
  for (int k = 0; k < 3; k++) { d.gravity[k] = (float)s->opt_gravity[k]; d.magnetic[k] = (float)s->opt_magnetic[k]; }
  d.tolerance = (float)s->opt_tolerance; d.ls_tolerance = (float)s->opt_ls_tolerance; d.impratio = (float)s->opt_impratio;
  d.ls_parallel_min_step = (float)(s->opt_ls_parallel_min_step > 0 ? s->opt_ls_parallel_min_step : 1e-6);
  d.meaninertia = (float)s->stat_meaninertia;
  d.integrator = s->opt_integrator; d.cone = s->opt_cone; d.solver = s->opt_solver;
  d.fric = 0;
  for (int i = 0; i < nv; i++) if (s->dof_frictionloss[i] > 0.0) d.fric = 1; d.iterations = s->opt_iterations; d.ls_iterations = s->opt_ls_iterations;
  d.disableflags = s->opt_disableflags; d.ls_parallel = s->opt_ls_parallel;
  d.contact_sensor_maxmatch = s->opt_contact_sensor_maxmatch > 0 ? s->opt_contact_sensor_maxmatch : 64;
  d.body_parentid = up.i(s->body_parentid, nb); d.body_rootid = up.i(s->body_rootid, nb); d.body_mocapid = up.i(s->body_mocapid, nb);
  d.body_jntnum = up.i(s->body_jntnum, nb); d.body_jntadr = up.i(s->body_jntadr, nb);
  d.body_dofnum = up.i(s->body_dofnum, nb); d.body_dofadr = up.i(s->body_dofadr, nb);
  d.body_subtree_end = up.raw(send); d.body_chain_adr = up.raw(chain_adr); d.body_chain = up.raw(chain);
  d.level_adr = up.raw(level_adr); d.level_body = up.raw(level_body);
  d.body_dofmask = up.raw(dofmask); d.body_ancmask = up.raw(ancmask);
  d.body_subtreemass = up.f(s->body_subtreemass, nb); d.body_invweight0 = up.f(s->body_invweight0, 2 * nb);
  d.jnt_type = up.i(s->jnt_type, s->njnt); d.jnt_qposadr = up.i(s->jnt_qposadr, s->njnt); d.jnt_dofadr = up.i(s->jnt_dofadr, s->njnt);
  d.jnt_bodyid = up.i(s->jnt_bodyid, s->njnt); d.jnt_limited = up.i(s->jnt_limited, s->njnt);
  d.jnt_pos = up.f(s->jnt_pos, 3 * s->njnt); d.jnt_axis = up.f(s->jnt_axis, 3 * s->njnt); d.jnt_margin = up.f(s->jnt_margin, s->njnt);
  d.jnt_solref = up.f(s->jnt_solref, 2 * s->njnt); d.jnt_solimp = up.f(s->jnt_solimp, 5 * s->njnt);
  d.dof_bodyid = up.i(s->dof_bodyid, nv); d.dof_jntid = up.i(s->dof_jntid, nv); d.dof_parentid = up.i(s->dof_parentid, nv);
  d.dof_invweight0 = up.f(s->dof_invweight0, nv);
  d.geom_type = up.i(s->geom_type, s->ngeom); d.geom_bodyid = up.i(s->geom_bodyid, s->ngeom);
  d.geom_condim = up.i(s->geom_condim, s->ngeom); d.geom_priority = up.i(s->geom_priority, s->ngeom);
  d.geom_size = up.f(s->geom_size, 3 * s->ngeom); d.geom_solmix = up.f(s->geom_solmix, s->ngeom);
  d.geom_solref = up.f(s->geom_solref, 2 * s->ngeom); d.geom_solimp = up.f(s->geom_solimp, 5 * s->ngeom);
  d.geom_margin = up.f(s->geom_margin, s->ngeom); d.geom_gap = up.f(s->geom_gap, s->ngeom); d.geom_rbound = up.f(s->geom_rbound, s->ngeom);
  d.pair_geom = up.raw(pairs);
  {
    // a body is static if it is welded to the world and is not a mocap body
    std::vector<int> pg, ps, sg, ss;
    auto is_static = [&](int b) { return s->body_weldid[b] == 0 && s->body_mocapid[b] < 0; };
    for (int g = 0; g < s->ngeom; g++) (is_static(s->geom_bodyid[g]) ? sg : pg).push_back(g);
    for (int k = 0; k < s->nsite; k++) (is_static(s->site_bodyid[k]) ? ss : ps).push_back(k);
    d.nposegeom = (int)pg.size(); d.nposesite = (int)ps.size(); d.nstatgeom = (int)sg.size(); d.nstatsite = (int)ss.size();
    d.pose_geom = up.raw(pg); d.pose_site = up.raw(ps); d.stat_geom = up.raw(sg); d.stat_site = up.raw(ss);
    // many static sites (mjlab's per-env origin markers, terrain_importer.py:98-120): store their frames once (H3)
    std::vector<int> slot((size_t)s->nsite);
    d.site_split = (int)ss.size() >= B2S_SITE_SPLIT_MIN ? 1 : 0;
    if (d.site_split) {
      for (size_t k = 0; k < ss.size(); k++) slot[ss[k]] = ~(int)k;
      for (size_t k = 0; k < ps.size(); k++) slot[ps[k]] = (int)k;
      d.nsite_world = (int)ps.size();
    } else {
      for (int k = 0; k < s->nsite; k++) slot[k] = k;
      d.nsite_world = s->nsite;
    }
    d.site_slot = up.raw(slot);
  }
  d.geom_contype = up.i(s->geom_contype, s->ngeom); d.geom_conaffinity = up.i(s->geom_conaffinity, s->ngeom);
  {
    // broadphase volumes: a box in the geom frame (centre offset + half sizes) and the radius of the sphere around it
    std::vector<float> bcen((size_t)3 * s->ngeom, 0.f), bsize((size_t)3 * s->ngeom, 0.f), brad((size_t)s->ngeom, 0.f);
    for (int g = 0; g < s->ngeom; g++) {
      const double* sz = s->geom_size + 3 * g;
      float* bs = &bsize[(size_t)3 * g];
      const int t = s->geom_type[g];
      brad[g] = (float)s->geom_rbound[g];
      bs[0] = bs[1] = bs[2] = brad[g];
      if (t == B2S_GEOM_SPHERE) { bs[0] = bs[1] = bs[2] = (float)sz[0]; }
      else if (t == B2S_GEOM_CAPSULE) { bs[0] = bs[1] = (float)sz[0]; bs[2] = (float)(sz[0] + sz[1]); }
      else if (t == B2S_GEOM_BOX || t == B2S_GEOM_ELLIPSOID) { bs[0] = (float)sz[0]; bs[1] = (float)sz[1]; bs[2] = (float)sz[2]; }
      else if (t == B2S_GEOM_CYLINDER) { bs[0] = bs[1] = (float)sz[0]; bs[2] = (float)sz[1]; }
      else if (t == B2S_GEOM_MESH && s->geom_dataid[g] >= 0 && s->geom_dataid[g] < s->nmesh) {
        const int id = s->geom_dataid[g], a0 = s->mesh_vertadr[id], nvt = s->mesh_vertnum[id];
        double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
        for (int i = a0; i < a0 + nvt; i++) for (int a = 0; a < 3; a++) { lo[a] = std::fmin(lo[a], s->mesh_vert[3*i + a]); hi[a] = std::fmax(hi[a], s->mesh_vert[3*i + a]); }
        double r2 = 0;
        for (int a = 0; a < 3; a++) { bcen[(size_t)3*g + a] = (float)(0.5*(lo[a] + hi[a])); bs[a] = (float)(0.5*(hi[a] - lo[a]) * 1.000001 + 1e-7); }
        for (int i = a0; i < a0 + nvt; i++) { double q = 0; for (int a = 0; a < 3; a++) { const double e = s->mesh_vert[3*i + a] - 0.5*(lo[a] + hi[a]); q += e*e; } r2 = std::fmax(r2, q); }
        brad[g] = (float)(std::sqrt(r2) * 1.000001 + 1e-7);
      }
    }
    d.geom_bcen = up.raw(bcen); d.geom_bsize = up.raw(bsize); d.geom_brad = up.raw(brad);
  }
  d.col_maxc = col_maxc <= 2 ? 2 : (col_maxc <= 4 ? 4 : 8);
  d.col_convex = col_convex; d.col_defer = col_defer; d.col_hf = col_hf; d.solve_rowcap = 128;
  // stage overlap by default (measured at 4096 worlds, tools/ab7.sh): the side-stream branches pay for the small models (Go1 +2.5 %, on
  // height fields +6 %, Asimov +1.4 %) and cost where the stage kernels fill the register file by themselves (G1: -0.6 %) or where
  // collision is three launches with a persistent task kernel that then fights dynamics for CTA slots (Asimov FULL: -11 %)
  m->overlap = (col_defer || d.nv >= 30) ? 2 : 1;
  d.geom_dataid = up.i(s->geom_dataid, s->ngeom);
  d.mesh_vertadr = up.i(s->mesh_vertadr, s->nmesh); d.mesh_vertnum = up.i(s->mesh_vertnum, s->nmesh);
  {
    std::vector<float> v4((size_t)4 * s->nmeshvert, 0.f);       // 16-byte vertices: one load per vertex in the hull scans
    for (int i = 0; i < s->nmeshvert; i++) for (int a = 0; a < 3; a++) v4[(size_t)4 * i + a] = (float)s->mesh_vert[3 * i + a];
    d.mesh_vert = up.raw(v4);
  }
  d.hfield_size = up.f(s->hfield_size, 4LL * s->nhfield); d.hfield_data = up.f(s->hfield_data, s->nhfielddata);
  d.hfield_nrow = up.i(s->hfield_nrow, s->nhfield); d.hfield_ncol = up.i(s->hfield_ncol, s->nhfield); d.hfield_adr = up.i(s->hfield_adr, s->nhfield);
  d.ngrid = use_grid ? (int)grid_geoms.size() : 0;
  d.ndyngeom = (int)dyn_geoms.size(); d.dyn_geom = up.raw(dyn_geoms);
  d.grid_nx = grid_nx; d.grid_ny = grid_ny; d.grid_x0 = (float)grid_x0; d.grid_y0 = (float)grid_y0;
  d.grid_inv = (float)(1.0 / grid_cell); d.grid_margin = (float)grid_margin;
  d.grid_adr = up.raw(grid_adr); d.grid_geom = up.raw(grid_list); d.grid_aabb = up.raw(grid_aabb);
  d.site_bodyid = up.i(s->site_bodyid, s->nsite);
  d.ntendon = s->ntendon; d.nwrap = s->nwrap;
  {
    std::vector<int> wdof((size_t)s->nwrap), wqpos((size_t)s->nwrap);
    for (int k = 0; k < s->nwrap; k++) { wdof[k] = s->jnt_dofadr[s->wrap_objid[k]]; wqpos[k] = s->jnt_qposadr[s->wrap_objid[k]]; }
    d.wrap_dof = up.raw(wdof); d.wrap_qpos = up.raw(wqpos);
  }
  d.tendon_adr = up.i(s->tendon_adr, s->ntendon); d.tendon_num = up.i(s->tendon_num, s->ntendon);
  d.wrap_coef = up.f(s->wrap_prm, s->nwrap);
  d.tendon_stiffness = up.f(s->tendon_stiffness, s->ntendon); d.tendon_damping = up.f(s->tendon_damping, s->ntendon);
  d.tendon_lengthspring = up.f(s->tendon_lengthspring, 2 * s->ntendon);
  d.actuator_trntype = up.i(s->actuator_trntype, s->nu);
  d.actuator_trnid = up.i(s->actuator_trnid, s->nu); d.actuator_gaintype = up.i(s->actuator_gaintype, s->nu);
  d.actuator_biastype = up.i(s->actuator_biastype, s->nu); d.actuator_ctrllimited = up.i(s->actuator_ctrllimited, s->nu);
  d.actuator_forcelimited = up.i(s->actuator_forcelimited, s->nu); d.actuator_gear = up.f(s->actuator_gear, s->nu);
  d.actuator_ctrlrange = up.f(s->actuator_ctrlrange, 2 * s->nu); d.actuator_forcerange = up.f(s->actuator_forcerange, 2 * s->nu);
  d.sensor_type = up.i(s->sensor_type, s->nsensor); d.sensor_objtype = up.i(s->sensor_objtype, s->nsensor);
  d.sensor_objid = up.i(s->sensor_objid, s->nsensor); d.sensor_reftype = up.i(s->sensor_reftype, s->nsensor);
  d.sensor_refid = up.i(s->sensor_refid, s->nsensor); d.sensor_intprm = up.i(s->sensor_intprm, 3 * s->nsensor);
  d.sensor_dim = up.i(s->sensor_dim, s->nsensor); d.sensor_adr = up.i(s->sensor_adr, s->nsensor); d.sensor_cutoff = up.f(s->sensor_cutoff, s->nsensor);
  for (int k = 0; k < s->nsensor; k++) {
    if (s->sensor_type[k] == B2S_SENS_SUBTREELINVEL || s->sensor_type[k] == B2S_SENS_SUBTREEANGMOM) d.need_subtreevel = 1;
    if (s->sensor_type[k] == B2S_SENS_CONTACT) d.need_concache = 1;
    if (s->sensor_type[k] == B2S_SENS_FORCE || s->sensor_type[k] == B2S_SENS_TORQUE) d.need_concache = d.need_cfrc = 1;
    if (s->sensor_type[k] == B2S_SENS_RANGEFINDER) d.need_ray = 1;
  }
  // expandable (per-world capable) fields: leading dim 1 with world stride 0 until b2s_model_expand
  std::vector<double> rgba((size_t)4 * s->ngeom, 0.5);
  for (int g = 0; g < s->ngeom; g++) rgba[4 * g + 3] = 1.0;
  struct { const char* name; WField<float>* wf; const double* src; std::vector<long long> shape; } ex[] = {
    {"body_pos", &d.body_pos, s->body_pos, {nb, 3}}, {"body_quat", &d.body_quat, s->body_quat, {nb, 4}},
    {"body_ipos", &d.body_ipos, s->body_ipos, {nb, 3}}, {"body_iquat", &d.body_iquat, s->body_iquat, {nb, 4}},
    {"body_mass", &d.body_mass, s->body_mass, {nb}}, {"body_inertia", &d.body_inertia, s->body_inertia, {nb, 3}},
    {"jnt_range", &d.jnt_range, s->jnt_range, {s->njnt, 2}}, {"jnt_stiffness", &d.jnt_stiffness, s->jnt_stiffness, {s->njnt}},
    {"qpos0", &d.qpos0, s->qpos0, {s->nq}}, {"qpos_spring", &d.qpos_spring, s->qpos_spring, {s->nq}},
    {"dof_armature", &d.dof_armature, s->dof_armature, {nv}}, {"dof_damping", &d.dof_damping, s->dof_damping, {nv}},
    {"dof_frictionloss", &d.dof_frictionloss, s->dof_frictionloss, {nv}},
    {"geom_pos", &d.geom_pos, s->geom_pos, {s->ngeom, 3}}, {"geom_quat", &d.geom_quat, s->geom_quat, {s->ngeom, 4}},
    {"geom_friction", &d.geom_friction, s->geom_friction, {s->ngeom, 3}}, {"geom_rgba", &d.geom_rgba, rgba.data(), {s->ngeom, 4}},
    {"site_pos", &d.site_pos, s->site_pos, {s->nsite, 3}}, {"site_quat", &d.site_quat, s->site_quat, {s->nsite, 4}},
    {"actuator_gainprm", &d.actuator_gainprm, s->actuator_gainprm, {s->nu, 3}},
    {"actuator_biasprm", &d.actuator_biasprm, s->actuator_biasprm, {s->nu, 3}},
  };
  for (auto& e : ex) {
    long long n = 1;
    for (long long v : e.shape) n *= v;
    e.wf->ptr = up.f(e.src, n); e.wf->wstride = 0;
    add_field(m->fields, m->index, e.name, B2S_F32, nullptr, e.wf, e.shape);
  }
  add_field(m->fields, m->index, "geom_bodyid", B2S_I32, (void*)d.geom_bodyid, nullptr, {s->ngeom});
  add_field(m->fields, m->index, "site_bodyid", B2S_I32, (void*)d.site_bodyid, nullptr, {s->nsite});
  add_field(m->fields, m->index, "site_slot", B2S_I32, (void*)d.site_slot, nullptr, {s->nsite});
  add_field(m->fields, m->index, "body_parentid", B2S_I32, (void*)d.body_parentid, nullptr, {nb});
  add_field(m->fields, m->index, "body_rootid", B2S_I32, (void*)d.body_rootid, nullptr, {nb});
  add_field(m->fields, m->index, "jnt_qposadr", B2S_I32, (void*)d.jnt_qposadr, nullptr, {s->njnt});
  add_field(m->fields, m->index, "jnt_dofadr", B2S_I32, (void*)d.jnt_dofadr, nullptr, {s->njnt});
  add_field(m->fields, m->index, "geom_type", B2S_I32, (void*)d.geom_type, nullptr, {s->ngeom});
  add_field(m->fields, m->index, "geom_size", B2S_F32, (void*)d.geom_size, nullptr, {s->ngeom, 3});
  add_field(m->fields, m->index, "body_subtreemass", B2S_F32, (void*)d.body_subtreemass, nullptr, {nb});
  add_field(m->fields, m->index, "dof_invweight0", B2S_F32, (void*)d.dof_invweight0, nullptr, {nv});
  add_field(m->fields, m->index, "body_invweight0", B2S_F32, (void*)d.body_invweight0, nullptr, {nb, 2});
  add_field(m->fields, m->index, "pair_geom", B2S_I32, (void*)d.pair_geom, nullptr, {d.npair, 2});
  if (!up.ok) { std::string e = up.err; delete m; return fail(e); }
  *out = m;
  return 0;
}

extern "C" int b2s_model_destroy(b2s_model* m) { delete m; return 0; }

extern "C" int b2s_model_set_opt(b2s_model* m, const char* key, double v) {
  if (!m || !key) return fail("null argument");
  ModelDev& d = m->dev;
  const std::string k(key);
  if (k == "ls_parallel") d.ls_parallel = v != 0;
  else if (k == "stage_overlap") m->overlap = (int)v;      // 0: serial, 1: collision / sensors / the solver's second launch on side streams, 2: only the solver's second launch
  else if (k == "solve_rowcap") { if (v < 1) return fail("solve_rowcap must be positive"); d.solve_rowcap = (int)v; }
  else if (k == "contact_sensor_maxmatch") d.contact_sensor_maxmatch = (int)v;
  else if (k == "iterations") d.iterations = (int)v;
  else if (k == "ls_iterations") d.ls_iterations = (int)v;
  else if (k == "tolerance") d.tolerance = (float)v;
  else if (k == "ls_tolerance") d.ls_tolerance = (float)v;
  else if (k == "timestep") d.timestep = (float)v;
  else if (k == "impratio") d.impratio = (float)v;
  else if (k == "ls_parallel_min_step") d.ls_parallel_min_step = (float)v;
  else if (k == "gravity_x") d.gravity[0] = (float)v;
  else if (k == "gravity_y") d.gravity[1] = (float)v;
  else if (k == "gravity_z") d.gravity[2] = (float)v;
  else if (k == "disableflags") d.disableflags = (int)v;
  else if (k == "integrator") {
    if ((int)v != B2S_INT_EULER && (int)v != B2S_INT_IMPLICITFAST) return fail("integrator must be euler or implicitfast");
    d.integrator = (int)v;
  } else if (k == "solver") {
    if ((int)v != B2S_SOL_NEWTON && (int)v != B2S_SOL_CG) return fail("solver must be newton or cg (PGS is not supported)");
    d.solver = (int)v;
  } else if (k == "cone") {
    if ((int)v != B2S_CONE_PYRAMIDAL && (int)v != B2S_CONE_ELLIPTIC) return fail("cone must be pyramidal or elliptic");
    d.cone = (int)v;
  } else return fail("unknown option " + k);
  return 0;
}
extern "C" int b2s_model_get_opt(const b2s_model* m, const char* key, double* v) {
  if (!m || !key || !v) return fail("null argument");
  const ModelDev& d = m->dev;
  const std::string k(key);
  if (k == "ls_parallel") *v = d.ls_parallel;
  else if (k == "stage_overlap") *v = m->overlap;
  else if (k == "solve_rowcap") *v = d.solve_rowcap;
  else if (k == "contact_sensor_maxmatch") *v = d.contact_sensor_maxmatch;
  else if (k == "iterations") *v = d.iterations;
  else if (k == "ls_iterations") *v = d.ls_iterations;
  else if (k == "tolerance") *v = d.tolerance;
  else if (k == "ls_tolerance") *v = d.ls_tolerance;
  else if (k == "timestep") *v = d.timestep;
  else if (k == "impratio") *v = d.impratio;
  else if (k == "ls_parallel_min_step") *v = d.ls_parallel_min_step;
  else if (k == "gravity_x") *v = d.gravity[0];
  else if (k == "gravity_y") *v = d.gravity[1];
  else if (k == "gravity_z") *v = d.gravity[2];
  else if (k == "disableflags") *v = d.disableflags;
  else if (k == "integrator") *v = d.integrator;
  else if (k == "cone") *v = d.cone;
  else if (k == "solver") *v = d.solver;
  else if (k == "npair") *v = d.npair;
  else if (k == "nvp") *v = d.nvp;
  else return fail("unknown option " + k);
  return 0;
}

extern "C" int b2s_model_expand(b2s_model* m, const char* name, int nworld) {
  if (!m || !name) return fail("null argument");
  auto it = m->index.find(name);
  if (it == m->index.end()) return fail(std::string("unknown model field ") + name);
  FieldRec& r = m->fields[it->second];
  if (!r.wf) return fail(std::string("model field ") + name + " cannot vary per world");
  if (r.wf->wstride != 0) return 0;   // already expanded
  if (r.elems == 0) return 0;
  {
    // the static-geom grid (terrain broadphase) is built once from the compile-time poses: a per-world pose of those geoms
    // would leave it stale, so refuse instead of colliding against the wrong boxes
    const std::string n(name);
    if (m->dev.site_split && (n == "site_pos" || n == "site_quat" || n == "body_pos" || n == "body_quat"))
      return fail("model field " + n + " cannot vary per world: this model's " + std::to_string(m->dev.nstatsite) + " static site frames are stored once (site_split)");
    if (m->dev.ngrid > 0 && (n == "geom_pos" || n == "geom_quat" || n == "body_pos" || n == "body_quat"))
      return fail("model field " + n + " cannot vary per world: this model's static terrain geoms are indexed by a broadphase grid built at put_model");
  }
  CU(cudaSetDevice(m->device));
  float* p = nullptr;
  CU(cudaMalloc(&p, (size_t)nworld * r.elems * sizeof(float)));
  m->allocs.push_back(p);
  const long long total = (long long)nworld * r.elems;
  k_tile<<<(int)((total + 255) / 256 > 4096 ? 4096 : (total + 255) / 256), 256>>>(p, r.wf->ptr, r.elems, total);
  b2s_count_launch();
  CU(cudaDeviceSynchronize());
  r.wf->ptr = p; r.wf->wstride = (int)r.elems;
  m->nexpanded++;
  // a per-world dof_frictionloss may be written later (domain randomisation, envs/mdp/events.py:231): switch the
  // friction-loss rows and the solver instantiation that understands them on
  if (std::string(name) == "dof_frictionloss") m->dev.fric = 1;
  return 0;
}

static void fill_info(const FieldRec& r, long long lead, bool broadcast, void* ptr, b2s_field_info* o) {
  o->ptr = ptr; o->dtype = r.dtype; o->ndim = r.ndim + 1;
  o->shape[0] = lead;
  for (int k = 0; k < r.ndim; k++) o->shape[k + 1] = r.shape[k];
  for (int k = r.ndim + 1; k < 4; k++) o->shape[k] = 1;
  long long st = 1;
  for (int k = o->ndim - 1; k >= 1; k--) { o->stride[k] = st; st *= o->shape[k]; }
  for (int k = o->ndim; k < 4; k++) o->stride[k] = 1;
  o->stride[0] = broadcast ? 0 : st;
}
extern "C" int b2s_model_field(b2s_model* m, const char* name, b2s_field_info* out) {
  if (!m || !name || !out) return fail("null argument");
  auto it = m->index.find(name);
  if (it == m->index.end()) return fail(std::string("unknown model field ") + name);
  const FieldRec& r = m->fields[it->second];
  if (r.wf) {
    const bool ex = r.wf->wstride != 0;
    fill_info(r, ex ? -1 : 1, !ex, (void*)r.wf->ptr, out);   // shape[0] = -1: caller substitutes nworld
  } else {
    out->ptr = r.ptr; out->dtype = r.dtype; out->ndim = r.ndim;
    long long st = 1;
    for (int k = 0; k < 4; k++) out->shape[k] = r.shape[k];
    for (int k = r.ndim - 1; k >= 0; k--) { out->stride[k] = st; st *= r.shape[k]; }
    for (int k = r.ndim; k < 4; k++) out->stride[k] = 1;
  }
  return 0;
}
extern "C" int b2s_model_num_fields(const b2s_model* m) { return m ? (int)m->fields.size() : 0; }
extern "C" const char* b2s_model_field_name(const b2s_model* m, int i) {
  return (m && i >= 0 && i < (int)m->fields.size()) ? m->fields[i].name.c_str() : nullptr;
}

// ------------------------------------------------------------------------------------------------ put_data
extern "C" int b2s_data_create(const b2s_model* m, const b2s_state_desc* init, int nworld, int nconmax, int njmax, b2s_data** out) {
  if (!m || !init || !out) return fail("null argument");
  if (nworld < 1) return fail("nworld must be positive");
  const ModelDev& md = m->dev;
  if (nconmax <= 0) nconmax = 64;
  if (njmax <= 0) njmax = 4 * nconmax + 2 * md.njnt;
  CU(cudaSetDevice(m->device));
  b2s_data* D = new b2s_data();
  D->device = m->device;
  DataDev& d = D->dev;
  memset(&d, 0, sizeof(d));
  d.nworld = nworld; d.nconmax = nconmax; d.njmax = njmax;
  const long long W = nworld, nb = md.nbody, nv = md.nv, nvp = md.nvp;
  struct Slot { const char* name; int dtype; void** dst; std::vector<long long> shape; long long shared = 0; };
  std::vector<Slot> slots = {
    {"qpos", B2S_F32, (void**)&d.qpos, {md.nq}}, {"qvel", B2S_F32, (void**)&d.qvel, {nv}}, {"ctrl", B2S_F32, (void**)&d.ctrl, {md.nu}},
    {"qacc_warmstart", B2S_F32, (void**)&d.qacc_warmstart, {nv}}, {"time", B2S_F32, (void**)&d.time, {}},
    {"xfrc_applied", B2S_F32, (void**)&d.xfrc_applied, {nb, 6}}, {"qfrc_applied", B2S_F32, (void**)&d.qfrc_applied, {nv}},
    {"mocap_pos", B2S_F32, (void**)&d.mocap_pos, {md.nmocap, 3}}, {"mocap_quat", B2S_F32, (void**)&d.mocap_quat, {md.nmocap, 4}},
    {"xpos", B2S_F32, (void**)&d.xpos, {nb, 3}}, {"xquat", B2S_F32, (void**)&d.xquat, {nb, 4}}, {"xmat", B2S_F32, (void**)&d.xmat, {nb, 3, 3}},
    {"xipos", B2S_F32, (void**)&d.xipos, {nb, 3}}, {"ximat", B2S_F32, (void**)&d.ximat, {nb, 3, 3}},
    {"xanchor", B2S_F32, (void**)&d.xanchor, {md.njnt, 3}}, {"xaxis", B2S_F32, (void**)&d.xaxis, {md.njnt, 3}},
    {"geom_xpos", B2S_F32, (void**)&d.geom_xpos, {md.ngeom, 3}}, {"geom_xmat", B2S_F32, (void**)&d.geom_xmat, {md.ngeom, 3, 3}},
    // dense: (nworld, nsite, .); split (H3): nstatsite frames once, then (nworld, nsite_world, .) -- see b2s_site_index
    {md.site_split ? "site_xpos_world" : "site_xpos", B2S_F32, (void**)&d.site_xpos, {md.nsite_world, 3}, md.site_split ? 3LL * md.nstatsite : 0},
    {md.site_split ? "site_xmat_world" : "site_xmat", B2S_F32, (void**)&d.site_xmat, {md.nsite_world, 3, 3}, md.site_split ? 9LL * md.nstatsite : 0},
    {"subtree_com", B2S_F32, (void**)&d.subtree_com, {nb, 3}}, {"cinert", B2S_F32, (void**)&d.cinert, {nb, 10}}, {"cdof", B2S_F32, (void**)&d.cdof, {nv, 6}},
    {"cvel", B2S_F32, (void**)&d.cvel, {nb, 6}}, {"cdof_dot", B2S_F32, (void**)&d.cdof_dot, {nv, 6}}, {"qM", B2S_F32, (void**)&d.qM, {nv, nvp}},
    {"qfrc_bias", B2S_F32, (void**)&d.qfrc_bias, {nv}}, {"qfrc_passive", B2S_F32, (void**)&d.qfrc_passive, {nv}},
    {"ten_length", B2S_F32, (void**)&d.ten_length, {md.ntendon}}, {"ten_velocity", B2S_F32, (void**)&d.ten_velocity, {md.ntendon}},
    {"actuator_length", B2S_F32, (void**)&d.actuator_length, {md.nu}}, {"actuator_velocity", B2S_F32, (void**)&d.actuator_velocity, {md.nu}},
    {"actuator_force", B2S_F32, (void**)&d.actuator_force, {md.nu}}, {"qfrc_actuator", B2S_F32, (void**)&d.qfrc_actuator, {nv}},
    {"qfrc_smooth", B2S_F32, (void**)&d.qfrc_smooth, {nv}}, {"qacc_smooth", B2S_F32, (void**)&d.qacc_smooth, {nv}},
    {"ncon", B2S_I32, (void**)&d.ncon, {}}, {"contact_dim", B2S_I32, (void**)&d.con_dim, {nconmax}}, {"contact_geom", B2S_I32, (void**)&d.con_geom, {nconmax, 2}},
    {"contact_efc_address", B2S_I32, (void**)&d.con_efc_address, {nconmax}}, {"contact_dist", B2S_F32, (void**)&d.con_dist, {nconmax}},
    {"contact_pos", B2S_F32, (void**)&d.con_pos, {nconmax, 3}}, {"contact_frame", B2S_F32, (void**)&d.con_frame, {nconmax, 3, 3}},
    {"contact_includemargin", B2S_F32, (void**)&d.con_includemargin, {nconmax}}, {"contact_friction", B2S_F32, (void**)&d.con_friction, {nconmax, 5}},
    {"contact_solref", B2S_F32, (void**)&d.con_solref, {nconmax, 2}}, {"contact_solimp", B2S_F32, (void**)&d.con_solimp, {nconmax, 5}},
    {"col_surv", B2S_I32, (void**)&d.col_surv, {md.col_defer ? B2S_MAX_SURVIVORS : 0}}, {"col_nsurv", B2S_I32, (void**)&d.col_nsurv, {md.col_defer ? 1 : 0}},
    {"col_pidx", B2S_I32, (void**)&d.col_pidx, {md.col_defer ? B2S_MAX_SURVIVORS : 0}}, {"cvx_ntask", B2S_I32, (void**)&d.cvx_ntask, {md.col_defer ? 4 : 0}}      /* [0] tasks emitted, [1] cursor of k_convex_tasks, [2] expensive tasks (queued from the end of the list) */,
    {"cvx_task", B2S_I32, (void**)&d.cvx_task, {md.col_defer ? B2S_MAX_SURVIVORS : 0, 8}}, {"cvx_stage", B2S_F32, (void**)&d.cvx_stage, {md.col_defer ? B2S_MAX_SURVIVORS : 0, 8}},
    {"col_axis", B2S_F32, (void**)&d.col_axis, {(md.col_convex && md.ngrid == 0) ? md.npair : 0, 3}},
    {"col_cost", B2S_I32, (void**)&d.col_cost, {(md.col_defer && md.ngrid == 0) ? md.npair : 0}},
    {"nefc", B2S_I32, (void**)&d.nefc, {}}, {"nl", B2S_I32, (void**)&d.nl, {}}, {"nf", B2S_I32, (void**)&d.nf, {}}, {"efc_type", B2S_I32, (void**)&d.efc_type, {njmax}}, {"efc_id", B2S_I32, (void**)&d.efc_id, {njmax}},
    {"efc_g", B2S_F32, (void**)&d.efc_g, {njmax, 6}}, {"efc_bp", B2S_I32, (void**)&d.efc_bp, {njmax}}, {"efc_pos", B2S_F32, (void**)&d.efc_pos, {njmax}}, {"efc_margin", B2S_F32, (void**)&d.efc_margin, {njmax}},
    {"efc_D", B2S_F32, (void**)&d.efc_D, {njmax}}, {"efc_vel", B2S_F32, (void**)&d.efc_vel, {njmax}}, {"efc_aref", B2S_F32, (void**)&d.efc_aref, {njmax}},
    {"efc_force", B2S_F32, (void**)&d.efc_force, {njmax}}, {"efc_frictionloss", B2S_F32, (void**)&d.efc_frictionloss, {njmax}},
    {"qacc", B2S_F32, (void**)&d.qacc, {nv}}, {"qfrc_constraint", B2S_F32, (void**)&d.qfrc_constraint, {nv}},
    {"solver_cost", B2S_F32, (void**)&d.solver_cost, {}}, {"solver_niter", B2S_I32, (void**)&d.solver_niter, {}},
    {"solve_order", B2S_I32, (void**)&d.solve_order, {B2S_SOLVE_CLASSES}}, {"solve_count", B2S_I32, (void**)&d.solve_count, {B2S_SOLVE_CLASSES}},
    {"cacc", B2S_F32, (void**)&d.cacc, {nb, 6}}, {"subtree_linvel", B2S_F32, (void**)&d.subtree_linvel, {nb, 3}},
    {"subtree_angmom", B2S_F32, (void**)&d.subtree_angmom, {nb, 3}}, {"sensordata", B2S_F32, (void**)&d.sensordata, {md.nsensordata}},
    {"snap_qpos", B2S_F32, (void**)&d.snap_qpos, {md.nq}}, {"snap_qvel", B2S_F32, (void**)&d.snap_qvel, {nv}}, {"snap_time", B2S_F32, (void**)&d.snap_time, {}},
  };
  // one arena; every field starts on a 256-byte boundary
  size_t total = 0;
  std::vector<size_t> offs;
  for (auto& s : slots) {
    long long n = 1;
    for (long long v : s.shape) n *= v;
    offs.push_back(total);
    size_t bytes = (size_t)(W * n + s.shared) * 4;
    if (bytes == 0) bytes = 4;
    total += (bytes + 255) & ~(size_t)255;
  }
  if (cudaMalloc(&D->arena, total) != cudaSuccess) { delete D; return fail("cudaMalloc of the data arena failed (" + std::to_string(total >> 20) + " MiB)"); }
  D->bytes = total;
  cudaMemset(D->arena, 0, total);
  for (size_t i = 0; i < slots.size(); i++) {
    *slots[i].dst = (char*)D->arena + offs[i];
    // a field with a shared block is exposed as two views: "<name>" = the per-world block, "<base>_static" = the shared one
    add_field(D->fields, D->index, slots[i].name, slots[i].dtype, (char*)*slots[i].dst + slots[i].shared * 4, nullptr, slots[i].shape);
    if (slots[i].shared) {
      std::string sn(slots[i].name); sn = sn.substr(0, sn.size() - 6) + "_static";      // ..._world -> ..._static
      std::vector<long long> sh = slots[i].shape; sh[0] = md.nstatsite;
      add_field(D->fields, D->index, sn.c_str(), slots[i].dtype, *slots[i].dst, nullptr, sh);
      D->fields.back().shared = 1;
    }
  }
  d.site_shared = md.site_split ? md.nstatsite : 0;
  // initial state, broadcast to every world
  auto bcast = [&](float* dst, const double* src, long long n) -> int {
    if (n == 0) return 0;
    std::vector<float> h((size_t)n);
    for (long long i = 0; i < n; i++) h[(size_t)i] = (float)src[i];
    float* tmp = nullptr;
    if (cudaMalloc(&tmp, (size_t)n * 4) != cudaSuccess) return -1;
    cudaMemcpy(tmp, h.data(), (size_t)n * 4, cudaMemcpyHostToDevice);
    const long long tot = W * n;
    k_tile<<<(int)((tot + 255) / 256 > 4096 ? 4096 : (tot + 255) / 256), 256>>>(dst, tmp, n, tot);
    b2s_count_launch();
    cudaDeviceSynchronize();
    cudaFree(tmp);
    return 0;
  };
  int rc = bcast(d.qpos, init->qpos, md.nq) | bcast(d.qvel, init->qvel, nv) | bcast(d.ctrl, init->ctrl, md.nu) |
           bcast(d.mocap_pos, init->mocap_pos, 3LL * md.nmocap) | bcast(d.mocap_quat, init->mocap_quat, 4LL * md.nmocap);
  double t = init->time;
  rc |= bcast(d.time, &t, 1);
  if (rc || cudaGetLastError() != cudaSuccess) { delete D; return fail("initialising the data arena failed"); }
  // raise the dynamic shared-memory limit of every stage once
  struct { const void* fn; size_t smem; } st[] = {
    {fn_position(), smem_position(md)}, {fn_dynamics(md), smem_dynamics(md)}, {fn_collision(md), smem_collision(md)},
    {fn_constraint(), smem_constraint(md)}, {fn_solve(md), smem_solve(md, d)}, {fn_sensors(), smem_sensors(md, d)},
    {fn_integrate(md), smem_integrate(md)}};
  for (auto& e : st) {
    if (e.smem > 227 * 1024) { delete D; return fail("model too large for the shared-memory working set (" + std::to_string(e.smem) + " B per CTA)"); }
    if (b2s_raise_smem(m->device, e.fn, e.smem) != 0) { delete D; return fail("cannot raise the dynamic shared-memory limit"); }
  }
  {
    size_t need = 0;
    const int rc2 = solve_prepare_all(md, d, &need, m->device);
    if (rc2) { delete D; return fail(rc2 == -1 ? "model too large for the solver's shared-memory working set (" + std::to_string(need) + " B per CTA)" : std::string("cannot raise the solver's dynamic shared-memory limit")); }
  }
  for (auto& e : D->ev) cudaEventCreate(&e);
  D->have_ev = true;
  if (cudaStreamCreateWithFlags(&D->side, cudaStreamNonBlocking) != cudaSuccess) { delete D; return fail("cannot create the side stream"); }
  {
    int lo = 0, hi = 0;
    cudaDeviceGetStreamPriorityRange(&lo, &hi);
    if (cudaStreamCreateWithPriority(&D->side_hi, cudaStreamNonBlocking, hi) != cudaSuccess) { delete D; return fail("cannot create the priority side stream"); }
  }
  for (cudaEvent_t* e : {&D->ev_fork, &D->ev_join, &D->ev_fork2, &D->ev_join2})
    if (cudaEventCreateWithFlags(e, cudaEventDisableTiming) != cudaSuccess) { delete D; return fail("cannot create the fork/join events"); }
  // frames of everything attached to static bodies, once
  launch_position(md, d, 0);
  launch_static_frames(md, d, 0);
  if (cudaDeviceSynchronize() != cudaSuccess) { delete D; return fail(std::string("initial position pass failed: ") + cudaGetErrorString(cudaGetLastError())); }
  *out = D;
  return 0;
}
// a launch that failed (bad configuration, shared-memory limit, ...) is reported, never silent: cudaPeekAtLastError is legal
// during stream capture and does not synchronise
static int launch_status(const char* what) {
  const cudaError_t e = cudaPeekAtLastError();
  if (e == cudaSuccess) return 0;
  cudaGetLastError();
  return fail(std::string(what) + ": kernel launch failed: " + cudaGetErrorString(e));
}
// Recompute the frames of geoms / sites on static bodies (after writing body_pos, body_quat, geom_pos, geom_quat,
// site_pos or site_quat of a static body, e.g. per-world terrain randomisation).  step/forward never touch them.
extern "C" int b2s_refresh_static(const b2s_model* m, b2s_data* d, void* stream) {
  if (!m || !d) return fail("null argument");
  launch_position(m->dev, d->dev, (cudaStream_t)stream);
  launch_static_frames(m->dev, d->dev, (cudaStream_t)stream);
  return launch_status("b2s_refresh_static");
}
extern "C" int b2s_data_destroy(b2s_data* d) { delete d; return 0; }
extern "C" int b2s_data_field(b2s_data* d, const char* name, b2s_field_info* out) {
  if (!d || !name || !out) return fail("null argument");
  auto it = d->index.find(name);
  if (it == d->index.end()) return fail(std::string("unknown data field ") + name);
  const FieldRec& r = d->fields[it->second];
  fill_info(r, r.shared ? 1 : d->dev.nworld, false, r.ptr, out);
  return 0;
}
// Dense constraint Jacobian (mjwarp's Data.efc.J) from the matrix-free rows: allocated on first use, refilled on
// every call.  Not part of step/forward and not graph-capturable (it may allocate).
extern "C" int b2s_efc_jacobian(const b2s_model* m, b2s_data* d, void* stream, b2s_field_info* out) {
  if (!m || !d || !out) return fail("null argument");
  CU(cudaSetDevice(d->device));
  const size_t n = (size_t)d->dev.nworld * d->dev.njmax * m->dev.nvp;
  if (!d->efc_J) CU(cudaMalloc(&d->efc_J, n * sizeof(float)));
  launch_efc_jacobian(m->dev, d->dev, d->efc_J, (cudaStream_t)stream);
  CU(cudaGetLastError());
  out->ptr = d->efc_J; out->dtype = B2S_F32; out->ndim = 3;
  out->shape[0] = d->dev.nworld; out->shape[1] = d->dev.njmax; out->shape[2] = m->dev.nvp; out->shape[3] = 1;
  out->stride[2] = 1; out->stride[1] = m->dev.nvp; out->stride[0] = (long long)d->dev.njmax * m->dev.nvp; out->stride[3] = 1;
  return 0;
}
extern "C" int b2s_data_num_fields(const b2s_data* d) { return d ? (int)d->fields.size() : 0; }
extern "C" const char* b2s_data_field_name(const b2s_data* d, int i) {
  return (d && i >= 0 && i < (int)d->fields.size()) ? d->fields[i].name.c_str() : nullptr;
}
extern "C" int64_t b2s_data_bytes(const b2s_data* d) { return d ? (int64_t)d->bytes : 0; }

// ------------------------------------------------------------------------------------------ forward / step
// One forward pass (and, with `integrate`, one step) as a DAG on the caller's stream `st` plus the data's side stream:
//
//   position -> dynamics ------------> constraint -> solve -> integrate ---> (next position)
//           \-> collision (side) --/                     \-> sensors (side) -/
//
// collision needs only the geom frames of `position`; sensors need the solve and the state snapshot k_dynamics took
// (DataDev::snap_*), so they run beside the integrator, which advances qpos / qvel / time.  Every stage kernel leaves
// most of the machine idle at 4096 worlds (one warp per world, 18-40 % warps active), so the branches fill each other's
// gaps.  The side stream is forked from and joined back into `st` with events: under stream capture this records a DAG
// in the same graph, eagerly it is plain stream concurrency.  Results are identical to the serial order.
// NVTX ranges around the passes and the stage launches (host-side markers: in a timeline they bracket the eager launches or
// the graph capture).  Off unless B2S_NVTX=1 is set when the library is loaded.
static const bool g_nvtx = [] { const char* e = getenv("B2S_NVTX"); return e && e[0] == '1'; }();
struct NvtxRange {
  explicit NvtxRange(const char* name) { if (g_nvtx) nvtxRangePushA(name); }
  ~NvtxRange() { if (g_nvtx) nvtxRangePop(); }
};
// the solver's launch for worlds with many rows (normally empty) runs beside the main one, on the priority stream
static void solve_stage(const b2s_model* M, b2s_data* D, cudaStream_t st, bool fork) {
  const ModelDev& m = M->dev; const DataDev& d = D->dev;
  NvtxRange r("solve");
  if (fork && solve_has_second_launch(m, d)) {
    cudaEventRecord(D->ev_fork, st); cudaStreamWaitEvent(D->side_hi, D->ev_fork, 0);
    launch_solve(m, d, st, D->side_hi);
    cudaEventRecord(D->ev_join, D->side_hi);
    cudaStreamWaitEvent(st, D->ev_join, 0);
  } else launch_solve(m, d, st);
}
static void run_pass(const b2s_model* M, b2s_data* D, cudaStream_t st, bool integrate) {
  const ModelDev& m = M->dev; const DataDev& d = D->dev;
  const bool fork = M->overlap == 1;
  NvtxRange pass(integrate ? "b2s step" : "b2s forward");
  { NvtxRange r("position"); launch_position(m, d, st); }
  if (fork) {
    cudaEventRecord(D->ev_fork, st); cudaStreamWaitEvent(D->side, D->ev_fork, 0);
    { NvtxRange r("collision (side stream)"); launch_collision(m, d, D->side); }
    cudaEventRecord(D->ev_join, D->side);
    { NvtxRange r("dynamics"); launch_dynamics(m, d, st); }
    cudaStreamWaitEvent(st, D->ev_join, 0);
  } else {
    { NvtxRange r("dynamics"); launch_dynamics(m, d, st); }
    { NvtxRange r("collision"); launch_collision(m, d, st); }
  }
  { NvtxRange r("constraint"); launch_constraint(m, d, st); }
  solve_stage(M, D, st, M->overlap != 0);
  if (fork && integrate) {
    cudaEventRecord(D->ev_fork2, st); cudaStreamWaitEvent(D->side, D->ev_fork2, 0);
    { NvtxRange r("sensors (side stream)"); launch_sensors(m, d, D->side); }
    cudaEventRecord(D->ev_join2, D->side);
    { NvtxRange r("integrate"); launch_integrate(m, d, st); }
    cudaStreamWaitEvent(st, D->ev_join2, 0);
  } else {
    { NvtxRange r("sensors"); launch_sensors(m, d, st); }
    if (integrate) { NvtxRange r("integrate"); launch_integrate(m, d, st); }
  }
}
extern "C" int b2s_forward(const b2s_model* m, b2s_data* d, void* stream) {
  if (!m || !d) return fail("null argument");
  run_pass(m, d, (cudaStream_t)stream, false);
  return launch_status("b2s_forward");
}
extern "C" int b2s_step_n(const b2s_model* m, b2s_data* d, int nsub, void* stream) {
  if (!m || !d) return fail("null argument");
  for (int i = 0; i < nsub; i++) run_pass(m, d, (cudaStream_t)stream, true);
  return launch_status("b2s_step_n");
}
extern "C" int b2s_step(const b2s_model* m, b2s_data* d, void* stream) { return b2s_step_n(m, d, 1, stream); }

extern "C" int b2s_step_host(const b2s_model* m, b2s_data* d, const float* ctrl_host, int nsub, float* qpos_host, float* qvel_host,
                             float* sensordata_host, void* stream) {
  if (!m || !d) return fail("null argument");
  cudaStream_t st = (cudaStream_t)stream;
  const ModelDev& md = m->dev;
  const size_t W = (size_t)d->dev.nworld;
  CU(cudaSetDevice(m->device));
  if (ctrl_host && md.nu) CU(cudaMemcpyAsync(d->dev.ctrl, ctrl_host, W * md.nu * 4, cudaMemcpyHostToDevice, st));
  if (b2s_step_n(m, d, nsub, stream) != 0) return -1;
  if (qpos_host) CU(cudaMemcpyAsync(qpos_host, d->dev.qpos, W * md.nq * 4, cudaMemcpyDeviceToHost, st));
  if (qvel_host) CU(cudaMemcpyAsync(qvel_host, d->dev.qvel, W * md.nv * 4, cudaMemcpyDeviceToHost, st));
  if (sensordata_host && md.nsensordata) CU(cudaMemcpyAsync(sensordata_host, d->dev.sensordata, W * md.nsensordata * 4, cudaMemcpyDeviceToHost, st));
  CU(cudaStreamSynchronize(st));
  return 0;
}

extern "C" int b2s_profile_step(const b2s_model* m, b2s_data* d, void* stream, int reps, const char** names, float* ms, int cap) {
  if (!m || !d) return fail("null argument");
  static const char* nm[] = {"position", "dynamics", "collision", "constraint", "solve", "sensors", "integrate"};
  typedef void (*fnT)(const ModelDev&, const DataDev&, cudaStream_t);
  fnT fn[] = {launch_position, launch_dynamics, launch_collision, launch_constraint,
              [](const ModelDev& a, const DataDev& b, cudaStream_t s) { launch_solve(a, b, s); }, launch_sensors, launch_integrate};
  cudaStream_t st = (cudaStream_t)stream;
  float acc[7] = {0, 0, 0, 0, 0, 0, 0};
  for (int r = 0; r < reps; r++) {
    cudaEventRecord(d->ev[0], st);
    for (int k = 0; k < 7; k++) {
      if (k == 4) solve_stage(m, d, st, m->overlap != 0); else fn[k](m->dev, d->dev, st);
      cudaEventRecord(d->ev[k + 1], st);
    }
    CU(cudaStreamSynchronize(st));
    for (int k = 0; k < 7; k++) { float t = 0; cudaEventElapsedTime(&t, d->ev[k], d->ev[k + 1]); acc[k] += t; }
  }
  int n = cap < 7 ? cap : 7;
  for (int k = 0; k < n; k++) { names[k] = nm[k]; ms[k] = acc[k] / (reps > 0 ? reps : 1); }
  return n;
}

// ------------------------------------------------------------------------------------------ allocation-free re-solve
// b2s_solve_noalloc (include/b200sim.h; no counterpart in the reference): the
// constraint solve (k_solve) on the rows the last forward / step left in `d`, with the solver's scheduling scratch -- the
// cost-ordered world queues k_constraint normally fills in the data arena -- in a file-scope device array instead.  For callers
// that may not allocate (CUDA graph capture, tight serving loops): no cudaMalloc / cudaFree / allocating library call here.
//
// The array is shared by every call on a device and keeps the last call's contents: the per-class counters are cleared at the
// start of each call (the order slots need no clearing: only [0, count) of a class is read), and two calls whose GPU work
// overlaps (different streams) race on it.  Every GPU that loads the library holds the array until exit
// (B2S_NOALLOC_MAX_WORLDS x B2S_SOLVE_CLASSES ints + the counters = 4 MiB + 64 B).
static __device__ __align__(256) int g_noalloc_scratch[B2S_SOLVE_CLASSES * (B2S_NOALLOC_MAX_WORLDS + 1)];

// per world: the cost class k_constraint would pick (previous Newton iteration count x rows), appended to that class's queue
__global__ void k_noalloc_queue(const DataDev d, int* order, int* count) {
  const int w = blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= d.nworld) return;
  const int nefc = d.nefc[w];
  const int cls = nefc == 0 ? 0 : min(B2S_SOLVE_CLASSES - 1, 1 + (1 + min(d.solver_niter[w], 15)) * (32 + nefc) / 64);
  order[(size_t)cls * d.nworld + atomicAdd(&count[cls], 1)] = w;
}

// device address of the scratch array, looked up once per device with that device current.  In host code the array's NAME is
// a host-side stand-in, not a device address: only cudaGetSymbolAddress gives the pointer a kernel may be handed.
static int noalloc_scratch_ptr(int device, int** out) {
  static std::atomic<int*> ptrs[64];
  if (device < 0 || device >= 64) return fail("b2s_solve_noalloc: device index out of range");
  int* p = ptrs[device].load(std::memory_order_acquire);
  if (!p) {
    CU(cudaSetDevice(device));
    void* sym = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&sym, g_noalloc_scratch);
    if (e != cudaSuccess || !sym) return fail(std::string("b2s_solve_noalloc: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e));
    p = (int*)sym;
    ptrs[device].store(p, std::memory_order_release);
  }
  *out = p;
  return 0;
}

extern "C" int b2s_solve_noalloc(const b2s_model* m, b2s_data* d, void* stream) {
  if (!m || !d) return fail("null argument");
  if (d->dev.nworld > B2S_NOALLOC_MAX_WORLDS)
    return fail("b2s_solve_noalloc: nworld " + std::to_string(d->dev.nworld) + " exceeds the static scratch capacity of " +
                std::to_string(B2S_NOALLOC_MAX_WORLDS) + " worlds (B2S_NOALLOC_MAX_WORLDS); use b2s_forward");
  int* scratch = nullptr;
  if (noalloc_scratch_ptr(d->device, &scratch) != 0) return -1;
  cudaStream_t st = (cudaStream_t)stream;
  int* count = scratch;                           // [B2S_SOLVE_CLASSES]
  int* order = scratch + B2S_SOLVE_CLASSES;       // [B2S_SOLVE_CLASSES][nworld]
  CU(cudaMemsetAsync(count, 0, B2S_SOLVE_CLASSES * sizeof(int), st));
  DataDev dv = d->dev;                            // same inputs / outputs, the queue in the static array
  dv.solve_order = order; dv.solve_count = count;
  k_noalloc_queue<<<(dv.nworld + 127) / 128, 128, 0, st>>>(dv, order, count);
  b2s_count_launch();
  launch_solve(m->dev, dv, st);
  return launch_status("b2s_solve_noalloc");
}
extern "C" int b2s_noalloc_max_worlds(void) { return B2S_NOALLOC_MAX_WORLDS; }
