// b2s_types.h -- device-side model/data descriptors shared by all kernels.
//
// Layout in HBM: every field is its own array (struct of
// arrays), world-major, fp32/int32: field[world][...].  One warp owns one
// world, so a warp's loads of a field are contiguous runs (coalesced), and
// every field is directly a zero-copy torch tensor of the shape the
// reference exposes (src/mjlab/sim/sim_data.py:21).
#pragma once
#include <stdint.h>

#define B2S_WARPS_PER_CTA 4
#define B2S_MAX_NV 56          // register-tiled Hessian: 7 x 7 tiles of up to 8 x 8
#define B2S_GRID_MIN_GEOMS 64   // static non-plane geoms beyond this count are indexed by a uniform grid
#define B2S_MAX_SURVIVORS 160    // broadphase survivors (candidate geom pairs) kept per world and step
#define B2S_SITE_SPLIT_MIN 1024  // static sites: from this many on their frames are stored once instead of per world
#define B2S_MINVAL 1e-15f
#define B2S_MINMU 1e-5f

// A model field that may be per-world (after expand_model_fields): element
// (w, i) lives at ptr[(w * wstride) + i]; wstride == 0 while un-expanded.
template <typename T>
struct WField {
  const T* ptr;
  int wstride;
  __host__ __device__ const T* at(int w) const { return ptr + (size_t)w * wstride; }
};

struct ModelDev {
  // sizes
  int nq, nv, nvp, nu, nbody, njnt, ngeom, nsite, nsensor, nsensordata, nmocap;
  int nlevel, npair, ncolgeom, maxchain;
  int nchain;                       // total length of body_chain (= body_chain_adr[nbody])
  int ref_root;                     // root body of the first kinematic tree: its subtree COM is the constraint reference point
  // options
  float timestep, gravity[3], magnetic[3], tolerance, ls_tolerance, impratio, ls_parallel_min_step, meaninertia;
  int fric;                         // 1: the model can have friction-loss rows (dof_frictionloss > 0 somewhere, or the field is per-world)
  int integrator, cone, solver, iterations, ls_iterations, disableflags, ls_parallel, contact_sensor_maxmatch;
  // ---- bodies
  const int *body_parentid, *body_rootid, *body_mocapid, *body_jntnum, *body_jntadr, *body_dofnum, *body_dofadr;
  const int *body_subtree_end;      // bodies are in DFS pre-order: subtree(b) = [b, body_subtree_end[b])
  const int *body_chain_adr;        // CSR over ancestor-or-self dofs of each body, root first  [nbody+1]
  const int *body_chain;            // dof ids
  const int *level_adr, *level_body; // bodies grouped by tree depth (level 0 = world) [nlevel+1], [nbody]
  const unsigned char *body_dofmask; // [nbody][nv] 1 if dof is an ancestor-or-self dof of the body
  const unsigned char *body_ancmask; // [nbody][nbody] 1 if column body is an ancestor-or-self of row body
  WField<float> body_pos, body_quat, body_ipos, body_iquat, body_mass, body_inertia;
  const float *body_subtreemass, *body_invweight0;
  // ---- joints / dofs
  const int *jnt_type, *jnt_qposadr, *jnt_dofadr, *jnt_bodyid, *jnt_limited;
  const float *jnt_pos, *jnt_axis, *jnt_margin, *jnt_solref, *jnt_solimp;
  WField<float> jnt_range, jnt_stiffness, qpos0, qpos_spring;
  const int *dof_bodyid, *dof_jntid, *dof_parentid;
  WField<float> dof_armature, dof_damping, dof_frictionloss;
  const float *dof_invweight0;
  // ---- geoms
  const int *geom_type, *geom_bodyid, *geom_condim, *geom_priority;
  const float *geom_size, *geom_solmix, *geom_solref, *geom_solimp, *geom_margin, *geom_gap, *geom_rbound;
  WField<float> geom_pos, geom_quat, geom_friction, geom_rgba;
  const int *geom_contype, *geom_conaffinity;
  // uniform xy grid over the static (world-welded) non-plane geoms, used when there are many of them (terrain)
  int ngrid, ndyngeom, grid_nx, grid_ny;
  float grid_x0, grid_y0, grid_inv, grid_margin;
  const int *grid_adr, *grid_geom, *dyn_geom;   // CSR cell -> static geom ids; dynamic geoms that can collide
  const float* grid_aabb;                       // [ngeom][6] world AABB (min xyz, max xyz) of grid geoms
  // geoms / sites on moving bodies (recomputed every step) and on static, world-welded bodies (computed once at
  // put_data and by b2s_refresh_static: thousands of terrain geoms would otherwise dominate the position stage)
  int nposegeom, nposesite, nstatgeom, nstatsite;
  const int *pose_geom, *pose_site, *stat_geom, *stat_site;
  // Site frames (mjlab adds one visualisation site PER ENV to the world body, an O(nworld^2) array in the
  // reference's layout).  site_slot[s] >= 0: slot of a per-world site; < 0: ~slot of a static site whose frame is stored ONCE
  // (site_split = 1, chosen at put_model when there are many static sites).  Dense models: site_slot[s] = s, site_split = 0.
  int site_split, nsite_world;      // nsite_world = sites stored per world (nsite when dense)
  const int* site_slot;
  const int *pair_geom;             // [npair][2] after type ordering (type1 <= type2), sorted by (min id, max id)
  // ---- sites
  const int *site_bodyid;
  WField<float> site_pos, site_quat;
  // ---- fixed tendons (wraps flattened to dof / qpos addresses) and actuators
  int col_maxc;      // narrowphase variant: most contacts per geom pair (2, 4 or 8)
  int col_convex;    // 1: some pair needs the GJK/EPA, plane-mesh or height-field colliders
  int col_hf;        // 1: the model has height-field pairs (0: the second collision phase is instantiated without GJK / EPA)
  int solve_rowcap;  // rows the DEFAULT solver variant's shared-memory layout holds (option "solve_rowcap", default 128): worlds with more rows go to a second launch with room for njmax
  int col_defer;     // 1: GJK/EPA pairs between two convex cores exist: they run as (world, pair) tasks of a second kernel (k_convex_tasks)
  const int *geom_dataid, *mesh_vertadr, *mesh_vertnum, *hfield_nrow, *hfield_ncol, *hfield_adr;
  const float *mesh_vert, *hfield_size, *hfield_data;
  const float *geom_bcen, *geom_bsize, *geom_brad;   // broadphase volume: local box (centre offset, half sizes) and its radius
  int ntendon, nwrap;
  const int *tendon_adr, *tendon_num, *wrap_dof, *wrap_qpos;
  const float *wrap_coef, *tendon_stiffness, *tendon_damping, *tendon_lengthspring;
  const int *actuator_trntype;
  const int *actuator_trnid, *actuator_gaintype, *actuator_biastype, *actuator_ctrllimited, *actuator_forcelimited;
  const float *actuator_gear, *actuator_ctrlrange, *actuator_forcerange;
  WField<float> actuator_gainprm, actuator_biasprm;   // 3 per actuator
  // ---- sensors
  const int *sensor_type, *sensor_objtype, *sensor_objid, *sensor_reftype, *sensor_refid, *sensor_intprm, *sensor_dim, *sensor_adr;
  const float *sensor_cutoff;
  int need_subtreevel;
  int need_concache;   // some sensor is a contact sensor: k_sensors stages the world's contacts in shared memory
  int need_ray;        // some sensor is a rangefinder
  int need_cfrc;       // some sensor is a force / torque sensor: k_sensors runs the force half of rne_postconstraint (needs the contact cache)
};

#define B2S_SOLVE_CLASSES 16
struct DataDev {
  int nworld, nconmax, njmax;
  // state / inputs
  float *qpos, *qvel, *ctrl, *qacc_warmstart, *time, *xfrc_applied, *qfrc_applied, *mocap_pos, *mocap_quat;
  // position stage
  float *xpos, *xquat, *xmat, *xipos, *ximat, *xanchor, *xaxis, *geom_xpos, *geom_xmat, *site_xpos, *site_xmat;
  float *subtree_com, *cinert, *cdof;
  // dynamics stage
  float *cvel, *cdof_dot, *qM, *qfrc_bias, *qfrc_passive, *ten_length, *ten_velocity, *actuator_length, *actuator_velocity, *actuator_force;
  float *qfrc_actuator, *qfrc_smooth, *qacc_smooth;
  // collision
  int *ncon, *con_dim, *con_geom, *con_efc_address;
  float *con_dist, *con_pos, *con_frame, *con_includemargin, *con_friction, *con_solref, *con_solimp;
  // deferred convex pairs (ModelDev::col_defer): survivor list of each world [B2S_MAX_SURVIVORS] (+ count, + pair-table index),
  // global task list (world, survivor index) with its counter, and one staged result (CvxStage, 8 floats) per survivor
  unsigned* col_surv; int *col_nsurv, *col_pidx, *cvx_ntask, *cvx_task; float* cvx_stage;
  float* col_axis;   // [npair][3] per world, convex models only: last separating direction of each table pair (k_convex.cuh gjk_epa)
  int* col_cost;     // [npair] per world, same models: GJK iterations (+ 64 if EPA ran) of the pair's last task: expensive pairs are queued first
  // constraints
  int *nefc, *nl, *nf, *efc_type, *efc_id, *efc_bp;   // rows [0, nl) act on one dof: nf friction-loss rows, then the limits;   // efc_bp: contact rows body1 | body2 << 16, limit rows -1 - dof
  float *efc_g;                                  // [njmax][6] row wrench about the reference point (limit rows: g[0] = +-1)
  float *efc_pos, *efc_margin, *efc_D, *efc_vel, *efc_aref, *efc_force, *efc_frictionloss;
  // solver
  float *qacc, *qfrc_constraint, *solver_cost;
  int *solver_niter;
  // scheduling only (never affects results): worlds binned by expected solve cost (previous Newton iteration count x rows), so that the
  // solver's CTAs start the expensive worlds first; solve_order[class][nworld], solve_count[class]
  int *solve_order, *solve_count;
  // post
  float *cacc, *subtree_linvel, *subtree_angmom, *sensordata;
  // qpos / qvel / time as k_dynamics saw them: k_sensors reads these copies, so that it can run beside k_integrate (which
  // advances the originals) on a forked graph branch
  float *snap_qpos, *snap_qvel, *snap_time;
  // element index (in sites) of site s of world w in site_xpos / site_xmat: [shared static block | nworld x nsite_world]
  int site_shared;                  // static sites stored once ahead of the per-world block (0 when dense)
};
__host__ __device__ inline size_t b2s_site_index(const ModelDev& m, const DataDev& d, int w, int s) {
  const int slot = m.site_slot[s];
  return slot < 0 ? (size_t)(~slot) : (size_t)d.site_shared + (size_t)w * m.nsite_world + slot;
}
