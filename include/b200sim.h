/* b200sim.h -- C ABI of the B200-native batched rigid-body step backend.
 *
 * This is the drop-in boundary beneath mjlab's `Simulation`
 * (reference: src/mjlab/sim/sim.py:97-199).  Each entry point replaces one call
 * the reference makes into its un-vendored physics dependency `mujoco_warp`
 * (pinned at google-deepmind/mujoco_warp @ e605c406, pyproject.toml:108):
 *
 *   b2s_model_create   <- mjwarp.put_model(mjm)                  sim.py:116
 *   b2s_model_set_opt  <- model.opt.ls_parallel = ... etc.       sim.py:117-118
 *   b2s_model_expand   <- expand_model_fields(model, nworld, f)  sim/randomization.py:20-54
 *   b2s_data_create    <- mjwarp.put_data(mjm, mjd, nworld, nconmax, njmax)  sim.py:120-126
 *   b2s_forward        <- mjwarp.forward(m, d)                   sim.py:146,191
 *   b2s_step           <- mjwarp.step(m, d)                      sim.py:143,199
 *   b2s_field          <- wp.to_torch(array) zero-copy views     sim/sim_data.py:21
 *
 * Conventions: plain pointers and sizes, no exceptions across the ABI, every
 * function returns 0 on success and a negative code on failure with a message
 * available from b2s_last_error().  `step`/`forward` never allocate, never
 * synchronise with the host and launch only on the stream they are given, so
 * they can be captured into a CUDA graph (sim.py:138-147).
 *
 * Model description: a flat struct of host arrays named after MuJoCo's MjModel
 * fields.  Real-valued arrays are double (MjModel's mjtNum); the engine
 * converts to fp32 at upload.  Array extents are given in the trailing comment
 * of each member as an expression of the size members (the Python binding
 * parses this header to build its ctypes mirror -- keep one member per line).
 */
#ifndef B200SIM_H_
#define B200SIM_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- enums (values follow MuJoCo's mjt* enums where those are stable) ---- */
enum { B2S_JNT_FREE = 0, B2S_JNT_BALL = 1, B2S_JNT_SLIDE = 2, B2S_JNT_HINGE = 3 };
enum { B2S_GEOM_PLANE = 0, B2S_GEOM_HFIELD = 1, B2S_GEOM_SPHERE = 2, B2S_GEOM_CAPSULE = 3,
       B2S_GEOM_ELLIPSOID = 4, B2S_GEOM_CYLINDER = 5, B2S_GEOM_BOX = 6, B2S_GEOM_MESH = 7 };
enum { B2S_INT_EULER = 0, B2S_INT_RK4 = 1, B2S_INT_IMPLICIT = 2, B2S_INT_IMPLICITFAST = 3 };
enum { B2S_CONE_PYRAMIDAL = 0, B2S_CONE_ELLIPTIC = 1 };
enum { B2S_SOL_PGS = 0, B2S_SOL_CG = 1, B2S_SOL_NEWTON = 2 };
enum { B2S_OBJ_UNKNOWN = 0, B2S_OBJ_BODY = 1, B2S_OBJ_XBODY = 2, B2S_OBJ_JOINT = 3, B2S_OBJ_DOF = 4,
       B2S_OBJ_GEOM = 5, B2S_OBJ_SITE = 6, B2S_OBJ_TENDON = 18, B2S_OBJ_ACTUATOR = 19 };
enum { B2S_BIAS_NONE = 0, B2S_BIAS_AFFINE = 1 };
enum { B2S_TRN_JOINT = 0, B2S_TRN_TENDON = 3 };
enum { B2S_GAIN_FIXED = 0, B2S_GAIN_AFFINE = 1 };
/* sensor types: own numbering (MuJoCo's mjtSensor values moved between
 * releases); the Python binding maps by NAME from a real MjModel. */
enum { B2S_SENS_ACCELEROMETER = 1, B2S_SENS_VELOCIMETER = 2, B2S_SENS_GYRO = 3,
       B2S_SENS_JOINTPOS = 9, B2S_SENS_JOINTVEL = 10, B2S_SENS_TENDONPOS = 11, B2S_SENS_TENDONVEL = 12,
       B2S_SENS_ACTUATORPOS = 13, B2S_SENS_ACTUATORVEL = 14, B2S_SENS_ACTUATORFRC = 15, B2S_SENS_JOINTACTFRC = 16,
       B2S_SENS_FRAMEPOS = 25, B2S_SENS_FRAMEQUAT = 26, B2S_SENS_FRAMEXAXIS = 27, B2S_SENS_FRAMEYAXIS = 28, B2S_SENS_FRAMEZAXIS = 29,
       B2S_SENS_FRAMELINVEL = 30, B2S_SENS_FRAMEANGVEL = 31, B2S_SENS_FRAMELINACC = 32, B2S_SENS_FRAMEANGACC = 33,
       B2S_SENS_CLOCK = 45, B2S_SENS_MAGNETOMETER = 6, B2S_SENS_TENDONACTFRC = 17,
       B2S_SENS_JOINTLIMITPOS = 20, B2S_SENS_JOINTLIMITVEL = 21, B2S_SENS_JOINTLIMITFRC = 22,
       B2S_SENS_E_POTENTIAL = 43, B2S_SENS_E_KINETIC = 44, B2S_SENS_FORCE = 4, B2S_SENS_TORQUE = 5, B2S_SENS_RANGEFINDER = 7,
       B2S_SENS_SUBTREECOM = 34, B2S_SENS_SUBTREELINVEL = 35, B2S_SENS_SUBTREEANGMOM = 36,
       B2S_SENS_CONTACT = 42 };
/* contact-sensor data bits (intprm[0]) and reduce modes (intprm[1]);
 * reference: src/mjlab/sensor/contact_sensor.py:16-41,480-482 */
enum { B2S_CON_FOUND = 1, B2S_CON_FORCE = 2, B2S_CON_TORQUE = 4, B2S_CON_DIST = 8,
       B2S_CON_POS = 16, B2S_CON_NORMAL = 32, B2S_CON_TANGENT = 64 };
enum { B2S_REDUCE_NONE = 0, B2S_REDUCE_MINDIST = 1, B2S_REDUCE_MAXFORCE = 2, B2S_REDUCE_NETFORCE = 3 };
/* disable flags (subset of mjtDisableBit that this path honours) */
enum { B2S_DSBL_CONSTRAINT = 1 << 0, B2S_DSBL_FRICTIONLOSS = 1 << 2, B2S_DSBL_LIMIT = 1 << 3, B2S_DSBL_CONTACT = 1 << 4,
       B2S_DSBL_PASSIVE = 1 << 5, B2S_DSBL_GRAVITY = 1 << 6, B2S_DSBL_CLAMPCTRL = 1 << 7,
       B2S_DSBL_WARMSTART = 1 << 8, B2S_DSBL_FILTERPARENT = 1 << 9, B2S_DSBL_ACTUATION = 1 << 10,
       B2S_DSBL_REFSAFE = 1 << 11, B2S_DSBL_EULERDAMP = 1 << 13 };

/* ---- model description (host side; mirrors MjModel field names) ---- */
typedef struct b2s_model_desc {
  /* sizes */
  int32_t nq;
  int32_t nv;
  int32_t nu;
  int32_t na;
  int32_t nbody;
  int32_t njnt;
  int32_t ngeom;
  int32_t nsite;
  int32_t nsensor;
  int32_t nsensordata;
  int32_t nexclude;
  int32_t nmocap;
  int32_t ntendon;
  int32_t nwrap;
  int32_t nmesh;
  int32_t nmeshvert;
  int32_t nhfield;
  int32_t nhfielddata;
  /* options (MjModel.opt; reference MujocoCfg.apply, sim.py:64-76) */
  double opt_timestep;
  double opt_gravity[3];
  double opt_magnetic[3];
  double opt_tolerance;
  double opt_ls_tolerance;
  double opt_impratio;
  double opt_ls_parallel_min_step;
  double stat_meaninertia;
  int32_t opt_integrator;
  int32_t opt_cone;
  int32_t opt_solver;
  int32_t opt_iterations;
  int32_t opt_ls_iterations;
  int32_t opt_disableflags;
  int32_t opt_ls_parallel;
  int32_t opt_contact_sensor_maxmatch;
  /* bodies */
  const int32_t* body_parentid;   /* nbody */
  const int32_t* body_rootid;     /* nbody */
  const int32_t* body_weldid;     /* nbody */
  const int32_t* body_mocapid;    /* nbody */
  const int32_t* body_jntnum;     /* nbody */
  const int32_t* body_jntadr;     /* nbody */
  const int32_t* body_dofnum;     /* nbody */
  const int32_t* body_dofadr;     /* nbody */
  const double* body_pos;         /* nbody*3 */
  const double* body_quat;        /* nbody*4 */
  const double* body_ipos;        /* nbody*3 */
  const double* body_iquat;       /* nbody*4 */
  const double* body_mass;        /* nbody */
  const double* body_subtreemass; /* nbody */
  const double* body_inertia;     /* nbody*3 */
  const double* body_invweight0;  /* nbody*2 */
  /* joints */
  const int32_t* jnt_type;        /* njnt */
  const int32_t* jnt_qposadr;     /* njnt */
  const int32_t* jnt_dofadr;      /* njnt */
  const int32_t* jnt_bodyid;      /* njnt */
  const int32_t* jnt_limited;     /* njnt */
  const double* jnt_pos;          /* njnt*3 */
  const double* jnt_axis;         /* njnt*3 */
  const double* jnt_stiffness;    /* njnt */
  const double* jnt_range;        /* njnt*2 */
  const double* jnt_margin;       /* njnt */
  const double* jnt_solref;       /* njnt*2 */
  const double* jnt_solimp;       /* njnt*5 */
  const double* qpos0;            /* nq */
  const double* qpos_spring;      /* nq */
  /* dofs */
  const int32_t* dof_bodyid;      /* nv */
  const int32_t* dof_jntid;       /* nv */
  const int32_t* dof_parentid;    /* nv */
  const double* dof_armature;     /* nv */
  const double* dof_damping;      /* nv */
  const double* dof_frictionloss; /* nv */
  const double* dof_invweight0;   /* nv */
  /* geoms */
  const int32_t* geom_type;       /* ngeom */
  const int32_t* geom_bodyid;     /* ngeom */
  const int32_t* geom_contype;    /* ngeom */
  const int32_t* geom_conaffinity;/* ngeom */
  const int32_t* geom_condim;     /* ngeom */
  const int32_t* geom_priority;   /* ngeom */
  const double* geom_size;        /* ngeom*3 */
  const double* geom_pos;         /* ngeom*3 */
  const double* geom_quat;        /* ngeom*4 */
  const double* geom_friction;    /* ngeom*3 */
  const double* geom_solmix;      /* ngeom */
  const double* geom_solref;      /* ngeom*2 */
  const double* geom_solimp;      /* ngeom*5 */
  const double* geom_margin;      /* ngeom */
  const double* geom_gap;         /* ngeom */
  const double* geom_rbound;      /* ngeom */
  const int32_t* geom_dataid;     /* ngeom */
  /* convex meshes (collision uses the convex hull; vertices in the geom frame) and height fields */
  const int32_t* mesh_vertadr;    /* nmesh */
  const int32_t* mesh_vertnum;    /* nmesh */
  const double* mesh_vert;        /* nmeshvert*3 */
  const double* hfield_size;      /* nhfield*4 */
  const int32_t* hfield_nrow;     /* nhfield */
  const int32_t* hfield_ncol;     /* nhfield */
  const int32_t* hfield_adr;      /* nhfield */
  const double* hfield_data;      /* nhfielddata */
  /* sites */
  const int32_t* site_bodyid;     /* nsite */
  const double* site_pos;         /* nsite*3 */
  const double* site_quat;        /* nsite*4 */
  /* fixed tendons: length = sum over wraps of wrap_prm * qpos[joint wrap_objid] (mjWRAP_JOINT only) */
  const int32_t* tendon_adr;           /* ntendon */
  const int32_t* tendon_num;           /* ntendon */
  const int32_t* tendon_limited;       /* ntendon */
  const double* tendon_stiffness;      /* ntendon */
  const double* tendon_damping;        /* ntendon */
  const double* tendon_lengthspring;   /* ntendon*2 */
  const int32_t* wrap_type;            /* nwrap */
  const int32_t* wrap_objid;           /* nwrap */
  const double* wrap_prm;              /* nwrap */
  /* actuators (joint or fixed-tendon transmission) */
  const int32_t* actuator_trntype;     /* nu */
  const int32_t* actuator_trnid;       /* nu */
  const int32_t* actuator_gaintype;    /* nu */
  const int32_t* actuator_biastype;    /* nu */
  const int32_t* actuator_ctrllimited; /* nu */
  const int32_t* actuator_forcelimited;/* nu */
  const double* actuator_gear;         /* nu */
  const double* actuator_gainprm;      /* nu*3 */
  const double* actuator_biasprm;      /* nu*3 */
  const double* actuator_ctrlrange;    /* nu*2 */
  const double* actuator_forcerange;   /* nu*2 */
  /* sensors */
  const int32_t* sensor_type;     /* nsensor */
  const int32_t* sensor_objtype;  /* nsensor */
  const int32_t* sensor_objid;    /* nsensor */
  const int32_t* sensor_reftype;  /* nsensor */
  const int32_t* sensor_refid;    /* nsensor */
  const int32_t* sensor_intprm;   /* nsensor*3 */
  const int32_t* sensor_dim;      /* nsensor */
  const int32_t* sensor_adr;      /* nsensor */
  const double* sensor_cutoff;    /* nsensor */
  /* contact excludes (body pairs) */
  const int32_t* exclude_body1;   /* nexclude */
  const int32_t* exclude_body2;   /* nexclude */
} b2s_model_desc;

/* ---- initial state (host side; mirrors the MjData fields put_data reads) ---- */
typedef struct b2s_state_desc {
  const double* qpos;        /* nq */
  const double* qvel;        /* nv */
  const double* ctrl;        /* nu */
  const double* mocap_pos;   /* nmocap*3 */
  const double* mocap_quat;  /* nmocap*4 */
  double time;
} b2s_state_desc;

/* ---- zero-copy field descriptor ---- */
enum { B2S_F32 = 0, B2S_I32 = 1 };
typedef struct b2s_field_info {
  void* ptr;          /* device pointer */
  int32_t dtype;      /* B2S_F32 | B2S_I32 */
  int32_t ndim;
  int64_t shape[4];   /* logical shape; dim 0 is nworld for Data fields */
  int64_t stride[4];  /* in elements; stride[0] == 0 marks a broadcast (un-expanded) Model field */
} b2s_field_info;

typedef struct b2s_model b2s_model;
typedef struct b2s_data b2s_data;

/* Last error message of the calling thread ("" if none). */
const char* b2s_last_error(void);
/* Library/ABI version and the SM architecture the kernels were built for (100). */
int b2s_version(void);
int b2s_built_arch(void);
/* Number of kernels this library has launched since load (all streams). */
int64_t b2s_launch_count(void);

/* put_model: validates (unsupported features are an error, never ignored),
 * derives the traversal tables and uploads to `device`. */
int b2s_model_create(const b2s_model_desc* desc, int device, b2s_model** out);
int b2s_model_destroy(b2s_model* m);
/* Settable options after creation (sim.py:117-118, tests/test_sim.py:71-82).
 * Keys: "ls_parallel", "contact_sensor_maxmatch", "iterations", "ls_iterations",
 * "tolerance", "ls_tolerance", "timestep", "impratio", "integrator", "cone",
 * "gravity_x", "gravity_y", "gravity_z", "disableflags", "ls_parallel_min_step", "solver",
 * and two scheduling knobs without a counterpart in the reference: "stage_overlap" (0 / 1 / 2, default per model) and
 * "solve_rowcap" (rows the default solver launch holds in shared memory, default 128; worlds with
 * more rows are solved by a second launch sized for njmax; results do not depend on it).
 * Changing an option after a graph was captured needs a new capture, as in the reference. */
int b2s_model_set_opt(b2s_model* m, const char* key, double value);
int b2s_model_get_opt(const b2s_model* m, const char* key, double* value);
/* expand_model_fields: give `name` per-world storage (leading dim nworld).
 * Invalidates previously captured graphs (pointers change), as in the
 * reference (manager_based_rl_env.py:107-109 re-captures). */
int b2s_model_expand(b2s_model* m, const char* name, int nworld);
/* Model field view (per-world-capable fields have leading dim 1 or nworld). */
int b2s_model_field(b2s_model* m, const char* name, b2s_field_info* out);
int b2s_model_num_fields(const b2s_model* m);
const char* b2s_model_field_name(const b2s_model* m, int i);

/* put_data: allocate all per-world arrays and initialise every world from
 * `init`.  nconmax = contacts per world, njmax = constraint rows per world
 * (<= 0: heuristic). */
int b2s_data_create(const b2s_model* m, const b2s_state_desc* init, int nworld,
                    int nconmax, int njmax, b2s_data** out);
int b2s_data_destroy(b2s_data* d);
int b2s_data_field(b2s_data* d, const char* name, b2s_field_info* out);
int b2s_data_num_fields(const b2s_data* d);
const char* b2s_data_field_name(const b2s_data* d, int i);
int64_t b2s_data_bytes(const b2s_data* d);

/* Dense constraint Jacobian (mjwarp's Data.efc.J, shape nworld x njmax x roundup4(nv)).  The step keeps
 * constraint rows in a matrix-free geometric form (efc_g, efc_bp); this materialises the dense matrix on
 * demand into storage owned by `d` (allocated on first use; call again after a step to refresh it). */
int b2s_efc_jacobian(const b2s_model* m, b2s_data* d, void* stream, b2s_field_info* out);

/* Frames of geoms and sites on static (world-welded) bodies are computed once at b2s_data_create; call this after
 * writing pose fields of a static body (per-world terrain randomisation).  The reference recomputes them every step. */
int b2s_refresh_static(const b2s_model* m, b2s_data* d, void* stream);

/* forward / step on `stream` (a cudaStream_t passed as void*). */
int b2s_forward(const b2s_model* m, b2s_data* d, void* stream);
int b2s_step(const b2s_model* m, b2s_data* d, void* stream);
/* nsub sub-steps with ctrl held (one control decimation; manager_based_rl_env.py:114-119). */
int b2s_step_n(const b2s_model* m, b2s_data* d, int nsub, void* stream);

/* Allocation-free re-solve.  NO COUNTERPART IN THE REFERENCE: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Runs the constraint
 * solve (the k_solve stage of b2s_forward: SURVEY.md row K13) on the constraint rows, smooth forces and warm start that the
 * last b2s_forward / b2s_step left in `d`, and rewrites d's qacc, qfrc_constraint, efc_force, solver_niter and solver_cost.
 * The solver's scheduling scratch (cost-ordered world queues, normally part of d's arena and filled by the constraint stage)
 * lives in a file-scope device array of fixed size instead, so the call allocates nothing: no cudaMalloc, cudaFree or
 * allocating library routine.  Limit: d's nworld <= B2S_NOALLOC_MAX_WORLDS (larger -> error, nothing launched); every GPU
 * that loads the library holds the array (4 MiB + 64 B) until exit.  The array is shared by all calls on a device: the call
 * clears the part that must start at zero, and calls whose GPU work overlaps (different streams) race on it -- serialise
 * them on one stream.  Results are identical to b2s_forward's solve (the queue only orders the work).  The first call on a
 * device looks the array's device address up (cudaGetSymbolAddress): make that call outside stream capture. */
#define B2S_NOALLOC_MAX_WORLDS 65536
int b2s_solve_noalloc(const b2s_model* m, b2s_data* d, void* stream);
int b2s_noalloc_max_worlds(void);

/* Host-buffer convenience used by the end-to-end benchmark and by FFI callers
 * without a device allocator: copy ctrl (nworld*nu floats) from host, run nsub
 * sub-steps, copy qpos|qvel|sensordata back.  Pinned buffers recommended. */
int b2s_step_host(const b2s_model* m, b2s_data* d, const float* ctrl_host, int nsub,
                  float* qpos_host, float* qvel_host, float* sensordata_host, void* stream);

/* Per-stage device timing of one step (cudaEvents on `stream`, synchronises):
 * fills up to `cap` (name, ms) pairs; returns the number of stages. */
int b2s_profile_step(const b2s_model* m, b2s_data* d, void* stream, int reps,
                     const char** names, float* ms, int cap);

/* ---------------------------------------------------------------------------------------------------------
 * Fused manager side of the velocity task's env step.  The reference runs ~60 manager terms as
 * ~500 small torch kernels per env step (envs/manager_based_rl_env.py:111-152, tasks/velocity/mdp/rewards.py:23-359,
 * velocity_env_cfg.py:138-240, sensor/contact_sensor.py:327-367); these four launches read the engine's state arrays
 * directly, one warp per env.  All pointers are device pointers owned by the caller (torch tensors); `U` is a
 * (nworld, nrand) buffer of uniform [0,1) numbers refreshed by the caller before every env step, consumed at fixed
 * slots (B2S_VEL_U_*), so that a host-side restatement can be checked against the kernels on identical draws.
 * Index fields address the engine's sensordata / body / site arrays.  Weights follow B2S_VEL_TERM_* order. */
enum {
  B2S_VEL_TERM_TRACK_LIN = 0, B2S_VEL_TERM_TRACK_ANG, B2S_VEL_TERM_UPRIGHT, B2S_VEL_TERM_POSE, B2S_VEL_TERM_BODY_ANG_VEL,
  B2S_VEL_TERM_ANGMOM, B2S_VEL_TERM_DOF_LIMITS, B2S_VEL_TERM_ACTION_RATE, B2S_VEL_TERM_SELF_COLLISIONS, B2S_VEL_TERM_AIR_TIME,
  B2S_VEL_TERM_FOOT_CLEARANCE, B2S_VEL_TERM_FOOT_SWING, B2S_VEL_TERM_FOOT_SLIP, B2S_VEL_TERM_SOFT_LANDING, B2S_VEL_NTERM
};
/* slots of U: observation noise (lin 3, gyro 3, gravity 3, joint pos nu, joint vel nu) starts at 0; the rest follows at
 * base = 9 + 2 nu: reset x, y, yaw, push timer; command resample at reset (vx, vy, wz, heading, is_heading, is_standing,
 * timer); the same 7 for the timer-driven resample; push kick x, y, next push timer.  nrand >= base + 21. */
typedef struct b2s_velocity_mdp {
  int nworld, nq, nv, nu, nbody, nsite, nsensordata, nrand;
  /* engine state (b2s_data fields) */
  float *qpos, *qvel, *qacc_warmstart, *ctrl;
  const float *time, *sensordata, *xquat, *cvel, *site_xpos, *subtree_com;
  /* task constants */
  const float *default_qpos, *action_scale, *soft_lo, *soft_hi, *std_standing, *std_walking, *std_running, *env_origins;
  int torso, feet_site[2], feet_body[2], feet_root[2];
  int s_gyro, s_linvel, s_angmom, s_found[2], s_force[2], s_self;
  int max_episode_length;
  float step_dt, landed_dt, fall_angle;      /* landed_dt = step_dt + 1e-8 rounded to float (first-contact window) */
  float weight[B2S_VEL_NTERM];              /* term weight x step_dt (0 = term not summed) */
  float cmd_lo[4], cmd_hi[4], cmd_span[4];  /* lin_vel_x, lin_vel_y, ang_vel_z, heading; span = hi - lo rounded once */
  /* env buffers */
  const float* actions_in;
  float *action, *prev_action, *command, *heading_target, *cmd_time_left, *push_time_left;
  unsigned char *is_heading, *is_standing, *dones, *time_outs;
  long long* episode_length_buf;
  float *air_cur, *air_last, *con_cur, *con_last, *air_last_time, *peak_heights;
  float *obs, *critic_obs, *rew, *term_sums;   /* term_sums: (B2S_VEL_NTERM, nworld) */
  const float* U;
} b2s_velocity_mdp;
/* process + apply the action: prev_action <- action <- actions_in, ctrl = default joint pos + scale * action */
int b2s_vel_apply_action(const b2s_velocity_mdp* p, void* stream);
/* contact air-time tracker after each physics sub-step (ContactSensor._update_air_time_tracking) */
int b2s_vel_air_time(const b2s_velocity_mdp* p, void* stream);
/* episode counter, terminations, the 14 reward terms, then the masked reset writes of finished envs */
int b2s_vel_post(const b2s_velocity_mdp* p, void* stream);
/* after forward(): command timers / heading control, push event, policy and critic observations */
int b2s_vel_observe(const b2s_velocity_mdp* p, void* stream);

/* ---------------------------------------------------------------------------------------------------------
 * Fused manager terms of the motion-tracking task (tasks/tracking/mdp/{rewards,terminations,observations}.py, the metrics and
 * the re-anchoring of commands.py:219-259,396-415).  Two launches per env step, one warp per env, one lane per tracked body /
 * per joint; phase sampling, resets and the push event stay host-side tensor ops.  Motion arrays are (T, nb, 3|4) for the nb
 * tracked bodies and (T, nu) for the joints; `time_steps` is each env's frame.  U: (nworld, nrand >= 15 + 2 nu) uniform [0,1)
 * numbers for the observation noise (anchor pos 3, anchor ori 6, base lin vel 3, base ang vel 3, joint pos nu, joint vel nu). */
enum { B2S_TRK_TERM_ROOT_POS = 0, B2S_TRK_TERM_ROOT_ORI, B2S_TRK_TERM_BODY_POS, B2S_TRK_TERM_BODY_ORI, B2S_TRK_TERM_BODY_LIN_VEL,
       B2S_TRK_TERM_BODY_ANG_VEL, B2S_TRK_TERM_ACTION_RATE, B2S_TRK_TERM_JOINT_LIMIT, B2S_TRK_TERM_SELF_COLLISIONS, B2S_TRK_NTERM };
enum { B2S_TRK_MET_ANCHOR_POS = 0, B2S_TRK_MET_ANCHOR_ROT, B2S_TRK_MET_BODY_POS, B2S_TRK_MET_BODY_ROT, B2S_TRK_MET_JOINT_POS,
       B2S_TRK_MET_JOINT_VEL, B2S_TRK_NMET };
typedef struct b2s_tracking_mdp {
  int nworld, nq, nv, nu, nbody, nsensordata, nrand, nb, anchor_k, ntime;
  const float *qpos, *qvel, *sensordata, *xpos, *xquat, *cvel, *subtree_com;
  const float *mo_jp, *mo_jv, *mo_bpos, *mo_bquat, *mo_blin, *mo_bang, *env_origins;
  const long long* time_steps;
  const int *body_ids, *body_root;          /* model body id / root body id of each tracked body */
  const unsigned char* ee_mask;             /* [nb] 1 for end-effector bodies (termination) */
  const float *default_joint_pos, *soft_lo, *soft_hi, *action, *prev_action;
  int s_gyro, s_linvel, s_self, max_episode_length;
  float weight[B2S_TRK_NTERM];              /* term weight x step_dt */
  float inv_std2[6];                        /* 1 / std^2 of the six exponential tracking terms */
  float *body_pos_relative_w, *body_quat_relative_w;
  long long* episode_length_buf;
  unsigned char *dones, *time_outs, *terminated;
  float *rew, *term_sums, *metrics;         /* term_sums (B2S_TRK_NTERM, nworld), metrics (B2S_TRK_NMET, nworld) */
  float *obs, *critic_obs;
  const float* U;
} b2s_tracking_mdp;
/* episode counter, terminations (time-out, anchor height, anchor orientation, end-effector height), the 9 reward terms, metrics */
int b2s_trk_terms(const b2s_tracking_mdp* p, void* stream);
/* re-anchor the motion's body poses at the robot's anchor (body_*_relative_w), then policy and critic observations */
int b2s_trk_observe(const b2s_tracking_mdp* p, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200SIM_H_ */
