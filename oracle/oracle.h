/* oracle.h -- per-world working set of the CPU oracle.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product
 * path; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may build, load or call it.
 *
 * PARITY UNPINNED: the algorithm this restates lives in third-party packages
 * that are absent from /root/reference and from this container
 * (mujoco-warp @ e605c406, mujoco 3.4.0.dev818719010), and the
 * reference's tests hold no golden vectors for the step arithmetic.  The
 * oracle is pinned instead by analytic known-answer tests
 * (tests/test_oracle_kat.py) and by an independent float64 numpy kinematics /
 * CRBA (asimov-mjlab_b200/hostmodel.py).
 *
 * One `orc_data` describes ONE world; all arrays are caller-allocated.  The
 * Python wrapper parses this header (one member per line, extent in the
 * trailing comment) -- keep that shape.  `real` is float (liboracle_f32) or
 * double (liboracle_f64).
 */
#ifndef ORACLE_H_
#define ORACLE_H_
#include <stdint.h>
#include "b200sim.h"

#ifdef ORC_DOUBLE
typedef double real;
#else
typedef float real;
#endif

typedef struct orc_data {
  int32_t nconmax;
  int32_t njmax;
  /* state / inputs */
  real* qpos;            /* nq */
  real* qvel;            /* nv */
  real* ctrl;            /* nu */
  real* qacc_warmstart;  /* nv */
  real* time;            /* 1 */
  real* xfrc_applied;    /* nbody*6 */
  real* qfrc_applied;    /* nv */
  real* mocap_pos;       /* nmocap*3 */
  real* mocap_quat;      /* nmocap*4 */
  /* position-dependent */
  real* xpos;            /* nbody*3 */
  real* xquat;           /* nbody*4 */
  real* xmat;            /* nbody*9 */
  real* xipos;           /* nbody*3 */
  real* ximat;           /* nbody*9 */
  real* xanchor;         /* njnt*3 */
  real* xaxis;           /* njnt*3 */
  real* geom_xpos;       /* ngeom*3 */
  real* geom_xmat;       /* ngeom*9 */
  real* site_xpos;       /* nsite*3 */
  real* site_xmat;       /* nsite*9 */
  real* subtree_com;     /* nbody*3 */
  real* cinert;          /* nbody*10 */
  real* cdof;            /* nv*6 */
  real* qM;              /* nv*nv */
  real* qL;              /* nv*nv */
  /* velocity-dependent */
  real* cvel;            /* nbody*6 */
  real* cdof_dot;        /* nv*6 */
  real* cacc;            /* nbody*6 */
  real* subtree_linvel;  /* nbody*3 */
  real* subtree_angmom;  /* nbody*3 */
  real* qfrc_bias;       /* nv */
  real* qfrc_passive;    /* nv */
  real* ten_length;      /* ntendon */
  real* ten_velocity;    /* ntendon */
  real* actuator_length; /* nu */
  real* actuator_velocity; /* nu */
  real* actuator_force;  /* nu */
  real* qfrc_actuator;   /* nv */
  real* qfrc_smooth;     /* nv */
  real* qacc_smooth;     /* nv */
  real* qacc;            /* nv */
  real* qfrc_constraint; /* nv */
  real* sensordata;      /* nsensordata */
  /* contacts */
  int32_t* ncon;         /* 1 */
  real* con_dist;        /* nconmax */
  real* con_pos;         /* nconmax*3 */
  real* con_frame;       /* nconmax*9 */
  real* con_includemargin; /* nconmax */
  real* con_friction;    /* nconmax*5 */
  real* con_solref;      /* nconmax*2 */
  real* con_solimp;      /* nconmax*5 */
  int32_t* con_dim;      /* nconmax */
  int32_t* con_geom;     /* nconmax*2 */
  int32_t* con_efc_address; /* nconmax */
  /* constraints */
  int32_t* nefc;         /* 1 */
  /* rows [0, nl) act on a single dof: first the nf friction-loss rows, then the joint-limit rows */
  int32_t* nl;           /* 1 */
  int32_t* nf;           /* 1 */
  int32_t* efc_type;     /* njmax */
  int32_t* efc_id;       /* njmax */
  real* efc_J;           /* njmax*nv */
  real* efc_pos;         /* njmax */
  real* efc_margin;      /* njmax */
  real* efc_D;           /* njmax */
  real* efc_vel;         /* njmax */
  real* efc_aref;        /* njmax */
  real* efc_force;       /* njmax */
  real* efc_frictionloss; /* njmax */
  /* solver statistics */
  int32_t* solver_niter; /* 1 */
  real* solver_cost;     /* 1 */
} orc_data;

enum { ORC_EFC_FRICTION_DOF = 1, ORC_EFC_LIMIT = 3, ORC_EFC_CONTACT_FRICTIONLESS = 4, ORC_EFC_CONTACT_PYRAMIDAL = 5, ORC_EFC_CONTACT_ELLIPTIC = 6 };

int orc_real_size(void);
/* mjwarp.forward / mjwarp.step for one world (reference call sites: sim.py:143,146). */
int orc_forward(const b2s_model_desc* m, orc_data* d);
int orc_step(const b2s_model_desc* m, orc_data* d);
/* individual stages, exposed for stage-level parity tests */
void orc_kinematics(const b2s_model_desc* m, orc_data* d);
void orc_com_pos(const b2s_model_desc* m, orc_data* d);
void orc_crb(const b2s_model_desc* m, orc_data* d);
int  orc_factor_m(const b2s_model_desc* m, orc_data* d);
void orc_collision(const b2s_model_desc* m, orc_data* d);
void orc_make_constraint(const b2s_model_desc* m, orc_data* d);
void orc_fwd_velocity(const b2s_model_desc* m, orc_data* d);
void orc_fwd_actuation(const b2s_model_desc* m, orc_data* d);
void orc_fwd_acceleration(const b2s_model_desc* m, orc_data* d);
void orc_solve(const b2s_model_desc* m, orc_data* d);
void orc_sensors(const b2s_model_desc* m, orc_data* d);
void orc_integrate(const b2s_model_desc* m, orc_data* d);
/* rollout: nstep steps for nworld worlds (array of orc_data), ctrl held; OpenMP over worlds if enabled. */
int orc_rollout(const b2s_model_desc* m, orc_data* worlds, int nworld, int nstep, int nthread);
#endif
