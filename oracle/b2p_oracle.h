/* TEST INFRASTRUCTURE -- NOT A PRODUCT PATH.
 *
 * CPU oracle: a plain-C, fp64, single-scene restatement of the hot path that
 * BASELINE.json:north_star names -- MuJoCo's mj_step (third-party `mujoco==3.2.7`,
 * un-vendored; reference pyproject.toml:42) as called from
 * mujoco_robot_environments/models/robot_arm.py:61-94, plus the operational-space
 * controller mujoco_controllers/osc.py:99-247 -- for the rearrangement model class.
 *
 * PARITY UNPINNED: the reference holds no golden vectors for this path and MuJoCo is
 * not installable here.  The oracle follows MuJoCo's published
 * algorithms and is pinned only by analytic known-answer tests
 * (tests/test_oracle_*.py).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library.  The package never does.
 */
#ifndef B2P_ORACLE_H
#define B2P_ORACLE_H
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_t orc_t;

enum { ORC_MAXCON = 192, ORC_MAXEFC = 640 };

/* field ids for orc_ptr() */
enum {
  ORC_QPOS, ORC_QVEL, ORC_CTRL, ORC_QACC_WARMSTART, ORC_TIME, ORC_QACC, ORC_QACC_SMOOTH,
  ORC_QFRC_BIAS, ORC_QFRC_PASSIVE, ORC_QFRC_ACTUATOR, ORC_QFRC_SMOOTH, ORC_QFRC_CONSTRAINT,
  ORC_XPOS, ORC_XQUAT, ORC_XMAT, ORC_XIPOS, ORC_GEOM_XPOS, ORC_GEOM_XMAT, ORC_SITE_XPOS, ORC_SITE_XMAT,
  ORC_QM, ORC_SUBTREE_COM, ORC_CINERT, ORC_CDOF, ORC_CVEL, ORC_CDOF_DOT,
  ORC_EFC_J, ORC_EFC_POS, ORC_EFC_D, ORC_EFC_R, ORC_EFC_AREF, ORC_EFC_FORCE, ORC_EFC_VEL, ORC_EFC_MARGIN,
  ORC_PROP_SIZE, ORC_ACT_FORCE, ORC_OSC_TAU, ORC_SOLVER_STATS
};

orc_t* orc_create(const void* blob, size_t nbytes);
void orc_destroy(orc_t* o);
double* orc_ptr(orc_t* o, int field);       /* NULL if unknown */
int orc_dim(orc_t* o, int what);            /* 0 nq 1 nv 2 nu 3 nbody 4 ngeom 5 nsite 6 ncon 7 nefc 8 nprop 9 solver iters */
void orc_reset(orc_t* o);                   /* mj_resetData: qpos=qpos0, rest 0 */
void orc_set_prop(orc_t* o, int prop, const double half[3], int active);
void orc_set_prop_type(orc_t* o, int prop, int geom_type); /* B2P_GEOM_BOX (as compiled) / B2P_GEOM_SPHERE / B2P_GEOM_CYLINDER: include/b2p_prop_types.h */
void orc_set_option(orc_t* o, int what, double v); /* 0 full contact list (1) / solver-active only (0); 1 integrator; 2 tolerance; 3 iterations; 4 solver (0 Newton, 1 PGS) */

void orc_step1(orc_t* o);                   /* mj_step1: position + velocity stages */
void orc_step2(orc_t* o);                   /* mj_step2: actuation, acceleration, constraint, integrate */
void orc_step(orc_t* o);                    /* mj_step = step1 ; step2 */
void orc_forward(orc_t* o);                 /* mj_forward (no integration) */
void orc_step_frozen_arm(orc_t* o);         /* one settle step under JointStaticIsolator(prop joints) */

/* contacts: returns ncon; arrays may be NULL. dist[n], pos[3n], frame[9n], geom[2n], excluded[n] */
int orc_contacts(orc_t* o, double* dist, double* pos, double* frame, int* geom, int* excluded);

/* OSC (mujoco_controllers/osc.py): target13 = pos3, vel3, quat4 (wxyz), angvel3; gains6 = kp,kd pos / ori / null */
void orc_osc_config(orc_t* o, const double gains6[6], const double nullspace7[7], double pos_thresh, double ori_thresh);
void orc_osc_torque(orc_t* o, const double target13[13], double tau7[7]);   /* needs step1-consistent data */
int orc_osc_converged(orc_t* o, const double target13[13]);
/* RobotArm.run_controller: n_ticks x (OSC + gripper + control_steps physics steps). Returns latched arm_converged */
int orc_run_controller(orc_t* o, const double target13[13], int grip_max, int n_ticks, int control_steps);

#ifdef __cplusplus
}
#endif
#endif
