/* b2p -- C-ABI of the B200-native batched physics + OSC backend (libb2p.so).
 *
 * Drop-in boundary for the hot path of peterdavidfagan/mujoco_robot_environments:
 * the calls below replace, with a leading env dimension, what the reference does through
 * dm_control / mujoco for ONE scene:
 *
 *   b2p_step_physics   <- physics.set_control(ctrl); physics.step()           models/robot_arm.py:78-79
 *   b2p_step_osc       <- RobotArm.run_controller(duration)                   models/robot_arm.py:61-94
 *                         (OSC.compute_control_output  mujoco_controllers/osc.py:180-238,
 *                          MinMax.compute_control_output  mujoco_controllers/min_max.py:14-19,
 *                          OSC.is_converged  osc.py:240-247)
 *                         LasaDrawEnv.move_to_draw_target / interactive_tuning (one tick)   tasks/lasa_draw.py:300-349
 *                         (b2p_step_physics on the position-servo scene: move_to_joint_position_target  tasks/lasa_draw.py:351-368)
 *   b2p_settle         <- PropPlacer settle loop under JointStaticIsolator    environment/prop_initializer.py:245-257
 *   b2p_forward        <- physics.forward() + reads of data.site_xpos / contact / qacc
 *                                                                           prop_initializer.py:190,221 ; tasks/rearrangement.py:653
 *   b2p_osc_config     <- config/robots/arm/controller_config/osc.yaml gains / nullspace / convergence
 *   b2p_set_env_params <- per-scene prop sizes drawn in environment/props.py:277-291 and the per-prop
 *                         contype/conaffinity toggling of prop_initializer.py:142-162
 *
 * Model class (b2p_create refuses anything else with B2P_ERR_MODEL and a message): one hinge-jointed tree of <= 15 dofs (the arm, with
 * or without the gripper's four-bar) under unit-gear joint actuators, <= 4 free box props (sphere / cylinder colliders per scene through
 * b2p_set_env_prop_types), static boxes / planes / convex hulls, a cylinder on the tree (the drawing scene's tool piece), condim 3,
 * Newton solver, elliptic or pyramidal friction cone, implicitfast or Euler integrator.
 *
 * Conventions (after the reference's own raw C boundary, madrona_mjx src/bindings.cpp:219-250): the caller owns
 * every I/O buffer (device pointers, e.g. torch tensors), the handle owns scratch, work is enqueued on the
 * caller's stream, calls return 0 or a negative error code (no exceptions cross the boundary), a handle is
 * not thread-safe, one handle per (GPU, model topology).  dtype: 0 = float64, 1 = float32; every "real"
 * pointer below is of the handle's dtype.  Quaternions are MuJoCo order (w, x, y, z).
 */
#ifndef B2P_H
#define B2P_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct b2p_handle_s* b2p_handle;

enum { B2P_F64 = 0, B2P_F32 = 1 };
enum { B2P_OK = 0, B2P_ERR_ARG = -1, B2P_ERR_MODEL = -2, B2P_ERR_CUDA = -3, B2P_ERR_STATE = -4 };

/* b2p_query ids */
enum {
  B2P_Q_NQ = 0, B2P_Q_NV, B2P_Q_NU, B2P_Q_NPROP, B2P_Q_STATE_STRIDE, B2P_Q_OUT_STRIDE, B2P_Q_MAXPROP,
  B2P_Q_REC_QPOS, B2P_Q_REC_QVEL, B2P_Q_REC_QACC_WARMSTART, B2P_Q_REC_TIME, B2P_Q_LAUNCHES, B2P_Q_SMEM_BYTES,
  B2P_Q_BLOCKS_PER_SM, B2P_Q_THREADS, B2P_Q_GRID, B2P_Q_MAXCON, B2P_Q_NGEOM,
  B2P_Q_ORDER_LAUNCHES   /* launches of the scene-queue ordering kernel (one before each step launch once a batch exceeds one round) */
};
/* derived-output fields (b2p_get / offsets inside the output record via b2p_out_offset) */
enum {
  B2P_O_SITE_XPOS = 0 /* [2][3] attachment_site, pinch */, B2P_O_SITE_XMAT /* [2][9] */, B2P_O_QACC /* [40] */,
  B2P_O_QFRC_BIAS /* [40] */, B2P_O_CTRL /* [8] */, B2P_O_NCON, B2P_O_NEFC, B2P_O_NITER,
  B2P_O_CONTACT /* [48][10]: dist,pos3,normal3,geom1,geom2,excluded */, B2P_O_MARM /* [16][16] */,
  B2P_O_GEOM_XPOS /* [32][3] */, B2P_O_GEOM_XMAT /* [32][9] */, B2P_O_QACC_SMOOTH /* [40] */, B2P_O_EFC_FORCE /* [160] */,
  B2P_O_NFIELD
};

int b2p_create(const void* model_blob, size_t nbytes, int n_env, int dtype, int device, b2p_handle* out);
int b2p_destroy(b2p_handle h);
int b2p_query(b2p_handle h, int what);
int b2p_out_offset(int field, int* offset, int* count);

/* state: [n_env][B2P_Q_STATE_STRIDE] reals, one record per env: qpos | qvel | qacc_warmstart | time (see B2P_Q_REC_*) */
int b2p_bind_state(b2p_handle h, void* state);
/* optional caller-owned derived-output buffer [n_env][B2P_Q_OUT_STRIDE]; NULL = use the handle's own */
int b2p_bind_out(b2p_handle h, void* out);
/* Which groups of the derived-output record the STEPPING calls (b2p_step_physics / b2p_step_osc / b2p_settle) write at the end of a
 * launch; b2p_forward always writes every group (that is what physics.forward() is for).  Default: all, so that any read of
 * data.* after a step sees current values, as with mjData.  A caller that only reads the end-effector site after run_controller
 * (RobotArm.eef_pos, models/robot_arm.py:55-58; OSC.current_position_error, osc.py:156-157) sets B2P_OUT_SITE and saves the
 * 5.7 kB (fp32) per scene per launch the other groups cost. */
enum {
  B2P_OUT_SITE = 1 /* SITE_XPOS, SITE_XMAT, NCON, NEFC */, B2P_OUT_CONTACT = 2, B2P_OUT_MARM = 4, B2P_OUT_GEOM = 8 /* GEOM_XPOS, GEOM_XMAT */,
  B2P_OUT_BIAS = 16 /* QFRC_BIAS */, B2P_OUT_DYN = 32 /* QACC, QACC_SMOOTH, CTRL, NITER */, B2P_OUT_EFC = 64 /* EFC_FORCE */, B2P_OUT_ALL = 127
};
int b2p_set_out_mask(b2p_handle h, int mask);
/* Constraint solver of the stepping calls (mjOption.solver; the reference never sets it, tasks/rearrangement.py:77-78, so MuJoCo's
 * default -- Newton -- is this library's default too).  PGS is MuJoCo's projected Gauss-Seidel on the dual problem (elliptic cones),
 * offered as the independent minimiser of the same objective and because BASELINE.json names it; it runs as its own kernel
 * instantiation and is several times slower than Newton on this model class.  Refused (B2P_ERR_ARG) for a model with a pyramidal cone or
 * a sphere / cylinder geom on the tree (the drawing scene of tasks/lasa_draw.py): those run in the class-capable Newton kernels only. */
enum { B2P_SOLVER_NEWTON = 0, B2P_SOLVER_PGS = 1 };
int b2p_set_solver(b2p_handle h, int solver);
/* prop_size: [n_env][MAXPROP][3] half sizes (NULL = model default); prop_active: [n_env][MAXPROP] (NULL = all) */
int b2p_set_env_params(b2p_handle h, const void* prop_size, const uint8_t* prop_active);
/* optional collider class per prop and scene: device [n_env][MAXPROP] bytes, B2P_GEOM_BOX (6, the compiled box with the scene's half
 * sizes), B2P_GEOM_SPHERE (2; prop_size[.][0] = radius) or B2P_GEOM_CYLINDER (5; prop_size[.][0] = radius, [1] = half height, axis z);
 * NULL = every prop is the compiled box.  Replaces what add_objects (environment/props.py:520-617) does when it draws a shape per
 * prop: the classes' contact parameters (cylinder: condim 6, friction 1 1 1, solref 0.01 1.1; GalaApple: a default sphere) are in
 * include/b2p_prop_types.h.  The pointer is borrowed. */
int b2p_set_env_prop_types(b2p_handle h, const uint8_t* prop_type);
/* optional per-scene OSC gains [n_env][6] reals (position kp, kd, orientation kp, kd, nullspace kp, kd): one candidate of the
 * controller tuner per scene (automated_controller_tuning/rearrangement_controller_tuning.py:145-168 mutates
 * `controller_gains` between roll-outs); NULL = every scene uses b2p_osc_config's gains.  The pointer is borrowed. */
int b2p_set_env_gains(b2p_handle h, const void* gains);
int b2p_osc_config(b2p_handle h, const double gains6[6], const double nullspace7[7], double pos_thresh, double ori_thresh);
/* 0 = only solver-active contacts while stepping (default), 1 = list every contact inside geom margin */
int b2p_set_full_contacts(b2p_handle h, int full);
/* sticky scene mask for b2p_step_physics / b2p_step_osc (device [n_env], borrowed; NULL = all): scenes with mask 0 are not
 * stepped and none of their outputs is written -- the batched form of "this scene's episode is over"
 * (transporter_network_data_generation.py:108-112 stops stepping an env once its demonstration is complete or has failed) */
int b2p_set_env_mask(b2p_handle h, const uint8_t* env_mask);

int b2p_step_physics(b2p_handle h, const void* ctrl /* [n_env][nu] */, int n_steps, void* stream);
/* b2p_step_physics for callers that must not allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same kernel, same results as b2p_step_physics on the same state.  The kernel's global-memory scratch (one Jacobian overflow area
 * per resident warp + the scene-queue counter) is a file-scope `static __device__` array of csrc/b2p_api.cu instead of the handle's
 * cudaMalloc'd pools: about 3.3 MB that every GPU which loads libb2p.so holds until the process exits.  Its device address is
 * resolved once per device by b2p_create (cudaGetSymbolAddress, that device current); this call itself makes no cudaMalloc /
 * cudaFree / allocating library call -- a cudaMemsetAsync of the queue counter and the kernel launch on `stream`, nothing else.
 *   state : device [n_env][B2P_Q_STATE_STRIDE] reals, 16-byte aligned, advanced in place (passed here, not taken from b2p_bind_state)
 *   ctrl  : device [n_env][nu] reals, or NULL for zero controls
 * Limits: the handle holds at most B2P_NOALLOC_MAX_ENV scenes (B2P_ERR_ARG otherwise, b2p_last_error says so), Newton solver,
 * box props only (no b2p_set_env_prop_types).  Status words, derived outputs (b2p_get), env mask and env params are the handle's.
 * The scene queue is not re-ordered by work and the handle's work figures are not touched (pure scheduling: results do not depend on it).
 * EVERY call on a device shares the one array, which keeps the previous call's contents (the counter is cleared by each call; the
 * overflow areas need no clearing, a row is written before it is read): calls that overlap on the GPU -- two streams, two handles,
 * two host threads -- RACE.  Serialise them on one stream or with events. */
enum { B2P_NOALLOC_MAX_ENV = 128 };
int b2p_step_physics_noalloc(b2p_handle h, void* state, const void* ctrl /* [n_env][nu] or NULL */, int n_steps, void* stream);
int b2p_step_osc(b2p_handle h, const void* target13 /* [n_env][13] pos vel quat angvel */, const uint8_t* grip /* [n_env] */,
                 int n_ticks, int control_steps, uint8_t* converged /* [n_env] or NULL */, void* stream);
int b2p_settle(b2p_handle h, const uint8_t* env_mask /* NULL = all */, int max_steps, int min_steps, double qvel_tol,
               double qacc_tol, int32_t* steps_done /* [n_env] or NULL */, void* stream);
int b2p_forward(b2p_handle h, int full_contacts, void* stream);
/* SortColours.prop_place (tasks/rearrangement.py:597-665) for every scene in ONE launch: up to max_attempts proposals per scene,
 * position ~ U(box lo, box hi) (box: device [n_env][6] reals lo3 | hi3), gripper-down orientation, accepted when the prop
 * (prop: device [n_env] prop index) has no contact closer than `dist` with any geom but `table_geom` after kinematics + the full
 * collision pass.  Proposals are a counter-based function of (seed, scene, attempt).  out: device [n_env][4] reals = accepted
 * xyz, attempts used (negative: nothing accepted).  The bound state is left untouched (the reference searches on a deepcopy
 * of the physics); scenes with env_mask 0 are skipped and their out row is not written. */
int b2p_place_search(b2p_handle h, const int32_t* prop, const void* box, const uint8_t* env_mask /* NULL = all */, int max_attempts,
                     double dist, int table_geom, unsigned seed, void* out, void* stream);
/* PropPlacer.__call__ for ONE prop of every scene in one launch (environment/prop_initializer.py:197-236, driven by the arena's
 * workspace box and `rotation = pi * U about z`, tasks/rearrangement.py:189-197): like b2p_place_search, but every proposal also
 * draws the yaw, and the accepted orientation comes back in quat_out (device [n_env][4], w x y z).  The collision rule is the
 * placer's "_has_collisions_with_prop": any listed contact of the prop closer than `dist` with anything but `table_geom`, with the
 * props not placed yet switched off through prop_active (prop_initializer.py:142-162,183-185).  The caller writes the accepted pose
 * into the state (the search itself runs on the warp's scratch copy) and moves on to the next prop: a reset places N props in N launches. */
int b2p_place_props(b2p_handle h, const int32_t* prop, const void* box, const uint8_t* env_mask /* NULL = all */, int max_attempts,
                    double dist, int table_geom, unsigned seed, void* out, void* quat_out, void* stream);
int b2p_get(b2p_handle h, int field, void* dst /* [n_env][count] */, void* stream);
int b2p_status(b2p_handle h, int32_t* flags /* device [n_env] */, int clear, void* stream);
/* overwrite the per-scene status words (device [n_env]): a masked reset clears the flags of the scenes it resets and of no other */
int b2p_status_set(b2p_handle h, const int32_t* flags, void* stream);

/* developer aid: per-stage cycle counters summed over scenes, [B2P_PROF_N] entries; all zero unless the library was built
 * with -DB2P_PROF (the product build has no counters in the kernel).  Synchronises the device. */
enum { B2P_PROF_N = 24 };
int b2p_prof_read(b2p_handle h, unsigned long long* out, int reset);

/* micro-benchmark: register-resident FMA chain, returns TFLOP/s of the handle-independent device (dtype 0/1) */
int b2p_fma_peak(int dtype, int device, double* tflops);
const char* b2p_last_error(void);

#ifdef __cplusplus
}
#endif
#endif
