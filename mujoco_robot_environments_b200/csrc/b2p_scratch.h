#pragma once
#include "b2p_math.h"

namespace b2p {

// ------------------------------------------------------------------------------------------ per-env scratch
// HBM state record of one env = the head of Scratch<T>: qpos[MAXNQ] | qvel[MAXNV] | qacc_warmstart[MAXNV] | time | pad
constexpr int REC_QPOS = 0, REC_QVEL = MAXNQ, REC_QACC = MAXNQ + MAXNV, REC_TIME = MAXNQ + 2 * MAXNV, REC_SIZE = 128;
// one hull-graph neighbour with its coordinates inline (one 16-byte load in fp32), and (first neighbour, count) per vertex
constexpr int W6_STRIDE = 24;        // reals per condim-6 contact in the w6 area: 21 (packed lower triangle) + torsional, rolling friction + pad
constexpr int SEP_N = 8;             // separation witnesses kept per scene
template <typename T> struct alignas(16) NbrPt { T x, y, z; int idx; };
struct NbrSpan { int adr, num; };
template <typename T>
struct Globals {                 // read-only global-memory tables
  const T* mesh_vert; const int* nbradr; const int* nbrnum; const int* nbr;
  const short* pair_g1; const short* pair_g2; const PairParam<T>* pair_param;
  const int* geom_start;         // [ngeom][27] hill-climb start vertex per quantised direction (mesh geoms)
  const NbrPt<T>* nbr_pt;        // [nmeshnbr] neighbour vertex coordinates + index, same order as nbr
  const NbrSpan* nbr_span;       // [nmeshvert]
  T* jovf;                       // [warp slots][Pool<T>::JP] Jacobian rows that do not fit a scene's shared-memory pool
  T* w6pool;                     // [warp slots][MAXCON][W6_STRIDE] condim-6 contacts: cone Hessian blocks + friction coefficients
  T* xpool;                      // [warp slots][Pool<T>::JP] PGS only: X = M^-1 J' in J's packed layout (NULL until b2p_set_solver(PGS))
  const GeomRaw<T>* geom_raw;    // [ngeom] unmixed contact parameters (pairs with a re-classed prop are mixed on the device)
  const T* geom_dop;             // [ngeom][27] hull support value along the 26 directions {-1,0,1}^3 (mesh geoms): 26-DOP bound
};

// Shared memory per scene is what bounds the number of resident warps (= scenes in flight) per SM, and the kernel is latency
// bound, so Scratch is kept small: buffers whose lifetimes never overlap share storage.  "pos" members are produced by step1
// (kinematics, collision, row construction) and are dead once make_rows / the derived-output write have run; "dyn" members
// live only inside step2 (solver, integrator).  Everything outside the two unions persists across both.
// C: the scene may hold re-classed props (sphere / cylinder colliders, condim-6 contacts).  The layout does not depend on it; the flag
// selects, at compile time, whether the stepping code carries those paths -- the box-only instantiation is the throughput kernel.
template <typename T, bool C = false>
struct alignas(16) Scratch {
  // persisted state (one HBM record per env: qpos | qvel | qacc_warmstart | time)
  T qpos[MAXNQ], qvel[MAXNV], qacc_ws[MAXNV], time, rec_pad[REC_SIZE - (MAXNQ + 2 * MAXNV + 1)], ctrl[MAXU];
  // per-env prop parameters
  T psize[MAXP][3], pinertia[MAXP][3], pinvw[MAXP][2]; int pactive[MAXP];
  T phalf[MAXP][3], pmass[MAXP]; unsigned char ptype[MAXP], gtype[MAXG]; int reclassed;   // collider class per prop (B2P_GEOM_*), bounding half extents, mass; gtype: this scene's geom types
  // tree (spatial quantities about the fixed reference point m.ref)
  T xpos[MAXTB][3], xmat[MAXTB][9], cdof[MAXTD][6];
  T ppos[MAXP][3], pmat[MAXP][9];
  T site_xpos[2][3], site_xmat[2][9]; unsigned char fkptr[MAXTB];   // fkptr: ancestor pointer of the pointer-jumping kinematics pass
  // dynamics: tree mass matrix (packed lower triangle; OSC, M*x) + its packed Cholesky factor; props diagonal
  T Mt[MPACK], Larm[MPACK], Mdiag[MAXNV];
  T qfrc_bias[MAXNV], qfrc_passive[MAXNV], qfrc_act[MAXNV], qfrc_smooth[MAXNV], qacc_smooth[MAXNV], qacc[MAXNV], qfrc_con[MAXNV];
  T act_force[MAXU];
  // contacts
  int ncand, ncon, nefc, flags, niter, jused, nmpr, nrow_nc; unsigned cpl_pairs;   // nrow_nc: rows ahead of the first contact row (equality + limit rows)
  T c_dist[MAXCON], c_mu[MAXCON], c_fric[MAXCON], c_incl[MAXCON];
  short c_pair[MAXCON], c_row[MAXCON]; unsigned char c_dim[MAXCON];   // c_dim: 3, or 6 for a contact of the cylinder class (torsional + rolling rows)
  // constraint rows.  Jacobian rows are packed: row r occupies Jp[joff[r] .. joff[r] + wa + wb) with segment A
  // (block row_a: ntd tree slots or 6 prop slots) followed by segment B (block row_b: 6 prop slots, or absent).
  T Jp[Pool<T>::JPS], D[Pool<T>::EFC], aref[Pool<T>::EFC], force[Pool<T>::EFC];
  T* xpool;                      // this warp's M^-1 J' area (PGS)
  T* w6;                         // this warp's condim-6 area (global memory, [MAXCON][W6_STRIDE]): packed 6 x 6 cone Hessian, torsional and rolling friction
  T* jovf;                       // this warp's overflow area (global memory, Pool<T>::JP reals): joff[r] < 0 <=> row r lives at jovf[-joff[r] - 1]
  signed char row_a[Pool<T>::EFC], row_b[Pool<T>::EFC], row_type[Pool<T>::EFC], row_state[Pool<T>::EFC];
  short row_id[Pool<T>::EFC], joff[Pool<T>::EFC], eq_row0[MAXEQ];   // eq_row0: first row of equality e
  // per-block row lists (block 0 = tree, 1+p = prop p): first row of every item (eq/limit row, or contact triple)
  unsigned char brow0[Pool<T>::EFC], browp[MAXP][MAXCON]; int nbrow[SP_MAXBLK];
  SpStruct sp;                   // block structure of the Newton Hessian in solver order (props first, tree last)
  // separation witnesses of hull pairs (b2p_collide.h: witness_separated): pair id (-1 = free), dA[3] | eB[3] | H ; ring cursor
  T sep_w[SEP_N][7]; short sep_pair[SEP_N]; int sep_next;
  // settle / frozen-arm support
  T hold_q[MAXTD], hold_v[MAXTD];
  union {                        // ---- storage shared between step1-only ("pos") and step2-only ("dyn") buffers, part 1
    T H[HPACK];                  // dyn: packed lower-triangular Newton Hessian / integrator matrix
    struct {                     // pos: composite inertias + rne sweeps, contact points, broad-phase lists
      T crb[MAXTB][10], cvel[MAXTB][6], cacc[MAXTB][6], cfrc[MAXTB][6], c_pos[MAXCON][3], xquat[MAXTB][4], grbound[MAXG];
      short cand[MAXCAND];
    };
  };
  union {                        // ---- part 2
    struct {                     // dyn: solver vectors, per-row residuals, per-contact cone Hessian (xx,yy,zz,xy,xz,yz)
      T grad[MAXNV], Mgrad[MAXNV], search[MAXNV], Ma[MAXNV], Mv[MAXNV], tmpv[MAXNV], Jaref[Pool<T>::EFC], Jv[Pool<T>::EFC], cW[MAXCON][6];
    };
    struct {                     // pos: geom poses, contact frames
      T gpos[MAXG][3], gmat[MAXG][9], c_frame[MAXCON][9];
    };
  };
};
static_assert(sizeof(((Scratch<float>*)0)->H) >= (MAXTB * 28 + MAXCON * 3 + MAXTB * 4 + MAXG) * sizeof(float) + MAXCAND * sizeof(short), "pos side must fit under H");


// block id of a dof: 0 = tree, 1+p = prop p
template <typename T> B2P_HD int dof_block(const DevModel<T>& m, int d) { return d < m.ntd ? 0 : 1 + (d - m.ntd) / 6; }
template <typename T> B2P_HD int block_start(const DevModel<T>& m, int blk) { return blk == 0 ? 0 : m.ntd + 6 * (blk - 1); }
template <typename T> B2P_HD int block_width(const DevModel<T>& m, int blk) { return blk == 0 ? m.ntd : 6; }
// first slot of packed Jacobian row r: shared-memory pool, or the warp's global overflow area
template <typename T, bool C> B2P_HD const T* jrow(const Scratch<T, C>& s, int r) { const int o = s.joff[r]; return o >= 0 ? s.Jp + o : s.jovf + (-o - 1); }
// pointer to the slots of block `blk` inside row r, or nullptr when the row does not touch the block
template <typename T, bool C> B2P_HD const T* row_seg(const DevModel<T>& m, const Scratch<T, C>& s, int r, int blk) {
  const int a = s.row_a[r];
  if (a == blk) return jrow(s, r);
  if (s.row_b[r] == blk) return jrow(s, r) + block_width(m, a);
  return nullptr;
}
template <typename T, bool C> B2P_HD const unsigned char* block_rows(const Scratch<T, C>& s, int blk) { return blk == 0 ? s.brow0 : s.browp[blk - 1]; }

// HBM state record of one env = the head of Scratch<T>: qpos[MAXNQ] | qvel[MAXNV] | qacc_warmstart[MAXNV] | time | pad
// launch parameters shared by kernel and host emulation
template <typename T>
struct StepArgs {
  int mode, n_env, n_steps, control_steps, state_stride, full_contacts, lockstep, lock_group, solver, need_classes;   // solver: 0 Newton, 1 PGS; need_classes: model_needs_classes() -- a pyramidal cone or a sphere / cylinder on the tree: launch the class-capable instantiations   // lock_group: warps per lockstep group (0 = whole CTA)
  T* state;                  // [n_env][state_stride]: qpos | qvel | qacc_warmstart | time
  const T* ctrl;             // MODE_PHYSICS / SETTLE: [n_env][nu]
  const T* target;           // MODE_OSC: [n_env][13] pos3 vel3 quat4(wxyz) angvel3
  const unsigned char* grip; // MODE_OSC: [n_env] 0 = min, 1 = max
  const T* prop_size;        // [n_env][MAXP][3] half sizes
  const unsigned char* prop_active; // [n_env][MAXP]
  const unsigned char* prop_type;   // optional [n_env][MAXP] collider class per prop: B2P_GEOM_BOX (as compiled) / SPHERE / CYLINDER; NULL = boxes
  const unsigned char* env_mask;    // optional [n_env]: 0 = skip env
  unsigned char* converged;  // MODE_OSC out [n_env]
  int* status;               // out [n_env] OR-ed flags
  int* steps_done;           // MODE_SETTLE out [n_env]
  T settle_qvel_tol, settle_qacc_tol; int settle_min_steps;
  T* out;                    // MODE_FORWARD / any: optional derived-output record per env (see OutLayout)
  int out_stride, out_mask;   // out_mask: groups of the record the stepping modes write (OUT_* of b2p_core.h); MODE_FORWARD writes all
  const int* order;          // optional [n_env]: queue position -> scene (scenes ordered by the work they did in the previous launch)
  int* work;                 // optional [n_env] out: work proxy of this launch, sum over steps of nefc * (2 + Newton iterations)
  // MODE_PLACE (expert prop_place): per scene the prop to place and the box [lo3 | hi3] proposals are drawn from
  const int* place_prop; const T* place_box; T place_dist; int place_table; unsigned place_seed;
  T* place_out;              // [n_env][4]: accepted position xyz, attempts used (negative: none accepted within n_steps attempts)
  int place_yaw; T* place_quat;   // place_yaw: draw a yaw about z per proposal (PropPlacer) instead of the gripper-down orientation; place_quat: optional [n_env][4] out
  const T* gains;            // MODE_OSC: optional [n_env][6] per-scene OSC gains (pos kp kd, ori kp kd, null kp kd); NULL = the model's
};
// derived-output record (reals): site_xpos[2][3] site_xmat[2][9] qacc[MAXNV] qfrc_bias[MAXNV] ctrl[MAXU] ncon nefc niter |
//   contacts[MAXCON] x (dist, pos3, normal3, geom1, geom2, excluded) | Marm[MAXTD*MAXTD] | gpos[MAXG][3] | gmat[MAXG][9]
// groups of the record (B2P_OUT_* of include/b2p.h)
enum { OUT_SITE = 1, OUT_CONTACT = 2, OUT_MARM = 4, OUT_GEOM = 8, OUT_BIAS = 16, OUT_DYN = 32, OUT_EFC = 64,
       OUT_POS = OUT_SITE | OUT_CONTACT | OUT_MARM | OUT_GEOM | OUT_BIAS, OUT_ALL = 127 };
struct OutLayout {
  static constexpr int SITE_XPOS = 0, SITE_XMAT = 6, QACC = 24, QFRC_BIAS = QACC + MAXNV, CTRL = QFRC_BIAS + MAXNV, NCON = CTRL + MAXU, NEFC = NCON + 1,
                       NITER = NEFC + 1, CONTACT = NITER + 1, CONTACT_W = 10, MARM = CONTACT + MAXCON * CONTACT_W, GPOS = MARM + MAXTD * MAXTD,
                       GMAT = GPOS + MAXG * 3, QACC_SMOOTH = GMAT + MAXG * 9, EFC_FORCE = QACC_SMOOTH + MAXNV, SIZE = EFC_FORCE + MAXEFC_OUT;
};

}  // namespace b2p
