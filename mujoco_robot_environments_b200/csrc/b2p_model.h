// Device-side model constants for the rearrangement model class, specialised:
//   * one kinematic tree (arm + gripper) whose moving bodies each carry exactly one hinge
//     (the compiler fuses welded bodies), tree body b <-> dof b-1;
//   * NP free "prop" bodies with identity inertial frame (box props), 6 dofs each;
//   * static geoms live on body 0.
// Unified body ids: 0 static/world, 1..ntb-1 tree bodies, PROP_BODY0+p prop p.
// DevModel<T> is a POD that the kernel stages into shared memory with one TMA bulk copy;
// large read-only tables (hull vertices, pair table) stay in global memory.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/b2p_model_layout.h"
#include "../../include/b2p_prop_types.h"

namespace b2p {

constexpr int MAXTD = 16;        // tree dofs (15 used)
constexpr int MAXTB = 16;        // tree bodies incl. world
constexpr int MAXP = 4;          // props
constexpr int MAXNV = MAXTD + 6 * MAXP;   // 40
constexpr int MAXNQ = MAXTD + 7 * MAXP;   // 44
constexpr int MAXG = 32;         // collidable geoms
constexpr int MAXCON = 48;       // contacts kept per env
constexpr int MAXCAND = 160;     // narrow-phase candidates per env
constexpr int MAXPAIR = 448;     // geom pairs of the static pair table (386 in the 4-prop scene)
constexpr int MAXEQ = 4;
constexpr int MAXU = 8;
constexpr int MAXLEVEL = 12;
constexpr int PROP_BODY0 = MAXTB;
// Constraint pools per scene, per dtype: rows (equality + limits + 3 per contact) and reals of packed Jacobian.  Shared memory per
// scene bounds the warps per SM: fp32 runs 9 scenes per SM with these sizes; fp64 runs 4 and has room for pools that hold
// MAXCON contacts x 3 rows + every equality / limit row, so the parity mode never drops a contact for lack of rows.
#ifndef B2P_EFC_F32
#define B2P_EFC_F32 128
#endif
#ifndef B2P_JPOOL_F32
#define B2P_JPOOL_F32 2304
#endif
// JP is the capacity of a scene's packed Jacobian; only the first JPS reals live in shared memory.  A scene at rest needs ~400
// (7 equality rows on the tree + 16 prop-table contacts x 3 rows x 6), so JPS is sized for that and the rows of heavier scenes
// (grasps, arm contacts) spill to a per-warp overflow area in global memory (L1/L2 resident): shared memory per scene is what
// bounds the scenes in flight per SM, and the kernel is latency bound.
#ifndef B2P_JPS_F32
#define B2P_JPS_F32 624
#endif
#ifndef B2P_JPS_F64
#define B2P_JPS_F64 576
#endif
template <typename T> struct Pool { static constexpr int EFC = B2P_EFC_F32, JP = B2P_JPOOL_F32, JPS = B2P_JPS_F32; };
template <> struct Pool<double> { static constexpr int EFC = 160, JP = 2944, JPS = B2P_JPS_F64; };
constexpr int MAXEFC_OUT = 160;  // efc_force slots in the derived-output record (both dtypes)
constexpr int SP_MAXBLK = 1 + MAXP;                  // solver blocks: props, then the tree
constexpr int HPACK = (MAXNV * (MAXNV + 1)) / 2;     // packed lower triangle of the Newton Hessian
constexpr int MPACK = (MAXTD * (MAXTD + 1)) / 2;     // packed lower triangle of the tree mass matrix
constexpr int TRI_N = ((MAXNV - 1) * MAXNV) / 2;     // pairs a >= c with a < MAXNV - 1 ... enumerated by DevModel::tri

// block-sparsity structure of a symmetric matrix in solver order (see b2p_linalg.h)
struct SpBlock { unsigned char start, len, nrows, nprev; };
struct SpStruct {
  int nblk; SpBlock blk[SP_MAXBLK]; unsigned char later[8];
  unsigned char rows[SP_MAXBLK][MAXNV];      // own rows, then rows of later coupled blocks (ascending)
  unsigned char prev[SP_MAXBLK][6 * MAXP];   // columns of earlier coupled blocks (ascending)
};

template <typename T>
struct PairParam {               // per candidate geom pair, mixed contact parameters (App. B item 10)
  T friction[3]; T solref[2]; T solimp[5]; T margin; T gap;
};
template <typename T>
struct GeomRaw {                 // a geom's own contact parameters: mixed on the device for pairs with a re-classed prop (sphere / cylinder)
  int priority, condim; T friction[3]; T solref[2]; T solimp[5]; T solmix, margin, gap;
};
constexpr int NPCLASS = 2;       // re-classable prop colliders beyond the box: 0 sphere, 1 cylinder (include/b2p_prop_types.h)

template <typename T>
struct alignas(16) DevModel {
  int nv, nq, ntd, ntb, nprop, ngeom, npair, neq, nu, nlevel;
  int iterations, ls_iterations, integrator, ccd_iterations, site_body[2], fk_rounds, cone;   // fk_rounds = ceil(log2(tree depth))
  T timestep, gravity[3], impratio, tolerance, ls_tolerance, ccd_tolerance, meaninertia, ref[3];
  // tree
  int tb_parent[MAXTB], level_start[MAXLEVEL + 1], level_body[MAXTB], child_start[MAXTB + 1], child_list[MAXTB];
  unsigned anc_mask[MAXTB];      // bit d set <=> tree dof d is on the chain of body b (incl. its own)
  T tb_pos[MAXTB][3], tb_quat[MAXTB][4], tb_ipos[MAXTB][3], tb_iquat[MAXTB][4], tb_mass[MAXTB], tb_inertia[MAXTB][3];
  T tb_invweight0[MAXTB][2];
  T tb_anchor[MAXTB][3], tb_axis[MAXTB][3];   // joint anchor and axis of body b's hinge in the PARENT body's frame: pos + R(quat) jnt_pos, R(quat) jnt_axis
  T jnt_pos[MAXTD][3], jnt_axis[MAXTD][3], jnt_stiffness[MAXTD], jnt_range[MAXTD][2], jnt_margin[MAXTD];
  T jnt_solref[MAXTD][2], jnt_solimp[MAXTD][5], qpos0[MAXTD], qpos_spring[MAXTD];
  T dof_armature[MAXTD], dof_damping[MAXTD], dof_invweight0[MAXTD];
  int jnt_limited[MAXTD];
  // props
  T prop_mass[MAXP];
  GeomRaw<T> pclass[NPCLASS]; T pclass_mass[NPCLASS];   // the sphere / cylinder prop classes (mass < 0: density 1000 x volume)
  // geoms
  int geom_type[MAXG], geom_body[MAXG], geom_vertadr[MAXG], geom_vertnum[MAXG], geom_prop[MAXG];
  T geom_size[MAXG][3], geom_pos[MAXG][3], geom_quat[MAXG][4], geom_rbound[MAXG], geom_invweight0[MAXG][2];
  T geom_aabb_c[MAXG][3];        // centre of the geom-frame bounding box (half extents = geom_size for boxes AND meshes)
  T geom_margin[MAXG], geom_gap[MAXG];
  unsigned short pair_pk[MAXPAIR];   // the pair table for the first broad-phase pass, g1 | g2 << 8: in shared memory with the rest of the model, so that pass has no global loads
  // sites: 0 attachment_site, 1 pinch
  T site_pos[2][3], site_quat[2][4];
  // equality
  int eq_type[MAXEQ], eq_obj1[MAXEQ], eq_obj2[MAXEQ];
  T eq_data[MAXEQ][6], eq_solref[MAXEQ][2], eq_solimp[MAXEQ][5];
  // actuators
  int act_trnnum[MAXU], act_trndof[MAXU][2], act_ctrllimited[MAXU], act_forcelimited[MAXU];
  T act_trncoef[MAXU][2], act_gain[MAXU], act_bias[MAXU][3], act_ctrlrange[MAXU][2], act_forcerange[MAXU][2];
  // OSC configuration (config/robots/arm/controller_config/osc.yaml)
  T osc_gains[6], osc_null[7], osc_pos_thresh, osc_ori_thresh;
  // linear-algebra tables: tri[e] = (a << 8) | c for the e-th pair a >= c ; dense structure of the tree block
  unsigned short tri[TRI_N + 3]; SpStruct sp_tree;
  T impl_hD[MPACK];             // packed lower triangle of -h * d(qfrc_passive + qfrc_actuator)/d(qvel) (integrator matrix = M + impl_hD)
};

// ------------------------------------------------------------------------------------------ host builder
// A model whose rows or geoms only the class-capable instantiations (C = true) carry: a pyramidal cone, or a sphere / cylinder geom on
// the tree.  Such a model is always launched with them, whether or not prop classes are bound (step_launch, tests/hostemu).
template <typename M> inline bool model_needs_classes(const M& m) {
  if (m.cone == B2P_CONE_PYRAMIDAL) return true;
  for (int g = 0; g < m.ngeom; g++) if (m.geom_type[g] == B2P_GEOM_SPHERE || m.geom_type[g] == B2P_GEOM_CYLINDER) return true;
  return false;
}
struct HostTables {              // global-memory tables, always fp64 on host; converted on upload
  std::vector<double> mesh_vert, geom_dop; std::vector<int> mesh_nbradr, mesh_nbrnum, mesh_nbr, geom_start;
  std::vector<int> pair_g1, pair_g2; std::vector<double> pair_param;  // 12 doubles per pair
  std::vector<GeomRaw<double>> geom_raw;
};

struct BlobView {
  const unsigned char* p; const int* hdr; const double* opt; const int* tab;
  explicit BlobView(const void* b) : p((const unsigned char*)b), hdr((const int*)b), opt((const double*)((const unsigned char*)b + B2PM_HEADER_BYTES)), tab((const int*)((const unsigned char*)b + B2PM_TABLE_OFFSET)) {}
  const int* I(int f) const { return (const int*)(p + tab[2 * f]); }
  const double* D(int f) const { return (const double*)(p + tab[2 * f]); }
  int n(int f) const { return tab[2 * f + 1]; }
  int C(int c) const { return hdr[c]; }
};

// Returns "" on success, else an error message.  Builds DevModel<double>; cast with convert<T>().
inline std::string build_model(const void* blob, size_t nbytes, DevModel<double>& m, HostTables& ht) {
  if (nbytes < 1024) return "blob too small";
  BlobView b(blob);
  if ((unsigned)b.hdr[0] != B2PM_MAGIC || b.hdr[1] != B2PM_VERSION) return "bad blob magic/version";
  std::memset(&m, 0, sizeof m);
  const int nbody = b.C(B2PM_C_NBODY), nv = b.C(B2PM_C_NV), ngeom = b.C(B2PM_C_NGEOM), nprop = b.C(B2PM_C_NPROP);
  const int ntd = b.C(B2PM_C_NARM) + b.C(B2PM_C_NGRIP), ntb = nbody - nprop;
  if (nprop > MAXP || ntd > MAXTD - 1 || ntb > MAXTB || ngeom > MAXG || nv != ntd + 6 * nprop) return "model exceeds kernel maxima";
  if (b.C(B2PM_C_NEQ) > MAXEQ || b.C(B2PM_C_NU) > MAXU) return "too many equalities / actuators";
  if (b.C(B2PM_C_NPAIR) > MAXPAIR) return "too many geom pairs";
  if ((int)b.opt[B2PM_O_CONE] != B2P_CONE_ELLIPTIC && (int)b.opt[B2PM_O_CONE] != B2P_CONE_PYRAMIDAL) return "unknown friction cone";
  m.nv = nv; m.nq = b.C(B2PM_C_NQ); m.ntd = ntd; m.ntb = ntb; m.nprop = nprop; m.ngeom = ngeom; m.npair = b.C(B2PM_C_NPAIR);
  m.neq = b.C(B2PM_C_NEQ); m.nu = b.C(B2PM_C_NU);
  m.timestep = b.opt[B2PM_O_TIMESTEP]; m.gravity[0] = b.opt[B2PM_O_GRAVITY_X]; m.gravity[1] = b.opt[B2PM_O_GRAVITY_Y]; m.gravity[2] = b.opt[B2PM_O_GRAVITY_Z];
  m.impratio = b.opt[B2PM_O_IMPRATIO]; m.tolerance = b.opt[B2PM_O_TOLERANCE]; m.ls_tolerance = b.opt[B2PM_O_LS_TOLERANCE];
  m.iterations = (int)b.opt[B2PM_O_ITERATIONS]; m.ls_iterations = (int)b.opt[B2PM_O_LS_ITERATIONS]; m.integrator = (int)b.opt[B2PM_O_INTEGRATOR]; m.cone = (int)b.opt[B2PM_O_CONE];
  m.ccd_iterations = (int)b.opt[B2PM_O_CCD_ITERATIONS]; m.ccd_tolerance = b.opt[B2PM_O_CCD_TOLERANCE]; m.meaninertia = b.opt[B2PM_O_MEANINERTIA];
  const int *par = b.I(B2PM_F_BODY_PARENTID), *jnum = b.I(B2PM_F_BODY_JNTNUM), *jadr = b.I(B2PM_F_BODY_JNTADR), *jtype = b.I(B2PM_F_JNT_TYPE);
  const int *dadr = b.I(B2PM_F_BODY_DOFADR), *bprop = b.I(B2PM_F_BODY_PROPID);
  // body layout check: bodies 1..ntb-1 = tree (one hinge each, dof = body-1), then props (free joint)
  for (int i = 1; i < nbody; i++) {
    if (i < ntb) {
      if (jnum[i] != 1 || jtype[jadr[i]] != B2P_JNT_HINGE || dadr[i] != i - 1 || par[i] >= i) return "tree body without exactly one hinge (compile with fuse=True)";
    } else {
      if (jnum[i] != 1 || jtype[jadr[i]] != B2P_JNT_FREE || bprop[i] != i - ntb || par[i] != 0) return "prop body layout unexpected";
      const double *ip = b.D(B2PM_F_BODY_IPOS) + 3 * i, *iq = b.D(B2PM_F_BODY_IQUAT) + 4 * i;
      if (std::fabs(ip[0]) + std::fabs(ip[1]) + std::fabs(ip[2]) > 1e-12 || std::fabs(std::fabs(iq[0]) - 1) > 1e-9) return "prop inertial frame must be identity";
    }
  }
  auto cp = [](double* d, const double* s, int n) { for (int i = 0; i < n; i++) d[i] = s[i]; };
  int depth[MAXTB] = {0}, maxd = 0;
  for (int i = 0; i < ntb; i++) {
    m.tb_parent[i] = i ? par[i] : 0;
    cp(m.tb_pos[i], b.D(B2PM_F_BODY_POS) + 3 * i, 3); cp(m.tb_quat[i], b.D(B2PM_F_BODY_QUAT) + 4 * i, 4);
    cp(m.tb_ipos[i], b.D(B2PM_F_BODY_IPOS) + 3 * i, 3); cp(m.tb_iquat[i], b.D(B2PM_F_BODY_IQUAT) + 4 * i, 4);
    m.tb_mass[i] = b.D(B2PM_F_BODY_MASS)[i]; cp(m.tb_inertia[i], b.D(B2PM_F_BODY_INERTIA) + 3 * i, 3);
    cp(m.tb_invweight0[i], b.D(B2PM_F_BODY_INVWEIGHT0) + 2 * i, 2);
    if (i) { depth[i] = depth[par[i]] + 1; if (depth[i] > maxd) maxd = depth[i]; m.anc_mask[i] = m.anc_mask[par[i]] | (1u << (i - 1)); }
  }
  if (maxd > MAXLEVEL) return "tree too deep";
  m.nlevel = maxd; int k = 0;
  m.fk_rounds = 0; while ((1 << m.fk_rounds) < maxd) m.fk_rounds++;
  for (int l = 1; l <= maxd; l++) { m.level_start[l - 1] = k; for (int i = 1; i < ntb; i++) if (depth[i] == l) m.level_body[k++] = i; }
  m.level_start[maxd] = k; k = 0;
  for (int i = 0; i < ntb; i++) { m.child_start[i] = k; for (int c = 1; c < ntb; c++) if (par[c] == i) m.child_list[k++] = c; }
  m.child_start[ntb] = k;
  for (int d = 0; d < ntd; d++) {
    cp(m.jnt_pos[d], b.D(B2PM_F_JNT_POS) + 3 * d, 3); cp(m.jnt_axis[d], b.D(B2PM_F_JNT_AXIS) + 3 * d, 3);
    m.jnt_stiffness[d] = b.D(B2PM_F_JNT_STIFFNESS)[d]; cp(m.jnt_range[d], b.D(B2PM_F_JNT_RANGE) + 2 * d, 2); m.jnt_margin[d] = b.D(B2PM_F_JNT_MARGIN)[d];
    cp(m.jnt_solref[d], b.D(B2PM_F_JNT_SOLREF) + 2 * d, 2); cp(m.jnt_solimp[d], b.D(B2PM_F_JNT_SOLIMP) + 5 * d, 5);
    m.qpos0[d] = b.D(B2PM_F_QPOS0)[d]; m.qpos_spring[d] = b.D(B2PM_F_QPOS_SPRING)[d];
    m.dof_armature[d] = b.D(B2PM_F_DOF_ARMATURE)[d]; m.dof_damping[d] = b.D(B2PM_F_DOF_DAMPING)[d]; m.dof_invweight0[d] = b.D(B2PM_F_DOF_INVWEIGHT0)[d];
    m.jnt_limited[d] = b.I(B2PM_F_JNT_LIMITED)[d];
  }
  for (int i = 1; i < ntb; i++) {   // constants of the kinematic pass: the hinge of body i as seen from its parent
    const double* q = m.tb_quat[i]; const int d = i - 1;
    const double w = q[0], x = q[1], y = q[2], z = q[3];
    const double R[9] = {w * w + x * x - y * y - z * z, 2 * (x * y - w * z), 2 * (x * z + w * y), 2 * (x * y + w * z), w * w - x * x + y * y - z * z, 2 * (y * z - w * x),
                         2 * (x * z - w * y), 2 * (y * z + w * x), w * w - x * x - y * y + z * z};
    for (int r = 0; r < 3; r++) {
      m.tb_anchor[i][r] = m.tb_pos[i][r] + R[3 * r] * m.jnt_pos[d][0] + R[3 * r + 1] * m.jnt_pos[d][1] + R[3 * r + 2] * m.jnt_pos[d][2];
      m.tb_axis[i][r] = R[3 * r] * m.jnt_axis[d][0] + R[3 * r + 1] * m.jnt_axis[d][1] + R[3 * r + 2] * m.jnt_axis[d][2];
    }
  }
  // spatial reference point: the (world-fixed) anchor of dof 0 at qpos0 = parent-chain of static frames
  cp(m.ref, b.D(B2PM_F_BODY_POS) + 3 * 1, 3);   // body 1's parent is the (fused) world
  for (int p = 0; p < nprop; p++) m.prop_mass[p] = b.D(B2PM_F_PROP_MASS)[p];
  {
    const b2p_prop_class* cls[NPCLASS] = {&B2P_PROP_CLASS_SPHERE, &B2P_PROP_CLASS_CYLINDER};
    for (int k = 0; k < NPCLASS; k++) {
      GeomRaw<double>& r = m.pclass[k]; r.priority = cls[k]->priority; r.condim = cls[k]->condim;
      for (int q = 0; q < 3; q++) r.friction[q] = cls[k]->friction[q];
      for (int q = 0; q < 2; q++) r.solref[q] = cls[k]->solref[q];
      for (int q = 0; q < 5; q++) r.solimp[q] = cls[k]->solimp[q];
      r.solmix = cls[k]->solmix; r.margin = cls[k]->margin; r.gap = cls[k]->gap; m.pclass_mass[k] = cls[k]->mass;
    }
  }
  const int *gt = b.I(B2PM_F_GEOM_TYPE), *gb = b.I(B2PM_F_GEOM_BODYID), *gm = b.I(B2PM_F_GEOM_MESHID), *gp = b.I(B2PM_F_GEOM_PROPID), *condim = b.I(B2PM_F_GEOM_CONDIM);
  const int *vadr = b.I(B2PM_F_MESH_VERTADR), *vnum = b.I(B2PM_F_MESH_VERTNUM);
  for (int g = 0; g < ngeom; g++) {
    if (condim[g] != 3) return "only condim 3 is implemented";
    m.geom_type[g] = gt[g]; m.geom_prop[g] = gp[g];
    m.geom_body[g] = gb[g] < ntb ? gb[g] : PROP_BODY0 + (gb[g] - ntb);
    m.geom_vertadr[g] = gm[g] >= 0 ? vadr[gm[g]] : 0; m.geom_vertnum[g] = gm[g] >= 0 ? vnum[gm[g]] : 0;
    cp(m.geom_size[g], b.D(B2PM_F_GEOM_SIZE) + 3 * g, 3); cp(m.geom_pos[g], b.D(B2PM_F_GEOM_POS) + 3 * g, 3); cp(m.geom_quat[g], b.D(B2PM_F_GEOM_QUAT) + 4 * g, 4);
    m.geom_rbound[g] = b.D(B2PM_F_GEOM_RBOUND)[g]; cp(m.geom_invweight0[g], b.D(B2PM_F_GEOM_INVWEIGHT0) + 2 * g, 2);
    cp(m.geom_aabb_c[g], b.D(B2PM_F_GEOM_AABB) + 6 * g, 3);
    // a cylinder on the tree (the drawing scene's tool piece): MuJoCo's size is (radius, half height, 0); stored here as (r, r, h) so that
    // the bounding-box tests read their half extents from geom_size like a box or a hull, and the narrow phase reads (r, h) at geom_size + 1
    if (gt[g] == B2P_GEOM_CYLINDER && gp[g] < 0) { m.geom_size[g][2] = m.geom_size[g][1]; m.geom_size[g][1] = m.geom_size[g][0]; }
    if (gt[g] == B2P_GEOM_MESH) for (int k = 0; k < 3; k++) if (std::fabs(b.D(B2PM_F_GEOM_AABB)[6 * g + 3 + k] - m.geom_size[g][k]) > 1e-9) return "mesh geom_size must equal its bounding-box half extents";
  }
  const int sid[2] = {b.C(B2PM_C_SITE_ATTACHMENT), b.C(B2PM_C_SITE_PINCH)};
  for (int s = 0; s < 2; s++) {
    m.site_body[s] = b.I(B2PM_F_SITE_BODYID)[sid[s]];
    if (m.site_body[s] >= ntb) return "site not on the arm tree";
    cp(m.site_pos[s], b.D(B2PM_F_SITE_POS) + 3 * sid[s], 3); cp(m.site_quat[s], b.D(B2PM_F_SITE_QUAT) + 4 * sid[s], 4);
  }
  for (int e = 0; e < m.neq; e++) {
    m.eq_type[e] = b.I(B2PM_F_EQ_TYPE)[e]; m.eq_obj1[e] = b.I(B2PM_F_EQ_OBJ1ID)[e]; m.eq_obj2[e] = b.I(B2PM_F_EQ_OBJ2ID)[e];
    if (m.eq_type[e] == B2P_EQ_CONNECT && (m.eq_obj1[e] >= ntb || m.eq_obj2[e] >= ntb)) return "connect equality must be on the tree";
    cp(m.eq_data[e], b.D(B2PM_F_EQ_DATA) + 11 * e, 6); cp(m.eq_solref[e], b.D(B2PM_F_EQ_SOLREF) + 2 * e, 2); cp(m.eq_solimp[e], b.D(B2PM_F_EQ_SOLIMP) + 5 * e, 5);
    if (m.eq_type[e] == B2P_EQ_JOINT) { const double* c = b.D(B2PM_F_EQ_DATA) + 11 * e; if (c[2] != 0 || c[3] != 0 || c[4] != 0) return "only linear joint equalities are implemented"; }
  }
  for (int a = 0; a < m.nu; a++) {
    m.act_trnnum[a] = b.I(B2PM_F_ACT_TRNNUM)[a];
    for (int x = 0; x < 2; x++) { m.act_trndof[a][x] = b.I(B2PM_F_ACT_TRNDOF)[2 * a + x]; m.act_trncoef[a][x] = b.D(B2PM_F_ACT_TRNCOEF)[2 * a + x]; if (x < m.act_trnnum[a] && m.act_trndof[a][x] >= ntd) return "actuator on a prop dof"; }
    m.act_ctrllimited[a] = b.I(B2PM_F_ACT_CTRLLIMITED)[a]; m.act_forcelimited[a] = b.I(B2PM_F_ACT_FORCELIMITED)[a];
    m.act_gain[a] = b.D(B2PM_F_ACT_GAIN)[a]; cp(m.act_bias[a], b.D(B2PM_F_ACT_BIAS) + 3 * a, 3);
    cp(m.act_ctrlrange[a], b.D(B2PM_F_ACT_CTRLRANGE) + 2 * a, 2); cp(m.act_forcerange[a], b.D(B2PM_F_ACT_FORCERANGE) + 2 * a, 2);
  }
  // The kernel folds osc.py:224-236 (tau @ pinv(actuator_moment)[arm, arm]) to the identity (b2p_osc.h).  That holds exactly when
  // the moment row of actuator a < narm is the unit vector of dof a (unit gear, joint transmission) and no other actuator reaches
  // an arm dof; anything else is refused here rather than silently mis-scaled.  (Gain and bias do not enter actuator_moment; the
  // reference does not compensate them either.)
  {
    const int narm = b.C(B2PM_C_NARM);
    if (m.nu < narm) return "OSC needs one actuator per arm joint";
    for (int a = 0; a < m.nu; a++) {
      if (a < narm) {
        if (m.act_trnnum[a] != 1 || m.act_trndof[a][0] != a || m.act_trncoef[a][0] != 1.0)
          return "arm actuators must be unit-gear joint actuators on dofs 0..narm-1 (OSC actuator-moment inverse is folded to identity)";
      } else {
        for (int x = 0; x < m.act_trnnum[a]; x++) if (m.act_trndof[a][x] < narm) return "a non-arm actuator drives an arm dof (OSC actuator-moment inverse is folded to identity)";
      }
    }
  }
  const double g6[6] = {350, 20, 500, 100, 200, 30}, q0[7] = {0, -0.785, 0, -2.356, 0, 1.571, 0.785};
  cp(m.osc_gains, g6, 6); cp(m.osc_null, q0, 7); m.osc_pos_thresh = 5e-3; m.osc_ori_thresh = 68e-3;
  { int e = 0; for (int a = 0; a < MAXNV - 1; a++) for (int cc = 0; cc <= a; cc++) m.tri[e++] = (unsigned short)((a << 8) | cc); }
  m.sp_tree.nblk = 1; m.sp_tree.blk[0].start = 0; m.sp_tree.blk[0].len = (unsigned char)ntd; m.sp_tree.blk[0].nrows = (unsigned char)ntd; m.sp_tree.blk[0].nprev = 0;
  for (int k = 0; k < ntd; k++) m.sp_tree.rows[0][k] = (unsigned char)k;
  {   // integrator matrix correction (App. B item 15): Euler = h*damping on the diagonal; implicitfast adds the affine actuator's
      // velocity term -h * bias_vel * coef_i * coef_j wherever (i, j) lies inside M's (ancestor) sparsity pattern
    const double h = m.timestep;
    for (int i = 0; i < ntd; i++) for (int j = 0; j <= i; j++) {
      double v = 0;
      if (i == j) v += h * m.dof_damping[i];
      if (m.integrator == B2P_INT_IMPLICITFAST && (i == j || (m.anc_mask[i + 1] & (1u << j))))
        for (int a = 0; a < m.nu; a++) if (m.act_bias[a][2] != 0) for (int x = 0; x < m.act_trnnum[a]; x++) for (int y = 0; y < m.act_trnnum[a]; y++)
          if (m.act_trndof[a][x] == i && m.act_trndof[a][y] == j) v -= h * m.act_trncoef[a][x] * m.act_trncoef[a][y] * m.act_bias[a][2];
      m.impl_hD[(i * (i + 1)) / 2 + j] = v;
    }
  }
  // global tables
  ht.mesh_vert.assign(b.D(B2PM_F_MESH_VERT), b.D(B2PM_F_MESH_VERT) + b.n(B2PM_F_MESH_VERT));
  ht.mesh_nbradr.assign(b.I(B2PM_F_MESH_NBRADR), b.I(B2PM_F_MESH_NBRADR) + b.n(B2PM_F_MESH_NBRADR));
  ht.mesh_nbrnum.assign(b.I(B2PM_F_MESH_NBRNUM), b.I(B2PM_F_MESH_NBRNUM) + b.n(B2PM_F_MESH_NBRNUM));
  ht.mesh_nbr.assign(b.I(B2PM_F_MESH_NBR), b.I(B2PM_F_MESH_NBR) + b.n(B2PM_F_MESH_NBR));
  // neighbour indices are mesh-local; make adr/num per global vertex and nbr global
  {
    const int nmesh = b.C(B2PM_C_NMESH);
    for (int me = 0; me < nmesh; me++) for (int v = vadr[me]; v < vadr[me] + vnum[me]; v++)
      for (int q = ht.mesh_nbradr[v]; q < ht.mesh_nbradr[v] + ht.mesh_nbrnum[v]; q++) ht.mesh_nbr[q] += vadr[me];
  }
  // hill-climb start table: for every mesh geom and each of the 26 quantised directions the hull's extreme vertex (global index)
  // ... and the support VALUE along the same 26 (unnormalised) directions: a 26-DOP around the hull.  The support function is
  // sublinear, so for any direction written as a non-negative combination of three of these the combination of the tabulated
  // values bounds the hull's support from above -- the broad phase rejects hull pairs with it without touching the hull graph.
  ht.geom_start.assign((size_t)MAXG * 27, 0); ht.geom_dop.assign((size_t)MAXG * 27, 0.0);
  for (int g = 0; g < m.ngeom; g++) {
    if (m.geom_vertnum[g] <= 0) continue;
    for (int q = 0; q < 27; q++) {
      const double d[3] = {double(q / 9 - 1), double((q / 3) % 3 - 1), double(q % 3 - 1)};
      int best = m.geom_vertadr[g]; double bd = -1e300;
      for (int v = m.geom_vertadr[g]; v < m.geom_vertadr[g] + m.geom_vertnum[g]; v++) {
        const double x = ht.mesh_vert[3 * v] * d[0] + ht.mesh_vert[3 * v + 1] * d[1] + ht.mesh_vert[3 * v + 2] * d[2];
        if (x > bd) { bd = x; best = v; }
      }
      ht.geom_start[(size_t)g * 27 + q] = best;
      ht.geom_dop[(size_t)g * 27 + q] = q == 13 ? 0.0 : bd * (1 + 1e-6) + 1e-9;   // rounded up: stays an upper bound after the cast to float
    }
  }
  const int *p1 = b.I(B2PM_F_PAIR_GEOM1), *p2 = b.I(B2PM_F_PAIR_GEOM2), *prio = b.I(B2PM_F_GEOM_PRIORITY);
  const double *fr = b.D(B2PM_F_GEOM_FRICTION), *sr = b.D(B2PM_F_GEOM_SOLREF), *si = b.D(B2PM_F_GEOM_SOLIMP), *sm = b.D(B2PM_F_GEOM_SOLMIX);
  const double *mg = b.D(B2PM_F_GEOM_MARGIN), *gg = b.D(B2PM_F_GEOM_GAP);
  ht.geom_raw.resize(m.ngeom);
  for (int q = 0; q < m.ngeom; q++) {
    GeomRaw<double>& r = ht.geom_raw[q]; r.priority = prio[q]; r.condim = condim[q];
    for (int k = 0; k < 3; k++) r.friction[k] = fr[3 * q + k];
    for (int k = 0; k < 2; k++) r.solref[k] = sr[2 * q + k];
    for (int k = 0; k < 5; k++) r.solimp[k] = si[5 * q + k];
    r.solmix = sm[q]; r.margin = mg[q]; r.gap = gg[q];
  }
  for (int q = 0; q < m.ngeom; q++) { m.geom_margin[q] = mg[q]; m.geom_gap[q] = gg[q]; }
  for (int p = 0; p < m.npair; p++) m.pair_pk[p] = (unsigned short)(p1[p] | (p2[p] << 8));
  for (int p = 0; p < m.npair; p++) {
    int g1 = p1[p], g2 = p2[p]; double pp[12];
    if (prio[g1] != prio[g2]) {
      int g = prio[g1] > prio[g2] ? g1 : g2;
      cp(pp, fr + 3 * g, 3); cp(pp + 3, sr + 2 * g, 2); cp(pp + 5, si + 5 * g, 5);
    } else {
      for (int q = 0; q < 3; q++) pp[q] = fr[3 * g1 + q] > fr[3 * g2 + q] ? fr[3 * g1 + q] : fr[3 * g2 + q];
      double mix;
      if (sm[g1] >= 1e-15 && sm[g2] >= 1e-15) mix = sm[g1] / (sm[g1] + sm[g2]);
      else if (sm[g1] < 1e-15 && sm[g2] < 1e-15) mix = 0.5; else mix = sm[g1] < 1e-15 ? 0 : 1;
      if (sr[2 * g1] > 0 && sr[2 * g2] > 0) for (int q = 0; q < 2; q++) pp[3 + q] = mix * sr[2 * g1 + q] + (1 - mix) * sr[2 * g2 + q];
      else for (int q = 0; q < 2; q++) pp[3 + q] = sr[2 * g1 + q] < sr[2 * g2 + q] ? sr[2 * g1 + q] : sr[2 * g2 + q];
      for (int q = 0; q < 5; q++) pp[5 + q] = mix * si[5 * g1 + q] + (1 - mix) * si[5 * g2 + q];
    }
    pp[10] = mg[g1] > mg[g2] ? mg[g1] : mg[g2]; pp[11] = gg[g1] > gg[g2] ? gg[g1] : gg[g2];
    ht.pair_g1.push_back(g1); ht.pair_g2.push_back(g2); ht.pair_param.insert(ht.pair_param.end(), pp, pp + 12);
  }
  return "";
}

template <typename T>
inline void convert_model(const DevModel<double>& s, DevModel<T>& d) {
  // field-wise cast: both structs share the member order, ints copy, reals cast
  std::memset(&d, 0, sizeof d);
#define CI(f) std::memcpy(&d.f, &s.f, sizeof d.f)
#define CR(f) do { const double* a = (const double*)&s.f; T* o = (T*)&d.f; for (size_t i = 0; i < sizeof(s.f) / sizeof(double); i++) o[i] = (T)a[i]; } while (0)
  CI(nv); CI(nq); CI(ntd); CI(ntb); CI(nprop); CI(ngeom); CI(npair); CI(neq); CI(nu); CI(nlevel);
  CI(iterations); CI(ls_iterations); CI(integrator); CI(ccd_iterations); CI(site_body); CI(fk_rounds); CI(cone);
  CR(timestep); CR(gravity); CR(impratio); CR(tolerance); CR(ls_tolerance); CR(ccd_tolerance); CR(meaninertia); CR(ref);
  CI(tb_parent); CI(level_start); CI(level_body); CI(child_start); CI(child_list); CI(anc_mask);
  CR(tb_pos); CR(tb_quat); CR(tb_ipos); CR(tb_iquat); CR(tb_mass); CR(tb_inertia); CR(tb_invweight0); CR(tb_anchor); CR(tb_axis);
  CR(jnt_pos); CR(jnt_axis); CR(jnt_stiffness); CR(jnt_range); CR(jnt_margin); CR(jnt_solref); CR(jnt_solimp); CR(qpos0); CR(qpos_spring);
  CR(dof_armature); CR(dof_damping); CR(dof_invweight0); CI(jnt_limited); CR(prop_mass); CR(pclass_mass);
  for (int k = 0; k < NPCLASS; k++) { d.pclass[k].priority = s.pclass[k].priority; d.pclass[k].condim = s.pclass[k].condim; for (int q = 0; q < 3; q++) d.pclass[k].friction[q] = (T)s.pclass[k].friction[q]; for (int q = 0; q < 2; q++) d.pclass[k].solref[q] = (T)s.pclass[k].solref[q]; for (int q = 0; q < 5; q++) d.pclass[k].solimp[q] = (T)s.pclass[k].solimp[q]; d.pclass[k].solmix = (T)s.pclass[k].solmix; d.pclass[k].margin = (T)s.pclass[k].margin; d.pclass[k].gap = (T)s.pclass[k].gap; }
  CI(geom_type); CI(geom_body); CI(geom_vertadr); CI(geom_vertnum); CI(geom_prop);
  CR(geom_size); CR(geom_pos); CR(geom_quat); CR(geom_rbound); CR(geom_invweight0); CR(geom_aabb_c); CR(geom_margin); CR(geom_gap); CI(pair_pk); CR(site_pos); CR(site_quat);
  CI(eq_type); CI(eq_obj1); CI(eq_obj2); CR(eq_data); CR(eq_solref); CR(eq_solimp);
  CI(act_trnnum); CI(act_trndof); CI(act_ctrllimited); CI(act_forcelimited);
  CR(act_trncoef); CR(act_gain); CR(act_bias); CR(act_ctrlrange); CR(act_forcerange);
  CR(osc_gains); CR(osc_null); CR(osc_pos_thresh); CR(osc_ori_thresh); CI(tri); CI(sp_tree); CR(impl_hD);
#undef CI
#undef CR
}

}  // namespace b2p
