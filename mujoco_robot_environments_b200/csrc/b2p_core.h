// B200-native batched physics + OSC for the rearrangement model class: per-env algorithms.
//
// One CTA steps one scene ("block-per-env"): all scratch of a scene lives in shared memory
// (Scratch<T>), model constants are staged once per CTA (DevModel<T>), parallel loops are written
// as `for (i = tid; i < n; i += nt)` separated by block barriers.  The same templates compile
// for the host with nt == 1 (barriers become no-ops): that build exists ONLY so the CPU test-suite
// can check the kernel logic against the oracle without a GPU (tests/_hostemu); the product path
// is the CUDA build in b2p_api.cu and has no CPU fallback.
//
// Replaces, for this model class, MuJoCo 3.2.7's mj_step1/mj_step2 as driven by the reference's
// RobotArm.run_controller (mujoco_robot_environments/models/robot_arm.py:61-94) and the OSC maths of
// mujoco_controllers/osc.py:99-247.  Stage semantics: SURVEY.md Appendix B.
#pragma once
#include "b2p_model.h"

#if defined(__CUDACC__)
#define B2P_HD __host__ __device__ __forceinline__
#define B2P_HDN __host__ __device__ __noinline__
#else
#define B2P_HD inline
#define B2P_HDN
#endif

namespace b2p {

enum { ROW_EQ = 0, ROW_LIMIT = 1, ROW_CONTACT = 2 };
enum { ST_SATISFIED = 0, ST_QUADRATIC = 1, ST_CONE = 2 };
enum { FLAG_BADQPOS = 1, FLAG_BADQVEL = 2, FLAG_BADQACC = 4, FLAG_CONTACTFULL = 8, FLAG_EFCFULL = 16, FLAG_CANDFULL = 32, FLAG_SOLVER_MAXITER = 64 };
enum { MODE_PHYSICS = 0, MODE_OSC = 1, MODE_SETTLE = 2, MODE_FORWARD = 3 };

struct Ctx { int tid, nt; };

template <typename T> struct Num;
template <> struct Num<double> { static B2P_HD double eps() { return 2.220446049250313e-16; } static B2P_HD double minval() { return 1e-15; } static B2P_HD double big() { return 1e300; } };
template <> struct Num<float> { static B2P_HD float eps() { return 1.1920929e-7f; } static B2P_HD float minval() { return 1e-15f; } static B2P_HD float big() { return 1e30f; } };

// ------------------------------------------------------------------------------------------ block primitives
B2P_HD void bsync(const Ctx&) {
#ifdef __CUDA_ARCH__
  __syncthreads();
#endif
}
B2P_HD void wsync() {
#ifdef __CUDA_ARCH__
  __syncwarp();
#endif
}
// deterministic block-wide sum of K values; red must hold K*32 T.  All threads return the same sums.
template <typename T, int K>
B2P_HD void block_sum(const Ctx& c, T (&v)[K], T* red) {
#ifdef __CUDA_ARCH__
  const int lane = c.tid & 31, w = c.tid >> 5, nw = (c.nt + 31) >> 5;
#pragma unroll
  for (int k = 0; k < K; k++) {
    T x = v[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
    if (lane == 0) red[k * 32 + w] = x;
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < K; k++) { T s = 0; for (int i = 0; i < nw; i++) s += red[k * 32 + i]; v[k] = s; }
  __syncthreads();
#else
  (void)c; (void)v; (void)red;
#endif
}
// exclusive prefix sum of a small non-negative int over the block (thread order); total returned to all
B2P_HD int block_scan(const Ctx& c, int v, int& total, int* sm /* >= 33 ints */) {
#ifdef __CUDA_ARCH__
  const int lane = c.tid & 31, w = c.tid >> 5, nw = (c.nt + 31) >> 5;
  int x = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
  if (lane == 31) sm[w] = x;
  __syncthreads();
  int base = 0, tot = 0;
  for (int i = 0; i < nw; i++) { if (i < w) base += sm[i]; tot += sm[i]; }
  __syncthreads();
  total = tot;
  return base + x - v;
#else
  (void)c; (void)sm; total = v; return 0;
#endif
}

// ------------------------------------------------------------------------------------------ small math
template <typename T> B2P_HD T dot3(const T* a, const T* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
template <typename T> B2P_HD void cross3(T* r, const T* a, const T* b) {
  T x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2], z = a[0] * b[1] - a[1] * b[0];
  r[0] = x; r[1] = y; r[2] = z;
}
template <typename T> B2P_HD T norm3(const T* a) { return sqrt(dot3(a, a)); }
template <typename T> B2P_HD T normalize3(T* a) {
  T n = norm3(a);
  if (n < Num<T>::minval()) { a[0] = 1; a[1] = 0; a[2] = 0; return 0; }
  T i = 1 / n; a[0] *= i; a[1] *= i; a[2] *= i; return n;
}
template <typename T> B2P_HD void cp3(T* d, const T* s) { d[0] = s[0]; d[1] = s[1]; d[2] = s[2]; }
template <typename T> B2P_HD void mulquat(T* r, const T* a, const T* b) {
  T w = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3], x = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
  T y = a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1], z = a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0];
  r[0] = w; r[1] = x; r[2] = y; r[3] = z;
}
template <typename T> B2P_HD void normquat(T* q) {
  T n = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  if (n < Num<T>::minval()) { q[0] = 1; q[1] = q[2] = q[3] = 0; } else { T i = 1 / n; q[0] *= i; q[1] *= i; q[2] *= i; q[3] *= i; }
}
template <typename T> B2P_HD void quat2mat(T* m, const T* q) {
  T w = q[0], x = q[1], y = q[2], z = q[3];
  m[0] = w * w + x * x - y * y - z * z; m[1] = 2 * (x * y - w * z); m[2] = 2 * (x * z + w * y);
  m[3] = 2 * (x * y + w * z); m[4] = w * w - x * x + y * y - z * z; m[5] = 2 * (y * z - w * x);
  m[6] = 2 * (x * z - w * y); m[7] = 2 * (y * z + w * x); m[8] = w * w - x * x - y * y + z * z;
}
template <typename T> B2P_HD void mulmv(T* r, const T* m, const T* v) {
  T x = m[0] * v[0] + m[1] * v[1] + m[2] * v[2], y = m[3] * v[0] + m[4] * v[1] + m[5] * v[2], z = m[6] * v[0] + m[7] * v[1] + m[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
template <typename T> B2P_HD void mulmtv(T* r, const T* m, const T* v) {
  T x = m[0] * v[0] + m[3] * v[1] + m[6] * v[2], y = m[1] * v[0] + m[4] * v[1] + m[7] * v[2], z = m[2] * v[0] + m[5] * v[1] + m[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
template <typename T> B2P_HD void mat2quat(T* q, const T* m) {   // mju_mat2Quat branch order (App. B item 17)
  if (m[0] + m[4] + m[8] > 0) {
    q[0] = T(0.5) * sqrt(1 + m[0] + m[4] + m[8]); q[1] = T(0.25) * (m[7] - m[5]) / q[0]; q[2] = T(0.25) * (m[2] - m[6]) / q[0]; q[3] = T(0.25) * (m[3] - m[1]) / q[0];
  } else if (m[0] > m[4] && m[0] > m[8]) {
    q[1] = T(0.5) * sqrt(1 + m[0] - m[4] - m[8]); q[0] = T(0.25) * (m[7] - m[5]) / q[1]; q[2] = T(0.25) * (m[1] + m[3]) / q[1]; q[3] = T(0.25) * (m[2] + m[6]) / q[1];
  } else if (m[4] > m[8]) {
    q[2] = T(0.5) * sqrt(1 - m[0] + m[4] - m[8]); q[0] = T(0.25) * (m[2] - m[6]) / q[2]; q[1] = T(0.25) * (m[1] + m[3]) / q[2]; q[3] = T(0.25) * (m[5] + m[7]) / q[2];
  } else {
    q[3] = T(0.5) * sqrt(1 - m[0] - m[4] + m[8]); q[0] = T(0.25) * (m[3] - m[1]) / q[3]; q[1] = T(0.25) * (m[2] + m[6]) / q[3]; q[2] = T(0.25) * (m[5] + m[7]) / q[3];
  }
  normquat(q);
}
template <typename T> B2P_HD void make_frame(T* f) {   // mju_makeFrame
  normalize3(f);
  f[3] = 0; f[4] = 0; f[5] = 0;
  if (f[1] < T(0.5) && f[1] > T(-0.5)) f[4] = 1; else f[5] = 1;
  T d = dot3(f, f + 3);
  f[3] -= d * f[0]; f[4] -= d * f[1]; f[5] -= d * f[2];
  normalize3(f + 3);
  cross3(f + 6, f, f + 3);
}
template <typename T> B2P_HD void inert_com(T* res, const T* inert, const T* mat, const T* dif, T mass) {
  T I[9];
  for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++)
    I[3 * r + c] = mat[3 * r] * inert[0] * mat[3 * c] + mat[3 * r + 1] * inert[1] * mat[3 * c + 1] + mat[3 * r + 2] * inert[2] * mat[3 * c + 2];
  res[0] = I[0] + mass * (dif[1] * dif[1] + dif[2] * dif[2]); res[1] = I[4] + mass * (dif[0] * dif[0] + dif[2] * dif[2]);
  res[2] = I[8] + mass * (dif[0] * dif[0] + dif[1] * dif[1]); res[3] = I[1] - mass * dif[0] * dif[1];
  res[4] = I[2] - mass * dif[0] * dif[2]; res[5] = I[5] - mass * dif[1] * dif[2];
  res[6] = mass * dif[0]; res[7] = mass * dif[1]; res[8] = mass * dif[2]; res[9] = mass;
}
template <typename T> B2P_HD void mul_inert_vec(T* r, const T* i, const T* v) {
  r[0] = i[0] * v[0] + i[3] * v[1] + i[4] * v[2] - i[8] * v[4] + i[7] * v[5];
  r[1] = i[3] * v[0] + i[1] * v[1] + i[5] * v[2] + i[8] * v[3] - i[6] * v[5];
  r[2] = i[4] * v[0] + i[5] * v[1] + i[2] * v[2] - i[7] * v[3] + i[6] * v[4];
  r[3] = i[8] * v[1] - i[7] * v[2] + i[9] * v[3];
  r[4] = i[6] * v[2] - i[8] * v[0] + i[9] * v[4];
  r[5] = i[7] * v[0] - i[6] * v[1] + i[9] * v[5];
}
template <typename T> B2P_HD void cross_motion(T* r, const T* vel, const T* v) {
  T a[3], b[3], c[3];
  cross3(a, vel, v); cross3(b, vel, v + 3); cross3(c, vel + 3, v);
  r[0] = a[0]; r[1] = a[1]; r[2] = a[2]; r[3] = b[0] + c[0]; r[4] = b[1] + c[1]; r[5] = b[2] + c[2];
}
template <typename T> B2P_HD void cross_force(T* r, const T* vel, const T* f) {
  T a[3], b[3], c[3];
  cross3(a, vel, f); cross3(b, vel + 3, f + 3); cross3(c, vel, f + 3);
  r[0] = a[0] + b[0]; r[1] = a[1] + b[1]; r[2] = a[2] + b[2]; r[3] = c[0]; r[4] = c[1]; r[5] = c[2];
}

// ------------------------------------------------------------------------------------------ per-env scratch
// HBM state record of one env = the head of Scratch<T>: qpos[MAXNQ] | qvel[MAXNV] | qacc_warmstart[MAXNV] | time | pad
constexpr int REC_QPOS = 0, REC_QVEL = MAXNQ, REC_QACC = MAXNQ + MAXNV, REC_TIME = MAXNQ + 2 * MAXNV, REC_SIZE = 128;
template <typename T>
struct Globals {                 // read-only global-memory tables
  const T* mesh_vert; const int* nbradr; const int* nbrnum; const int* nbr;
  const short* pair_g1; const short* pair_g2; const PairParam<T>* pair_param;
};

template <typename T>
struct Scratch {
  // persisted state (one HBM record per env: qpos | qvel | qacc_warmstart | time)
  T qpos[MAXNQ], qvel[MAXNV], qacc_ws[MAXNV], time, rec_pad[REC_SIZE - (MAXNQ + 2 * MAXNV + 1)], ctrl[MAXU];
  // per-env prop parameters
  T psize[MAXP][3], pinertia[MAXP][3], pinvw[MAXP][2]; int pactive[MAXP];
  // tree
  T xpos[MAXTB][3], xquat[MAXTB][4], xmat[MAXTB][9], xanchor[MAXTD][3], xaxis[MAXTD][3];
  T cinert[MAXTB][10], crb[MAXTB][10], cdof[MAXTD][6], cdof_dot[MAXTD][6], cvel[MAXTB][6], cacc[MAXTB][6], cfrc[MAXTB][6];
  T ppos[MAXP][3], pmat[MAXP][9];
  T gpos[MAXG][3], gmat[MAXG][9], grbound[MAXG], site_xpos[2][3], site_xmat[2][9];
  // dynamics
  T Marm[MAXTD * MAXTD], Larm[MAXTD * MAXTD], Mdiag[MAXNV];
  T qfrc_bias[MAXNV], qfrc_passive[MAXNV], qfrc_act[MAXNV], qfrc_smooth[MAXNV], qacc_smooth[MAXNV], qacc[MAXNV], qfrc_con[MAXNV];
  T act_force[MAXU];
  // contacts
  int ncand, ncon, nefc, flags, niter;
  short cand[MAXCAND];
  T c_dist[MAXCON], c_pos[MAXCON][3], c_frame[MAXCON][9], c_mu[MAXCON], c_fric[MAXCON][2], c_incl[MAXCON];
  short c_pair[MAXCON], c_row[MAXCON];
  // constraint rows (compressed Jacobian: slots 0..14 segment A, 15..20 segment B)
  T J[MAXEFC][JW], D[MAXEFC], aref[MAXEFC], Jaref[MAXEFC], Jv[MAXEFC], force[MAXEFC];
  signed char row_a[MAXEFC], row_b[MAXEFC], row_type[MAXEFC], row_state[MAXEFC];
  short row_id[MAXEFC];
  T cW[MAXCON][6];               // per-contact 3x3 symmetric Hessian weight (xx,yy,zz,xy,xz,yz) in row space
  // solver
  T H[MAXNV * MAXNV], grad[MAXNV], Mgrad[MAXNV], search[MAXNV], Ma[MAXNV], Mv[MAXNV], tmpv[MAXNV];
  // settle / frozen-arm support
  T hold_q[MAXTD], hold_v[MAXTD];
  // reductions
  T red[3 * 32]; int iscan[40];
};

// block id of a dof: 0 = tree, 1+p = prop p
template <typename T> B2P_HD int dof_block(const DevModel<T>& m, int d) { return d < m.ntd ? 0 : 1 + (d - m.ntd) / 6; }
template <typename T> B2P_HD int block_start(const DevModel<T>& m, int blk) { return blk == 0 ? 0 : m.ntd + 6 * (blk - 1); }
// slot of dof d inside row r, or -1
template <typename T> B2P_HD int row_slot(const DevModel<T>& m, const Scratch<T>& s, int r, int d) {
  int blk = dof_block(m, d), a = s.row_a[r], b = s.row_b[r];
  if (a == blk) return d - block_start(m, blk);
  if (b == blk) return SEGB + d - block_start(m, blk);
  return -1;
}

// ------------------------------------------------------------------------------------------ mj_kinematics + comPos + comVel + rne (forward sweep)
template <typename T>
B2P_HD void forward_sweep(const DevModel<T>& m, Scratch<T>& s, const Ctx& c) {
  if (c.tid == 0) {
    for (int k = 0; k < 3; k++) s.xpos[0][k] = 0;
    s.xquat[0][0] = 1; s.xquat[0][1] = 0; s.xquat[0][2] = 0; s.xquat[0][3] = 0;
    quat2mat(s.xmat[0], s.xquat[0]);
    for (int k = 0; k < 6; k++) { s.cvel[0][k] = 0; s.cacc[0][k] = 0; s.cfrc[0][k] = 0; }
    for (int k = 0; k < 3; k++) s.cacc[0][3 + k] = -m.gravity[k];
    for (int k = 0; k < 10; k++) s.crb[0][k] = 0;
  }
  for (int p = c.tid; p < m.nprop; p += c.nt) {     // free bodies: xpos = qpos[0:3], xquat = normalize(qpos[3:7])
    T* q = s.qpos + m.ntd + 7 * p;
    normquat(q + 3); cp3(s.ppos[p], q); quat2mat(s.pmat[p], q + 3);
  }
  bsync(c);
  for (int l = 0; l < m.nlevel; l++) {
    for (int idx = m.level_start[l] + c.tid; idx < m.level_start[l + 1]; idx += c.nt) {
      const int b = m.level_body[idx], p = m.tb_parent[b], d = b - 1;
      T pos[3], quat[4], a[3];
      mulmv(pos, s.xmat[p], m.tb_pos[b]);
      for (int k = 0; k < 3; k++) pos[k] += s.xpos[p][k];
      mulquat(quat, s.xquat[p], m.tb_quat[b]);
      T mat[9]; quat2mat(mat, quat);
      mulmv(a, mat, m.jnt_pos[d]);
      for (int k = 0; k < 3; k++) s.xanchor[d][k] = pos[k] + a[k];
      mulmv(s.xaxis[d], mat, m.jnt_axis[d]);
      const T ang = s.qpos[d] - m.qpos0[d], sn = sin(T(0.5) * ang);
      T qloc[4] = {cos(T(0.5) * ang), sn * m.jnt_axis[d][0], sn * m.jnt_axis[d][1], sn * m.jnt_axis[d][2]};
      mulquat(quat, quat, qloc); normquat(quat);
      quat2mat(mat, quat);
      mulmv(a, mat, m.jnt_pos[d]);
      for (int k = 0; k < 3; k++) { pos[k] = s.xanchor[d][k] - a[k]; s.xpos[b][k] = pos[k]; }
      for (int k = 0; k < 4; k++) s.xquat[b][k] = quat[k];
      for (int k = 0; k < 9; k++) s.xmat[b][k] = mat[k];
      // inertia about the fixed reference point, motion axis
      T xipos[3], qi[4], ximat[9], off[3];
      mulmv(xipos, mat, m.tb_ipos[b]);
      for (int k = 0; k < 3; k++) off[k] = pos[k] + xipos[k] - m.ref[k];
      mulquat(qi, quat, m.tb_iquat[b]); quat2mat(ximat, qi);
      inert_com(s.cinert[b], m.tb_inertia[b], ximat, off, m.tb_mass[b]);
      for (int k = 0; k < 10; k++) s.crb[b][k] = s.cinert[b][k];
      T* cd = s.cdof[d];
      cp3(cd, s.xaxis[d]);
      for (int k = 0; k < 3; k++) off[k] = m.ref[k] - s.xanchor[d][k];
      cross3(cd + 3, s.xaxis[d], off);
      // velocity: cdof_dot = cvel_parent x cdof ; cvel = cvel_parent + cdof*qvel
      T v[6];
      for (int k = 0; k < 6; k++) v[k] = s.cvel[p][k];
      cross_motion(s.cdof_dot[d], v, cd);
      for (int k = 0; k < 6; k++) { v[k] += cd[k] * s.qvel[d]; s.cvel[b][k] = v[k]; }
      // rne forward: cacc = cacc_parent + cdof_dot*qvel ; cfrc = I*cacc + cvel x* (I*cvel)
      T acc[6], Ia[6], Iv[6], cf[6];
      for (int k = 0; k < 6; k++) { acc[k] = s.cacc[p][k] + s.cdof_dot[d][k] * s.qvel[d]; s.cacc[b][k] = acc[k]; }
      mul_inert_vec(Ia, s.cinert[b], acc); mul_inert_vec(Iv, s.cinert[b], v); cross_force(cf, v, Iv);
      for (int k = 0; k < 6; k++) s.cfrc[b][k] = Ia[k] + cf[k];
    }
    bsync(c);
  }
  // geoms + sites
  for (int g = c.tid; g < m.ngeom; g += c.nt) {
    const int b = m.geom_body[g];
    const T *bp, *bm; T bq[4];
    if (b >= PROP_BODY0) { const int p = b - PROP_BODY0; bp = s.ppos[p]; bm = s.pmat[p]; for (int k = 0; k < 4; k++) bq[k] = s.qpos[m.ntd + 7 * p + 3 + k]; }
    else { bp = s.xpos[b]; bm = s.xmat[b]; for (int k = 0; k < 4; k++) bq[k] = s.xquat[b][k]; }
    T t[3], q[4];
    mulmv(t, bm, m.geom_pos[g]);
    for (int k = 0; k < 3; k++) s.gpos[g][k] = bp[k] + t[k];
    mulquat(q, bq, m.geom_quat[g]); quat2mat(s.gmat[g], q);
  }
  for (int i = c.tid; i < 2; i += c.nt) {
    const int b = m.site_body[i]; T t[3], q[4];
    mulmv(t, s.xmat[b], m.site_pos[i]);
    for (int k = 0; k < 3; k++) s.site_xpos[i][k] = s.xpos[b][k] + t[k];
    mulquat(q, s.xquat[b], m.site_quat[i]); quat2mat(s.site_xmat[i], q);
  }
  bsync(c);
}

// backward sweep: composite inertias (mj_crb) and rne force accumulation, parent gathers children
template <typename T>
B2P_HD void backward_sweep(const DevModel<T>& m, Scratch<T>& s, const Ctx& c) {
  for (int l = m.nlevel - 2; l >= 0; l--) {
    for (int idx = m.level_start[l] + c.tid; idx < m.level_start[l + 1]; idx += c.nt) {
      const int b = m.level_body[idx];
      for (int q = m.child_start[b]; q < m.child_start[b + 1]; q++) {
        const int ch = m.child_list[q];
        for (int k = 0; k < 10; k++) s.crb[b][k] += s.crb[ch][k];
        for (int k = 0; k < 6; k++) s.cfrc[b][k] += s.cfrc[ch][k];
      }
    }
    bsync(c);
  }
}

// dense Cholesky (lower, in place) of an n x n matrix with leading dimension ld; block-parallel
template <typename T>
B2P_HD void block_cholesky(T* A, int n, int ld, const Ctx& c) {
  for (int k = 0; k < n; k++) {
    T d = A[k * ld + k];
    d = d > Num<T>::minval() ? sqrt(d) : sqrt(Num<T>::minval());
    bsync(c);   // everyone has read the pivot before it is overwritten
    for (int i = k + c.tid; i < n; i += c.nt) A[i * ld + k] = (i == k) ? d : A[i * ld + k] / d;
    bsync(c);
    const int rem = n - k - 1, cnt = rem * (rem + 1) / 2;
    for (int e = c.tid; e < cnt; e += c.nt) {      // e -> (i >= j) in the trailing block
      int i = (int)((sqrt(T(8) * e + 1) - 1) * T(0.5));
      while ((i + 1) * (i + 2) / 2 <= e) i++;
      while (i * (i + 1) / 2 > e) i--;
      const int j = e - i * (i + 1) / 2, ii = k + 1 + i, jj = k + 1 + j;
      A[ii * ld + jj] -= A[ii * ld + k] * A[jj * ld + k];
    }
    bsync(c);
  }
}
// x <- (L L')^-1 x  by the first warp (others idle); caller syncs afterwards
template <typename T>
B2P_HD void warp_chol_solve(const T* L, int n, int ld, T* x, const Ctx& c) {
#ifdef __CUDA_ARCH__
  if (c.tid < 32) {
    const int lane = c.tid;
    for (int k = 0; k < n; k++) {
      const T xk = x[k] / L[k * ld + k];
      __syncwarp();
      if (lane == 0) x[k] = xk;
      for (int i = k + 1 + lane; i < n; i += 32) x[i] -= L[i * ld + k] * xk;
      __syncwarp();
    }
    for (int k = n - 1; k >= 0; k--) {
      const T xk = x[k] / L[k * ld + k];
      __syncwarp();
      if (lane == 0) x[k] = xk;
      for (int i = lane; i < k; i += 32) x[i] -= L[k * ld + i] * xk;
      __syncwarp();
    }
  }
#else
  (void)c;
  for (int k = 0; k < n; k++) { const T xk = x[k] / L[k * ld + k]; x[k] = xk; for (int i = k + 1; i < n; i++) x[i] -= L[i * ld + k] * xk; }
  for (int k = n - 1; k >= 0; k--) { const T xk = x[k] / L[k * ld + k]; x[k] = xk; for (int i = 0; i < k; i++) x[i] -= L[k * ld + i] * xk; }
#endif
}

// mass matrix (tree block dense, props diagonal), bias forces, passive forces
template <typename T>
B2P_HD void dynamics_terms(const DevModel<T>& m, Scratch<T>& s, const Ctx& c) {
  const int n = m.ntd;
  for (int e = c.tid; e < n * n; e += c.nt) {
    const int i = e / n, j = e % n;
    if (j > i) continue;
    T v = 0;
    if (m.anc_mask[i + 1] & (1u << j)) {
      T buf[6]; mul_inert_vec(buf, s.crb[i + 1], s.cdof[i]);
      for (int k = 0; k < 6; k++) v += s.cdof[j][k] * buf[k];
      if (i == j) v += m.dof_armature[i];
    }
    s.Marm[i * MAXTD + j] = v; s.Marm[j * MAXTD + i] = v;
    s.Larm[i * MAXTD + j] = v; s.Larm[j * MAXTD + i] = v;
  }
  for (int d = c.tid; d < m.nv; d += c.nt) {
    if (d < n) {
      T v = 0;
      for (int k = 0; k < 6; k++) v += s.cdof[d][k] * s.cfrc[d + 1][k];
      s.qfrc_bias[d] = v;
      s.qfrc_passive[d] = -m.dof_damping[d] * s.qvel[d] - m.jnt_stiffness[d] * (s.qpos[d] - m.qpos_spring[d]);
    } else {
      const int p = (d - n) / 6, k = (d - n) % 6; const T mass = m.prop_mass[p];
      if (k < 3) { s.Mdiag[d] = mass; s.qfrc_bias[d] = -mass * m.gravity[k]; }
      else {
        const T* w = s.qvel + n + 6 * p + 3; const T* I = s.pinertia[p];
        T Iw[3] = {I[0] * w[0], I[1] * w[1], I[2] * w[2]}, cr[3];
        cross3(cr, w, Iw);
        s.Mdiag[d] = I[k - 3]; s.qfrc_bias[d] = cr[k - 3];
      }
      s.qfrc_passive[d] = 0;
    }
  }
  bsync(c);
  block_cholesky(s.Larm, n, MAXTD, c);
}

// ------------------------------------------------------------------------------------------ collision
template <typename T>
struct CObj { const T* pos; const T* mat; const T* size; const T* vert; int type, vadr, vnum, cur; T margin; };

// support point in world direction dir (|dir| = 1), inflated by margin/2.  Meshes: hill climbing over the
// hull graph with warm start inside one query (what MuJoCo does for meshes with a graph).
template <typename T>
B2P_HD void support(CObj<T>& o, const Globals<T>& g, const T* dir, T* res) {
  T ld[3], lp[3];
  mulmtv(ld, o.mat, dir);
  if (o.type == B2P_GEOM_BOX) {
    for (int k = 0; k < 3; k++) lp[k] = ld[k] >= 0 ? o.size[k] : -o.size[k];
  } else {
    int v = o.cur; T best = dot3(g.mesh_vert + 3 * v, ld);
    for (int guard = 0; guard < 4096; guard++) {
      int nb = v; const int a = g.nbradr[v], n = g.nbrnum[v];
      for (int q = 0; q < n; q++) { const int u = g.nbr[a + q]; const T d = dot3(g.mesh_vert + 3 * u, ld); if (d > best) { best = d; nb = u; } }
      if (nb == v) break;
      v = nb;
    }
    o.cur = v; cp3(lp, g.mesh_vert + 3 * v);
  }
  mulmv(res, o.mat, lp);
  const T hm = T(0.5) * o.margin;
  for (int k = 0; k < 3; k++) res[k] += o.pos[k] + hm * dir[k];
}
template <typename T> struct Sup { T v[3], v1[3], v2[3]; };
template <typename T>
B2P_HD void mink_support(CObj<T>& a, CObj<T>& b, const Globals<T>& g, const T* dir, Sup<T>& s) {
  T nd[3] = {-dir[0], -dir[1], -dir[2]};
  support(a, g, dir, s.v1); support(b, g, nd, s.v2);
  for (int k = 0; k < 3; k++) s.v[k] = s.v1[k] - s.v2[k];
}
template <typename T> B2P_HD bool is_zero(T x) { return fabs(x) < Num<T>::eps(); }
template <typename T> B2P_HD void portal_dir(const Sup<T>* p, T* dir) {
  T a[3], b[3];
  for (int k = 0; k < 3; k++) { a[k] = p[2].v[k] - p[1].v[k]; b[k] = p[3].v[k] - p[1].v[k]; }
  cross3(dir, a, b); normalize3(dir);
}
template <typename T> B2P_HD bool portal_reach_tol(const Sup<T>* p, const Sup<T>& v4, const T* dir, T tol) {
  T dv4 = dot3(v4.v, dir), mn = dv4 - dot3(p[1].v, dir), d2 = dv4 - dot3(p[2].v, dir), d3 = dv4 - dot3(p[3].v, dir);
  if (d2 < mn) mn = d2;
  if (d3 < mn) mn = d3;
  return is_zero(mn - tol) || mn < tol;
}
template <typename T> B2P_HD void expand_portal(Sup<T>* p, const Sup<T>& v4) {
  T cr[3]; cross3(cr, v4.v, p[0].v);
  if (dot3(p[1].v, cr) > 0) { if (dot3(p[2].v, cr) > 0) p[1] = v4; else p[3] = v4; }
  else { if (dot3(p[3].v, cr) > 0) p[2] = v4; else p[1] = v4; }
}
template <typename T> B2P_HD T point_tri_dist2(const T* a, const T* b, const T* c, T* w) {
  T ab[3], ac[3], ap[3], bp[3], cpt[3];
  for (int k = 0; k < 3; k++) { ab[k] = b[k] - a[k]; ac[k] = c[k] - a[k]; ap[k] = -a[k]; bp[k] = -b[k]; cpt[k] = -c[k]; }
  T d1 = dot3(ab, ap), d2 = dot3(ac, ap), d3 = dot3(ab, bp), d4 = dot3(ac, bp), d5 = dot3(ab, cpt), d6 = dot3(ac, cpt);
  if (d1 <= 0 && d2 <= 0) { cp3(w, a); }
  else if (d3 >= 0 && d4 <= d3) { cp3(w, b); }
  else if (d1 * d4 - d3 * d2 <= 0 && d1 >= 0 && d3 <= 0) { T v = d1 / (d1 - d3); for (int k = 0; k < 3; k++) w[k] = a[k] + v * ab[k]; }
  else if (d6 >= 0 && d5 <= d6) { cp3(w, c); }
  else if (d5 * d2 - d1 * d6 <= 0 && d2 >= 0 && d6 <= 0) { T v = d2 / (d2 - d6); for (int k = 0; k < 3; k++) w[k] = a[k] + v * ac[k]; }
  else if (d3 * d6 - d5 * d4 <= 0 && (d4 - d3) >= 0 && (d5 - d6) >= 0) { T v = (d4 - d3) / ((d4 - d3) + (d5 - d6)); for (int k = 0; k < 3; k++) w[k] = b[k] + v * (c[k] - b[k]); }
  else {
    T va = d3 * d6 - d5 * d4, vb = d5 * d2 - d1 * d6, vc = d1 * d4 - d3 * d2, den = 1 / (va + vb + vc);
    T v = vb * den, u = vc * den;
    for (int k = 0; k < 3; k++) w[k] = a[k] + ab[k] * v + ac[k] * u;
  }
  return dot3(w, w);
}
// Minkowski portal refinement (libccd ccdMPRPenetration, what mjc_Convex runs in MuJoCo 3.2.x)
template <typename T>
B2P_HDN bool mpr_penetration(CObj<T>& o1, CObj<T>& o2, const Globals<T>& g, int maxiter, T tol, T& depth, T* dir, T* pos) {
  Sup<T> p[4], v4; T d[3], va[3], vb[3];
  for (int k = 0; k < 3; k++) { p[0].v1[k] = o1.pos[k]; p[0].v2[k] = o2.pos[k]; p[0].v[k] = p[0].v1[k] - p[0].v2[k]; }
  if (is_zero(p[0].v[0]) && is_zero(p[0].v[1]) && is_zero(p[0].v[2])) p[0].v[0] += T(1e-5);
  for (int k = 0; k < 3; k++) d[k] = -p[0].v[k];
  normalize3(d);
  mink_support(o1, o2, g, d, p[1]);
  T dt = dot3(p[1].v, d);
  if (is_zero(dt) || dt < 0) return false;
  cross3(d, p[0].v, p[1].v);
  if (is_zero(dot3(d, d))) {
    if (is_zero(dot3(p[1].v, p[1].v))) { depth = 0; dir[0] = dir[1] = dir[2] = 0; }
    else { cp3(dir, p[1].v); depth = normalize3(dir); }
    for (int k = 0; k < 3; k++) pos[k] = T(0.5) * (p[1].v1[k] + p[1].v2[k]);
    return true;
  }
  normalize3(d);
  mink_support(o1, o2, g, d, p[2]);
  dt = dot3(p[2].v, d);
  if (is_zero(dt) || dt < 0) return false;
  for (int k = 0; k < 3; k++) { va[k] = p[1].v[k] - p[0].v[k]; vb[k] = p[2].v[k] - p[0].v[k]; }
  cross3(d, va, vb); normalize3(d);
  if (dot3(d, p[0].v) > 0) { Sup<T> t = p[1]; p[1] = p[2]; p[2] = t; for (int k = 0; k < 3; k++) d[k] = -d[k]; }
  for (int guard = 0;; guard++) {
    if (guard > 100) return false;
    mink_support(o1, o2, g, d, p[3]);
    dt = dot3(p[3].v, d);
    if (is_zero(dt) || dt < 0) return false;
    bool cont = false;
    cross3(va, p[1].v, p[3].v);
    dt = dot3(va, p[0].v);
    if (dt < 0 && !is_zero(dt)) { p[2] = p[3]; cont = true; }
    if (!cont) {
      cross3(va, p[3].v, p[2].v);
      dt = dot3(va, p[0].v);
      if (dt < 0 && !is_zero(dt)) { p[1] = p[3]; cont = true; }
    }
    if (!cont) break;
    for (int k = 0; k < 3; k++) { va[k] = p[1].v[k] - p[0].v[k]; vb[k] = p[2].v[k] - p[0].v[k]; }
    cross3(d, va, vb); normalize3(d);
  }
  for (int guard = 0;; guard++) {
    portal_dir(p, d);
    dt = dot3(d, p[1].v);
    if (is_zero(dt) || dt > 0) break;
    mink_support(o1, o2, g, d, v4);
    dt = dot3(v4.v, d);
    if (!(is_zero(dt) || dt > 0) || portal_reach_tol(p, v4, d, tol) || guard > maxiter) return false;
    expand_portal(p, v4);
  }
  for (int it = 0;; it++) {
    portal_dir(p, d);
    mink_support(o1, o2, g, d, v4);
    if (portal_reach_tol(p, v4, d, tol) || it > maxiter) {
      T pd[3];
      depth = sqrt(point_tri_dist2(p[1].v, p[2].v, p[3].v, pd));
      if (is_zero(pd[0]) && is_zero(pd[1]) && is_zero(pd[2])) cp3(pd, d);
      normalize3(pd); cp3(dir, pd);
      T b[4], cr[3], sum;
      cross3(cr, p[1].v, p[2].v); b[0] = dot3(cr, p[3].v);
      cross3(cr, p[3].v, p[2].v); b[1] = dot3(cr, p[0].v);
      cross3(cr, p[0].v, p[1].v); b[2] = dot3(cr, p[3].v);
      cross3(cr, p[2].v, p[1].v); b[3] = dot3(cr, p[0].v);
      sum = b[0] + b[1] + b[2] + b[3];
      if (is_zero(sum) || sum < 0) {
        b[0] = 0;
        cross3(cr, p[2].v, p[3].v); b[1] = dot3(cr, d);
        cross3(cr, p[3].v, p[1].v); b[2] = dot3(cr, d);
        cross3(cr, p[1].v, p[2].v); b[3] = dot3(cr, d);
        sum = b[1] + b[2] + b[3];
      }
      for (int k = 0; k < 3; k++) {
        T p1 = 0, p2 = 0;
        for (int i = 0; i < 4; i++) { p1 += b[i] * p[i].v1[k]; p2 += b[i] * p[i].v2[k]; }
        pos[k] = T(0.5) * (p1 + p2) / sum;
      }
      return true;
    }
    expand_portal(p, v4);
  }
}

template <typename T> struct RawCon { T dist[8]; T pos[8][3]; T normal[3]; };

template <typename T>
B2P_HD int plane_box(const T* ppos, const T* pmat, const T* bpos, const T* bmat, const T* bsize, T margin, RawCon<T>& out) {
  T n[3] = {pmat[2], pmat[5], pmat[8]};
  int cnt = 0;
  cp3(out.normal, n);
  for (int i = 0; i < 8 && cnt < 4; i++) {
    T loc[3] = {(i & 1 ? bsize[0] : -bsize[0]), (i & 2 ? bsize[1] : -bsize[1]), (i & 4 ? bsize[2] : -bsize[2])}, w[3], dif[3];
    mulmv(w, bmat, loc);
    for (int k = 0; k < 3; k++) { w[k] += bpos[k]; dif[k] = w[k] - ppos[k]; }
    const T dist = dot3(dif, n);
    if (dist > margin) continue;
    out.dist[cnt] = dist;
    for (int k = 0; k < 3; k++) out.pos[cnt][k] = w[k] - T(0.5) * dist * n[k];
    cnt++;
  }
  return cnt;
}

// SAT (15 axes) + reference-face clipping / edge-edge point.  Returns #contacts, or -1 when the boxes are
// separated by 0 < s <= margin (caller then runs the margin-inflated convex query: one closest-feature contact).
template <typename T>
B2P_HDN int box_box(const T* p1, const T* m1, const T* s1, const T* p2, const T* m2, const T* s2, T margin, RawCon<T>& out) {
  T R[9], aR[9], t[3], dp[3];
  for (int k = 0; k < 3; k++) dp[k] = p2[k] - p1[k];
  mulmtv(t, m1, dp);
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) {
    R[3 * i + j] = m1[i] * m2[j] + m1[3 + i] * m2[3 + j] + m1[6 + i] * m2[6 + j];
    aR[3 * i + j] = fabs(R[3 * i + j]) + T(1e-12);
  }
  T best = -Num<T>::big(); int code = -1; T bn[3] = {0, 0, 0};
  for (int i = 0; i < 3; i++) {
    T s = fabs(t[i]) - (s1[i] + s2[0] * aR[3 * i] + s2[1] * aR[3 * i + 1] + s2[2] * aR[3 * i + 2]);
    if (s > margin) return 0;
    if (s > best) { best = s; code = i; }
  }
  for (int j = 0; j < 3; j++) {
    T tj = t[0] * R[j] + t[1] * R[3 + j] + t[2] * R[6 + j];
    T s = fabs(tj) - (s2[j] + s1[0] * aR[j] + s1[1] * aR[3 + j] + s1[2] * aR[6 + j]);
    if (s > margin) return 0;
    if (s > best) { best = s; code = 3 + j; }
  }
  T best_edge = -Num<T>::big(); int ecode = -1;
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) {
    const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
    const T l = sqrt(R[3 * i1 + j] * R[3 * i1 + j] + R[3 * i2 + j] * R[3 * i2 + j]);
    if (l < T(1e-6)) continue;
    const T tl = t[i2] * R[3 * i1 + j] - t[i1] * R[3 * i2 + j];
    const T ra = s1[i1] * aR[3 * i2 + j] + s1[i2] * aR[3 * i1 + j], rb = s2[j1] * aR[3 * i + j2] + s2[j2] * aR[3 * i + j1];
    const T s = (fabs(tl) - (ra + rb)) / l;
    if (s > margin) return 0;
    if (s > best_edge) {
      best_edge = s; ecode = 6 + 3 * i + j;
      T a[3] = {m1[i], m1[3 + i], m1[6 + i]}, b[3] = {m2[j], m2[3 + j], m2[6 + j]};
      cross3(bn, a, b); normalize3(bn);
      if (dot3(bn, dp) < 0) for (int k = 0; k < 3; k++) bn[k] = -bn[k];
    }
  }
  if (ecode >= 0 && best_edge > best + T(1e-6) * (1 + fabs(best))) { best = best_edge; code = ecode; }
  if (best > 0) return -1;
  if (code >= 6) {
    const int i = (code - 6) / 3, j = (code - 6) % 3;
    T pa[3], pb[3], ua[3] = {m1[i], m1[3 + i], m1[6 + i]}, ub[3] = {m2[j], m2[3 + j], m2[6 + j]};
    cp3(pa, p1); cp3(pb, p2);
    for (int k = 0; k < 3; k++) {
      if (k != i) { T ax[3] = {m1[k], m1[3 + k], m1[6 + k]}; T sg = dot3(bn, ax) > 0 ? T(1) : T(-1); for (int q = 0; q < 3; q++) pa[q] += sg * s1[k] * ax[q]; }
      if (k != j) { T ax[3] = {m2[k], m2[3 + k], m2[6 + k]}; T sg = dot3(bn, ax) > 0 ? T(-1) : T(1); for (int q = 0; q < 3; q++) pb[q] += sg * s2[k] * ax[q]; }
    }
    T w[3] = {pb[0] - pa[0], pb[1] - pa[1], pb[2] - pa[2]};
    T uaub = dot3(ua, ub), q1 = dot3(ua, w), q2 = -dot3(ub, w), den = 1 - uaub * uaub;
    T al = den > T(1e-12) ? (q1 + uaub * q2) / den : T(0), be = den > T(1e-12) ? (uaub * q1 + q2) / den : T(0);
    for (int k = 0; k < 3; k++) { pa[k] += al * ua[k]; pb[k] += be * ub[k]; }
    out.dist[0] = best; cp3(out.normal, bn);
    for (int k = 0; k < 3; k++) out.pos[0][k] = T(0.5) * (pa[k] + pb[k]);
    return 1;
  }
  const T *pa = p1, *ma = m1, *sa = s1, *pb = p2, *mb = m2, *sb = s2; int ai = code; bool flip = false;
  if (code >= 3) { pa = p2; ma = m2; sa = s2; pb = p1; mb = m1; sb = s1; ai = code - 3; flip = true; }
  T nref[3] = {ma[ai], ma[3 + ai], ma[6 + ai]}, d[3] = {pb[0] - pa[0], pb[1] - pa[1], pb[2] - pa[2]};
  if (dot3(nref, d) < 0) for (int k = 0; k < 3; k++) nref[k] = -nref[k];
  int bj = 0; T bd = -1;
  for (int j = 0; j < 3; j++) { T ax[3] = {mb[j], mb[3 + j], mb[6 + j]}; T v = fabs(dot3(nref, ax)); if (v > bd) { bd = v; bj = j; } }
  T nb[3] = {mb[bj], mb[3 + bj], mb[6 + bj]};
  if (dot3(nb, nref) > 0) for (int k = 0; k < 3; k++) nb[k] = -nb[k];
  const int k1 = (bj + 1) % 3, k2 = (bj + 2) % 3;
  T poly[16][3], tmp[16][3]; int np = 4;
  for (int q = 0; q < 4; q++) {
    const T sg1 = (q == 0 || q == 3) ? T(1) : T(-1), sg2 = (q < 2) ? T(1) : T(-1);
    for (int k = 0; k < 3; k++) poly[q][k] = pb[k] + nb[k] * sb[bj] + sg1 * sb[k1] * mb[3 * k + k1] + sg2 * sb[k2] * mb[3 * k + k2];
  }
  const int u1 = (ai + 1) % 3, u2 = (ai + 2) % 3;
  for (int pl = 0; pl < 4; pl++) {
    const int u = pl < 2 ? u1 : u2; const T sg = (pl & 1) ? T(-1) : T(1);
    T ax[3] = {sg * ma[u], sg * ma[3 + u], sg * ma[6 + u]}; const T lim = sa[u];
    int nn = 0;
    for (int v = 0; v < np; v++) {
      const T *A = poly[v], *Bp = poly[(v + 1) % np];
      const T da = (A[0] - pa[0]) * ax[0] + (A[1] - pa[1]) * ax[1] + (A[2] - pa[2]) * ax[2] - lim;
      const T db = (Bp[0] - pa[0]) * ax[0] + (Bp[1] - pa[1]) * ax[1] + (Bp[2] - pa[2]) * ax[2] - lim;
      if (da <= 0) { cp3(tmp[nn], A); nn++; }
      if ((da < 0 && db > 0) || (da > 0 && db < 0)) { const T sx = da / (da - db); for (int k = 0; k < 3; k++) tmp[nn][k] = A[k] + sx * (Bp[k] - A[k]); nn++; }
    }
    np = nn;
    for (int v = 0; v < np; v++) cp3(poly[v], tmp[v]);
    if (np == 0) break;
  }
  int cnt = 0;
  for (int k = 0; k < 3; k++) out.normal[k] = flip ? -nref[k] : nref[k];
  for (int v = 0; v < np && cnt < 8; v++) {
    const T dist = (poly[v][0] - pa[0]) * nref[0] + (poly[v][1] - pa[1]) * nref[1] + (poly[v][2] - pa[2]) * nref[2] - sa[ai];
    if (dist > margin) continue;
    out.dist[cnt] = dist;
    for (int k = 0; k < 3; k++) out.pos[cnt][k] = poly[v][k] - T(0.5) * dist * nref[k];
    cnt++;
  }
  return cnt;
}

// broad phase (bounding spheres over the static pair table) + narrow phase, deterministic contact order
template <typename T>
B2P_HD void collide(const DevModel<T>& m, const Globals<T>& g, Scratch<T>& s, const Ctx& c, bool full_contacts) {
  if (c.tid == 0) { s.ncand = 0; s.ncon = 0; }
  for (int q = c.tid; q < m.ngeom; q += c.nt) {
    const int p = m.geom_prop[q];
    s.grbound[q] = p >= 0 ? norm3(s.psize[p]) : m.geom_rbound[q];
  }
  bsync(c);
  int base = 0;
  for (int p0 = 0; p0 < m.npair; p0 += c.nt) {
    const int p = p0 + c.tid; int flag = 0;
    if (p < m.npair) {
      const int g1 = g.pair_g1[p], g2 = g.pair_g2[p], pr1 = m.geom_prop[g1], pr2 = m.geom_prop[g2];
      if (!((pr1 >= 0 && !s.pactive[pr1]) || (pr2 >= 0 && !s.pactive[pr2]))) {
        const PairParam<T>& pp = g.pair_param[p];
        const T md = full_contacts ? pp.margin : pp.margin - pp.gap;
        T dif[3] = {s.gpos[g2][0] - s.gpos[g1][0], s.gpos[g2][1] - s.gpos[g1][1], s.gpos[g2][2] - s.gpos[g1][2]};
        if (m.geom_type[g1] == B2P_GEOM_PLANE) {
          T nrm[3] = {s.gmat[g1][2], s.gmat[g1][5], s.gmat[g1][8]};
          flag = (m.geom_type[g2] == B2P_GEOM_BOX) && !(dot3(dif, nrm) - s.grbound[g2] > md);
        } else flag = !(norm3(dif) > s.grbound[g1] + s.grbound[g2] + md);
      }
    }
    int tot; const int ofs = block_scan(c, flag, tot, s.iscan);
    if (flag && base + ofs < MAXCAND) s.cand[base + ofs] = (short)p;
    base += tot;
  }
  if (c.tid == 0) { s.ncand = base < MAXCAND ? base : MAXCAND; if (base > MAXCAND) s.flags |= FLAG_CANDFULL; }
  bsync(c);
  const int ncand = s.ncand;
  const T tol = m.ccd_tolerance > 16 * Num<T>::eps() ? m.ccd_tolerance : 16 * Num<T>::eps();
  int cbase = 0;
  for (int c0 = 0; c0 < ncand; c0 += c.nt) {
    const int ci = c0 + c.tid; int n = 0; RawCon<T> raw; int p = -1; T md = 0;
    if (ci < ncand) {
      p = s.cand[ci];
      const int g1 = g.pair_g1[p], g2 = g.pair_g2[p], t1 = m.geom_type[g1], t2 = m.geom_type[g2];
      const PairParam<T>& pp = g.pair_param[p];
      md = full_contacts ? pp.margin : pp.margin - pp.gap;
      const int pr1 = m.geom_prop[g1], pr2 = m.geom_prop[g2];
      const T* sz1 = pr1 >= 0 ? s.psize[pr1] : m.geom_size[g1];
      const T* sz2 = pr2 >= 0 ? s.psize[pr2] : m.geom_size[g2];
      if (t1 == B2P_GEOM_PLANE) n = plane_box(s.gpos[g1], s.gmat[g1], s.gpos[g2], s.gmat[g2], sz2, md, raw);
      else {
        bool convex = true;
        if (t1 == B2P_GEOM_BOX && t2 == B2P_GEOM_BOX) { n = box_box(s.gpos[g1], s.gmat[g1], sz1, s.gpos[g2], s.gmat[g2], sz2, md, raw); convex = n < 0; if (n < 0) n = 0; }
        if (convex) {
          CObj<T> a = {s.gpos[g1], s.gmat[g1], sz1, nullptr, t1, m.geom_vertadr[g1], m.geom_vertnum[g1], m.geom_vertadr[g1], md};
          CObj<T> b = {s.gpos[g2], s.gmat[g2], sz2, nullptr, t2, m.geom_vertadr[g2], m.geom_vertnum[g2], m.geom_vertadr[g2], md};
          T depth, dir[3], pos[3];
          if (mpr_penetration(a, b, g, m.ccd_iterations, tol, depth, dir, pos) && depth >= 0) {
            raw.dist[0] = md - depth; cp3(raw.normal, dir); cp3(raw.pos[0], pos); n = 1;
          }
        }
      }
      int keep = 0;   // drop points not strictly inside the detection margin
      for (int i = 0; i < n; i++) if (raw.dist[i] < md) { if (keep != i) { raw.dist[keep] = raw.dist[i]; cp3(raw.pos[keep], raw.pos[i]); } keep++; }
      n = keep;
    }
    int tot; const int ofs = block_scan(c, n, tot, s.iscan);
    for (int i = 0; i < n; i++) {
      const int k = cbase + ofs + i;
      if (k >= MAXCON) break;
      const PairParam<T>& pp = g.pair_param[p];
      s.c_dist[k] = raw.dist[i]; cp3(s.c_pos[k], raw.pos[i]); cp3(s.c_frame[k], raw.normal); make_frame(s.c_frame[k]);
      s.c_pair[k] = (short)p; s.c_incl[k] = pp.margin - pp.gap; s.c_fric[k][0] = pp.friction[0]; s.c_fric[k][1] = pp.friction[0]; s.c_row[k] = -1;
    }
    cbase += tot;
  }
  if (c.tid == 0) { s.ncon = cbase < MAXCON ? cbase : MAXCON; if (cbase > MAXCON) s.flags |= FLAG_CONTACTFULL; }
  bsync(c);
}

// ------------------------------------------------------------------------------------------ constraints
template <typename T> B2P_HD T clampT(T x, T lo, T hi) { return x < lo ? lo : (x > hi ? hi : x); }
template <typename T>
B2P_HD T impedance(const T* si, T pos, T margin) {
  const T d0 = clampT(si[0], T(0.0001), T(0.9999)), d1 = clampT(si[1], T(0.0001), T(0.9999)), w = si[2] > 0 ? si[2] : T(0);
  const T mid = clampT(si[3], T(0.0001), T(0.9999)), pw = si[4] > 1 ? si[4] : T(1);
  if (d0 == d1 || w <= Num<T>::minval()) return T(0.5) * (d0 + d1);
  const T x = fabs((pos - margin) / w); T y;
  if (x >= 1) return d1;
  if (x <= 0) return d0;
  if (pw == 1) y = x;
  else if (x <= mid) y = pow(x, pw) / pow(mid, pw - 1);
  else y = 1 - pow(1 - x, pw) / pow(1 - mid, pw - 1);
  return d0 + y * (d1 - d0);
}
// K, B, imp, R for one row (mj_makeImpedance, App. B item 9)
template <typename T>
B2P_HD void row_params(const DevModel<T>& m, const T* solref_in, const T* solimp, T pos_imp, T margin, T diag, bool friction_row, T& K, T& B, T& imp, T& R) {
  T sr0 = solref_in[0], sr1 = solref_in[1];
  if (sr0 > 0 && sr0 < 2 * m.timestep) sr0 = 2 * m.timestep;
  imp = impedance(solimp, pos_imp, margin);
  const T dmax = clampT(solimp[1], T(0.0001), T(0.9999)), mv = Num<T>::minval();
  if (sr0 > 0) { T a = dmax * dmax * sr0 * sr0 * sr1 * sr1; K = 1 / (a > mv ? a : mv); T bb = dmax * sr0; B = 2 / (bb > mv ? bb : mv); }
  else { T a = dmax * dmax; K = -sr0 / (a > mv ? a : mv); B = -sr1 / (dmax > mv ? dmax : mv); }
  if (friction_row) K = 0;
  R = (1 - imp) / imp * diag; if (R < mv) R = mv;
}
// add sign * ax . d(point velocity of unified body `body` at world point pt)/d(qvel) into a compressed row
template <typename T>
B2P_HD void add_jac(const DevModel<T>& m, const Scratch<T>& s, T* row, int seg0, int body, const T* pt, const T* ax, T sign) {
  if (body == 0) return;
  if (body >= PROP_BODY0) {
    const int p = body - PROP_BODY0; T r[3] = {pt[0] - s.ppos[p][0], pt[1] - s.ppos[p][1], pt[2] - s.ppos[p][2]};
    for (int k = 0; k < 3; k++) row[seg0 + k] += sign * ax[k];
    for (int k = 0; k < 3; k++) { T a[3] = {s.pmat[p][k], s.pmat[p][3 + k], s.pmat[p][6 + k]}, cr[3]; cross3(cr, a, r); row[seg0 + 3 + k] += sign * dot3(ax, cr); }
  } else {
    T off[3] = {pt[0] - m.ref[0], pt[1] - m.ref[1], pt[2] - m.ref[2]};
    const unsigned mask = m.anc_mask[body];
    for (int d = 0; d < m.ntd; d++) if (mask & (1u << d)) {
      T cr[3]; cross3(cr, s.cdof[d], off);
      row[seg0 + d] += sign * (ax[0] * (s.cdof[d][3] + cr[0]) + ax[1] * (s.cdof[d][4] + cr[1]) + ax[2] * (s.cdof[d][5] + cr[2]));
    }
  }
}
template <typename T>
B2P_HD T row_vel(const DevModel<T>& m, const Scratch<T>& s, int r) {
  T v = 0; const int a = s.row_a[r], b = s.row_b[r];
  if (a >= 0) { const int st = block_start(m, a), n = a == 0 ? m.ntd : 6; for (int k = 0; k < n; k++) v += s.J[r][k] * s.qvel[st + k]; }
  if (b >= 0) { const int st = block_start(m, b); for (int k = 0; k < 6; k++) v += s.J[r][SEGB + k] * s.qvel[st + k]; }
  return v;
}
template <typename T>
B2P_HD T row_dot(const DevModel<T>& m, const Scratch<T>& s, int r, const T* x) {
  T v = 0; const int a = s.row_a[r], b = s.row_b[r];
  if (a >= 0) { const int st = block_start(m, a), n = a == 0 ? m.ntd : 6; for (int k = 0; k < n; k++) v += s.J[r][k] * x[st + k]; }
  if (b >= 0) { const int st = block_start(m, b); for (int k = 0; k < 6; k++) v += s.J[r][SEGB + k] * x[st + k]; }
  return v;
}

// mj_makeConstraint + mj_referenceConstraint: equality -> limits -> contacts, rows complete incl. aref
template <typename T>
B2P_HD void make_rows(const DevModel<T>& m, const Globals<T>& g, Scratch<T>& s, const Ctx& c) {
  const int nlim = 2 * m.ntd, nitem = m.neq + nlim + s.ncon;
  int base = 0;
  for (int i0 = 0; i0 < nitem; i0 += c.nt) {
    const int it = i0 + c.tid; int nr = 0, kind = -1, id = 0; T dist = 0;
    if (it < nitem) {
      if (it < m.neq) { kind = ROW_EQ; id = it; nr = m.eq_type[it] == B2P_EQ_CONNECT ? 3 : 1; }
      else if (it < m.neq + nlim) {
        const int j = (it - m.neq) >> 1, side = ((it - m.neq) & 1) ? 1 : -1;
        if (m.jnt_limited[j]) { dist = side * (m.jnt_range[j][(side + 1) / 2] - s.qpos[j]); if (dist < m.jnt_margin[j]) { kind = ROW_LIMIT; id = j * 2 + (side > 0); nr = 1; } }
      } else { id = it - m.neq - nlim; if (s.c_dist[id] < s.c_incl[id]) { kind = ROW_CONTACT; nr = 3; } }
    }
    int tot; const int ofs = block_scan(c, nr, tot, s.iscan);
    int r0 = base + ofs;
    if (nr > 0 && r0 + nr > MAXEFC) { nr = 0; if (kind == ROW_CONTACT) s.c_row[id] = -1; s.flags |= FLAG_EFCFULL; }
    if (nr > 0) {
      for (int q = 0; q < nr; q++) { for (int k = 0; k < JW; k++) s.J[r0 + q][k] = 0; s.row_type[r0 + q] = (signed char)kind; s.row_id[r0 + q] = (short)id; s.row_b[r0 + q] = -1; s.row_a[r0 + q] = 0; }
      if (kind == ROW_EQ && nr == 3) {
        const int b1 = m.eq_obj1[id], b2 = m.eq_obj2[id]; T g1[3], g2[3], res[3];
        mulmv(g1, s.xmat[b1], m.eq_data[id]); mulmv(g2, s.xmat[b2], m.eq_data[id] + 3);
        for (int k = 0; k < 3; k++) { g1[k] += s.xpos[b1][k]; g2[k] += s.xpos[b2][k]; res[k] = g1[k] - g2[k]; }
        const T diag = m.tb_invweight0[b1][0] + m.tb_invweight0[b2][0], nrm = norm3(res);
        for (int q = 0; q < 3; q++) {
          T ax[3] = {T(q == 0), T(q == 1), T(q == 2)};
          add_jac(m, s, s.J[r0 + q], 0, b1, g1, ax, T(1)); add_jac(m, s, s.J[r0 + q], 0, b2, g2, ax, T(-1));
          T K, B, imp, R; row_params(m, m.eq_solref[id], m.eq_solimp[id], nrm, T(0), diag, false, K, B, imp, R);
          s.D[r0 + q] = 1 / R; s.aref[r0 + q] = -B * row_vel(m, s, r0 + q) - K * imp * res[q];
        }
      } else if (kind == ROW_EQ) {
        const int j1 = m.eq_obj1[id], j2 = m.eq_obj2[id]; const T* cf = m.eq_data[id];
        const T x = s.qpos[j2] - m.qpos0[j2], pos = s.qpos[j1] - m.qpos0[j1] - (cf[0] + cf[1] * x);
        s.J[r0][j1] = 1; s.J[r0][j2] = -cf[1];
        T K, B, imp, R; row_params(m, m.eq_solref[id], m.eq_solimp[id], pos, T(0), m.dof_invweight0[j1] + m.dof_invweight0[j2], false, K, B, imp, R);
        s.D[r0] = 1 / R; s.aref[r0] = -B * row_vel(m, s, r0) - K * imp * pos;
      } else if (kind == ROW_LIMIT) {
        const int j = id >> 1, side = (id & 1) ? 1 : -1;
        s.J[r0][j] = T(-side);
        T K, B, imp, R; row_params(m, m.jnt_solref[j], m.jnt_solimp[j], dist, m.jnt_margin[j], m.dof_invweight0[j], false, K, B, imp, R);
        s.D[r0] = 1 / R; s.aref[r0] = -B * row_vel(m, s, r0) - K * imp * (dist - m.jnt_margin[j]);
      } else {
        const int p = s.c_pair[id], g1 = g.pair_g1[p], g2 = g.pair_g2[p], b1 = m.geom_body[g1], b2 = m.geom_body[g2];
        const PairParam<T>& pp = g.pair_param[p];
        // segment assignment: tree -> A; a prop -> B if A is taken (by the tree or by the other prop)
        int ba = -1, bb = -1, seg1 = 0, seg2 = 0;
        const bool t1 = b1 > 0 && b1 < PROP_BODY0, t2 = b2 > 0 && b2 < PROP_BODY0;
        if (t1 || t2) ba = 0;
        if (b1 >= PROP_BODY0) { if (ba < 0) { ba = 1 + b1 - PROP_BODY0; seg1 = 0; } else { bb = 1 + b1 - PROP_BODY0; seg1 = SEGB; } }
        if (b2 >= PROP_BODY0) { if (ba < 0) { ba = 1 + b2 - PROP_BODY0; seg2 = 0; } else { bb = 1 + b2 - PROP_BODY0; seg2 = SEGB; } }
        T iw1[2], iw2[2];
        for (int k = 0; k < 2; k++) { iw1[k] = b1 >= PROP_BODY0 ? s.pinvw[b1 - PROP_BODY0][k] : m.geom_invweight0[g1][k]; iw2[k] = b2 >= PROP_BODY0 ? s.pinvw[b2 - PROP_BODY0][k] : m.geom_invweight0[g2][k]; }
        const T tran = iw1[0] + iw2[0];
        T K, B, imp, R0;
        row_params(m, pp.solref, pp.solimp, s.c_dist[id], s.c_incl[id], tran, false, K, B, imp, R0);
        const T mv = Num<T>::minval(), R1 = R0 / (m.impratio > mv ? m.impratio : mv);
        s.c_mu[id] = s.c_fric[id][0] * sqrt(R1 / R0); s.c_row[id] = (short)r0;
        for (int q = 0; q < 3; q++) {
          s.row_a[r0 + q] = (signed char)ba; s.row_b[r0 + q] = (signed char)bb;
          const T* ax = s.c_frame[id] + 3 * q;
          add_jac(m, s, s.J[r0 + q], seg2, b2, s.c_pos[id], ax, T(1)); add_jac(m, s, s.J[r0 + q], seg1, b1, s.c_pos[id], ax, T(-1));
          const T vel = row_vel(m, s, r0 + q);
          if (q == 0) { s.D[r0] = 1 / R0; s.aref[r0] = -B * vel - K * imp * (s.c_dist[id] - s.c_incl[id]); }
          else { s.D[r0 + q] = 1 / R1; s.aref[r0 + q] = -B * vel; }
        }
      }
    }
    base += tot;
  }
  if (c.tid == 0) s.nefc = base < MAXEFC ? base : MAXEFC;
  bsync(c);
}

// ------------------------------------------------------------------------------------------ actuation + smooth acceleration
template <typename T>
B2P_HD void actuation_and_smooth(const DevModel<T>& m, Scratch<T>& s, const Ctx& c) {
  for (int d = c.tid; d < m.nv; d += c.nt) s.qfrc_act[d] = 0;
  bsync(c);
  if (c.tid == 0) {
    for (int a = 0; a < m.nu; a++) {
      T len = 0, vel = 0;
      for (int k = 0; k < m.act_trnnum[a]; k++) { const int d = m.act_trndof[a][k]; len += m.act_trncoef[a][k] * s.qpos[d]; vel += m.act_trncoef[a][k] * s.qvel[d]; }
      T u = s.ctrl[a];
      if (m.act_ctrllimited[a]) u = clampT(u, m.act_ctrlrange[a][0], m.act_ctrlrange[a][1]);
      T f = m.act_gain[a] * u + m.act_bias[a][0] + m.act_bias[a][1] * len + m.act_bias[a][2] * vel;
      if (m.act_forcelimited[a]) f = clampT(f, m.act_forcerange[a][0], m.act_forcerange[a][1]);
      s.act_force[a] = f;
      for (int k = 0; k < m.act_trnnum[a]; k++) s.qfrc_act[m.act_trndof[a][k]] += m.act_trncoef[a][k] * f;
    }
  }
  bsync(c);
  for (int d = c.tid; d < m.nv; d += c.nt) {
    bool active = d < m.ntd || s.pactive[(d - m.ntd) / 6];
    const T f = active ? s.qfrc_passive[d] - s.qfrc_bias[d] + s.qfrc_act[d] : T(0);
    s.qfrc_smooth[d] = f;
    s.qacc_smooth[d] = d < m.ntd ? f : f / s.Mdiag[d];
  }
  bsync(c);
  warp_chol_solve(s.Larm, m.ntd, MAXTD, s.qacc_smooth, c);
  bsync(c);
}

// ------------------------------------------------------------------------------------------ Newton solver
template <typename T>
B2P_HD void mul_M(const DevModel<T>& m, const Scratch<T>& s, T* r, const T* v, const Ctx& c) {
  for (int d = c.tid; d < m.nv; d += c.nt) {
    if (d < m.ntd) { T a = 0; for (int j = 0; j < m.ntd; j++) a += s.Marm[d * MAXTD + j] * v[j]; r[d] = a; }
    else r[d] = s.Mdiag[d] * v[d];
  }
}
// per-item (rows of eq/limit, whole contacts) force/state/cost from s.Jaref; returns this thread's cost part
template <typename T>
B2P_HD T constraint_update(const DevModel<T>& m, Scratch<T>& s, const Ctx& c, bool want_hess) {
  T cost = 0;
  for (int r = c.tid; r < s.nefc; r += c.nt) {
    const int tp = s.row_type[r];
    if (tp == ROW_CONTACT) continue;
    const T j = s.Jaref[r];
    if (tp == ROW_EQ || j < 0) { s.force[r] = -s.D[r] * j; s.row_state[r] = ST_QUADRATIC; cost += T(0.5) * s.D[r] * j * j; }
    else { s.force[r] = 0; s.row_state[r] = ST_SATISFIED; }
  }
  for (int k = c.tid; k < s.ncon; k += c.nt) {
    const int r = s.c_row[k];
    if (r < 0) continue;
    const T mu = s.c_mu[k], f1 = s.c_fric[k][0], f2 = s.c_fric[k][1];
    const T j0 = s.Jaref[r], j1 = s.Jaref[r + 1], j2 = s.Jaref[r + 2];
    const T N = j0 * mu, U1 = j1 * f1, U2 = j2 * f2, Tt = sqrt(U1 * U1 + U2 * U2);
    T* W = s.cW[k];
    if (N >= mu * Tt || (Tt <= 0 && N >= 0)) {
      for (int q = 0; q < 3; q++) { s.force[r + q] = 0; s.row_state[r + q] = ST_SATISFIED; }
      if (want_hess) for (int q = 0; q < 6; q++) W[q] = 0;
    } else if (mu * N + Tt <= 0 || (Tt <= 0 && N < 0)) {
      for (int q = 0; q < 3; q++) { const T jq = s.Jaref[r + q]; s.force[r + q] = -s.D[r + q] * jq; s.row_state[r + q] = ST_QUADRATIC; cost += T(0.5) * s.D[r + q] * jq * jq; }
      if (want_hess) { W[0] = s.D[r]; W[1] = s.D[r + 1]; W[2] = s.D[r + 2]; W[3] = W[4] = W[5] = 0; }
    } else {
      const T mv = Num<T>::minval(), dn = mu * mu * (1 + mu * mu), Dm = s.D[r] / (dn > mv ? dn : mv), NmT = N - mu * Tt;
      cost += T(0.5) * Dm * NmT * NmT;
      const T fn = -Dm * NmT * mu;
      s.force[r] = fn; s.force[r + 1] = -fn / Tt * U1 * f1; s.force[r + 2] = -fn / Tt * U2 * f2;
      for (int q = 0; q < 3; q++) s.row_state[r + q] = ST_CONE;
      if (want_hess) {
        const T iT = 1 / Tt, a = Dm * mu * mu * iT * iT, b = Dm * mu * NmT, iT3 = iT * iT * iT;
        const T h00 = Dm, h01 = -Dm * mu * U1 * iT, h02 = -Dm * mu * U2 * iT;
        const T h11 = a * U1 * U1 - b * (iT - U1 * U1 * iT3), h22 = a * U2 * U2 - b * (iT - U2 * U2 * iT3), h12 = a * U1 * U2 + b * U1 * U2 * iT3;
        W[0] = h00 * mu * mu; W[1] = h11 * f1 * f1; W[2] = h22 * f2 * f2; W[3] = h01 * mu * f1; W[4] = h02 * mu * f2; W[5] = h12 * f1 * f2;
      }
    }
  }
  return cost;
}
// qfrc_con = J' force
template <typename T>
B2P_HD void mul_JT_force(const DevModel<T>& m, Scratch<T>& s, const Ctx& c) {
  for (int d = c.tid; d < m.nv; d += c.nt) {
    T a = 0;
    for (int r = 0; r < s.nefc; r++) { const T f = s.force[r]; if (f == 0) continue; const int sl = row_slot(m, s, r, d); if (sl >= 0) a += s.J[r][sl] * f; }
    s.qfrc_con[d] = a;
  }
}
// H = M + J' W J (lower triangle, dense nv x nv with leading dimension MAXNV)
template <typename T>
B2P_HD void assemble_H(const DevModel<T>& m, Scratch<T>& s, const Ctx& c) {
  const int nv = m.nv, cnt = nv * (nv + 1) / 2;
  for (int e = c.tid; e < cnt; e += c.nt) {
    int i = (int)((sqrt(T(8) * e + 1) - 1) * T(0.5));
    while ((i + 1) * (i + 2) / 2 <= e) i++;
    while (i * (i + 1) / 2 > e) i--;
    const int j = e - i * (i + 1) / 2;
    const int bi = dof_block(m, i), bj = dof_block(m, j), oi = i - block_start(m, bi), oj = j - block_start(m, bj);
    T h = 0;
    if (bi == 0 && bj == 0) h = s.Marm[i * MAXTD + j]; else if (i == j) h = s.Mdiag[i];
    for (int r = 0; r < s.nefc; r++) {
      const int a = s.row_a[r], b = s.row_b[r];
      const int si = a == bi ? oi : (b == bi ? SEGB + oi : -1);
      if (si < 0) { if (s.row_type[r] == ROW_CONTACT) r += 2; continue; }
      const int sj = a == bj ? oj : (b == bj ? SEGB + oj : -1);
      if (sj < 0) { if (s.row_type[r] == ROW_CONTACT) r += 2; continue; }
      if (s.row_type[r] != ROW_CONTACT) { if (s.row_state[r] == ST_QUADRATIC) h += s.D[r] * s.J[r][si] * s.J[r][sj]; }
      else {
        const int st = s.row_state[r];
        if (st != ST_SATISFIED) {
          const T* W = s.cW[s.row_id[r]];
          const T a0 = s.J[r][si], a1 = s.J[r + 1][si], a2 = s.J[r + 2][si], b0 = s.J[r][sj], b1 = s.J[r + 1][sj], b2 = s.J[r + 2][sj];
          h += W[0] * a0 * b0 + W[1] * a1 * b1 + W[2] * a2 * b2 + W[3] * (a0 * b1 + a1 * b0) + W[4] * (a0 * b2 + a2 * b0) + W[5] * (a1 * b2 + a2 * b1);
        }
        r += 2;
      }
    }
    s.H[i * MAXNV + j] = h;
  }
}
template <typename T> struct LsPt { T alpha, cost, d1, d2; };
template <typename T>
B2P_HD LsPt<T> ls_eval(const DevModel<T>& m, Scratch<T>& s, const Ctx& c, T a, const T* qg) {
  T q0 = 0, q1 = 0, q2 = 0, ec = 0, e1 = 0, e2 = 0;
  for (int r = c.tid; r < s.nefc; r += c.nt) {
    const int tp = s.row_type[r];
    if (tp == ROW_CONTACT) continue;
    const T D = s.D[r], j0 = s.Jaref[r], jv = s.Jv[r];
    if (tp == ROW_EQ || j0 + a * jv < 0) { q0 += T(0.5) * D * j0 * j0; q1 += D * j0 * jv; q2 += T(0.5) * D * jv * jv; }
  }
  for (int k = c.tid; k < s.ncon; k += c.nt) {
    const int r = s.c_row[k];
    if (r < 0) continue;
    const T mu = s.c_mu[k], f1 = s.c_fric[k][0], f2 = s.c_fric[k][1];
    T t0 = 0, t1 = 0, t2 = 0;
    for (int q = 0; q < 3; q++) { const T D = s.D[r + q], j0 = s.Jaref[r + q], jv = s.Jv[r + q]; t0 += T(0.5) * D * j0 * j0; t1 += D * j0 * jv; t2 += T(0.5) * D * jv * jv; }
    const T U0 = s.Jaref[r] * mu, V0 = s.Jv[r] * mu, u1 = s.Jaref[r + 1] * f1, u2 = s.Jaref[r + 2] * f2, v1 = s.Jv[r + 1] * f1, v2 = s.Jv[r + 2] * f2;
    const T UU = u1 * u1 + u2 * u2, UV = u1 * v1 + u2 * v2, VV = v1 * v1 + v2 * v2;
    const T N = U0 + a * V0, Tsqr = UU + a * (2 * UV + a * VV);
    if (Tsqr <= 0) { if (N < 0) { q0 += t0; q1 += t1; q2 += t2; } }
    else {
      const T Tt = sqrt(Tsqr);
      if (N >= mu * Tt) { }
      else if (mu * N + Tt <= 0) { q0 += t0; q1 += t1; q2 += t2; }
      else {
        const T mv = Num<T>::minval(), dn = mu * mu * (1 + mu * mu), Dm = s.D[r] / (dn > mv ? dn : mv);
        const T N1 = V0, T1 = (UV + a * VV) / Tt, T2 = VV / Tt - (UV + a * VV) * T1 / (Tt * Tt), NmT = N - mu * Tt;
        ec += T(0.5) * Dm * NmT * NmT; e1 += Dm * NmT * (N1 - mu * T1); e2 += Dm * ((N1 - mu * T1) * (N1 - mu * T1) + NmT * (-mu * T2));
      }
    }
  }
  T v[3] = {ec + a * a * q2 + a * q1 + q0, e1 + 2 * a * q2 + q1, e2 + 2 * q2};
  block_sum(c, v, s.red);
  LsPt<T> p; p.alpha = a; p.cost = v[0] + a * a * qg[2] + a * qg[1] + qg[0]; p.d1 = v[1] + 2 * a * qg[2] + qg[1]; p.d2 = v[2] + 2 * qg[2];
  if (p.d2 < Num<T>::minval()) p.d2 = Num<T>::minval();
  return p;
}
// Jaref = J*x - aref, Ma = M*x, returns total cost (all threads)
template <typename T>
B2P_HD T cost_at(const DevModel<T>& m, Scratch<T>& s, const Ctx& c, const T* x) {
  for (int r = c.tid; r < s.nefc; r += c.nt) s.Jaref[r] = row_dot(m, s, r, x) - s.aref[r];
  mul_M(m, s, s.Ma, x, c);
  bsync(c);
  T v[1] = {constraint_update(m, s, c, false)};
  for (int d = c.tid; d < m.nv; d += c.nt) v[0] += T(0.5) * (s.Ma[d] - s.qfrc_smooth[d]) * (x[d] - s.qacc_smooth[d]);
  block_sum(c, v, s.red);
  return v[0];
}
// cost, gradient, Hessian factor, Newton direction at s.qacc (s.Jaref / s.Ma consistent with it)
template <typename T>
B2P_HD T solver_update(const DevModel<T>& m, Scratch<T>& s, const Ctx& c, T& gradnorm2) {
  T v[2] = {constraint_update(m, s, c, true), 0};
  bsync(c);
  mul_JT_force(m, s, c);
  assemble_H(m, s, c);
  bsync(c);
  for (int d = c.tid; d < m.nv; d += c.nt) {
    const T gd = s.Ma[d] - s.qfrc_smooth[d] - s.qfrc_con[d];
    s.grad[d] = gd; s.Mgrad[d] = gd; v[1] += gd * gd;
    v[0] += T(0.5) * (s.Ma[d] - s.qfrc_smooth[d]) * (s.qacc[d] - s.qacc_smooth[d]);
  }
  block_sum(c, v, s.red);
  block_cholesky(s.H, m.nv, MAXNV, c);
  warp_chol_solve(s.H, m.nv, MAXNV, s.Mgrad, c);
  bsync(c);
  gradnorm2 = v[1];
  return v[0];
}
template <typename T>
B2P_HD T line_search(const DevModel<T>& m, Scratch<T>& s, const Ctx& c, T scale) {
  T v[4] = {0, 0, 0, 0};
  mul_M(m, s, s.Mv, s.search, c);
  for (int r = c.tid; r < s.nefc; r += c.nt) s.Jv[r] = row_dot(m, s, r, s.search);
  bsync(c);
  for (int d = c.tid; d < m.nv; d += c.nt) {
    const T sd = s.search[d], res = s.Ma[d] - s.qfrc_smooth[d];
    v[0] += sd * sd; v[1] += sd * res; v[2] += T(0.5) * sd * s.Mv[d]; v[3] += T(0.5) * res * (s.qacc[d] - s.qacc_smooth[d]);
  }
  block_sum(c, v, s.red);
  const T snorm = sqrt(v[0]);
  if (snorm < Num<T>::minval()) return 0;
  const T gtol = m.tolerance * m.ls_tolerance * snorm / scale;
  const T qg[3] = {v[3], v[1], v[2]};
  LsPt<T> p0 = ls_eval(m, s, c, T(0), qg), p1 = ls_eval(m, s, c, -p0.d1 / p0.d2, qg), p2;
  if (p0.cost < p1.cost) p1 = p0;
  if (fabs(p1.d1) < gtol) return p1.alpha;
  const int dir = p1.d1 < 0 ? 1 : -1; int k = 0;
  p2 = p1;
  while (p1.d1 * dir <= -gtol && k < m.ls_iterations) {
    p2 = p1; p1 = ls_eval(m, s, c, p1.alpha - p1.d1 / p1.d2, qg); k++;
    if (fabs(p1.d1) < gtol) return p1.alpha;
  }
  if (k >= m.ls_iterations) return p1.alpha;
  LsPt<T> lo = p2, hi = p1, best = p1.cost < p2.cost ? p1 : p2;
  while (k < m.ls_iterations) {
    const LsPt<T>& from = fabs(lo.d1) < fabs(hi.d1) ? lo : hi;
    T a = from.alpha - from.d1 / from.d2; const T amin = lo.alpha < hi.alpha ? lo.alpha : hi.alpha, amax = lo.alpha < hi.alpha ? hi.alpha : lo.alpha;
    if (!(a > amin && a < amax)) a = T(0.5) * (lo.alpha + hi.alpha);
    LsPt<T> pm = ls_eval(m, s, c, a, qg); k++;
    if (pm.cost < best.cost) best = pm;
    if (fabs(pm.d1) < gtol) return pm.alpha;
    if (pm.d1 * dir < 0) lo = pm; else hi = pm;
    if (fabs(hi.alpha - lo.alpha) < 64 * Num<T>::eps() * (1 + fabs(hi.alpha))) break;
  }
  return best.alpha;
}
template <typename T>
B2P_HD void solve_constraints(const DevModel<T>& m, Scratch<T>& s, const Ctx& c) {
  const int nv = m.nv;
  if (s.nefc == 0) {
    for (int d = c.tid; d < nv; d += c.nt) { s.qacc[d] = s.qacc_smooth[d]; s.qfrc_con[d] = 0; }
    if (c.tid == 0) s.niter = 0;
    bsync(c);
    return;
  }
  const T scale = 1 / (m.meaninertia * (nv > 1 ? nv : 1));
  const T cw = cost_at(m, s, c, s.qacc_ws), cs = cost_at(m, s, c, s.qacc_smooth);
  for (int d = c.tid; d < nv; d += c.nt) s.qacc[d] = cw < cs ? s.qacc_ws[d] : s.qacc_smooth[d];
  bsync(c);
  cost_at(m, s, c, s.qacc);
  T gn2, cost = solver_update(m, s, c, gn2);
  for (int d = c.tid; d < nv; d += c.nt) s.search[d] = -s.Mgrad[d];
  bsync(c);
  int it = 0;
  for (; it < m.iterations; it++) {
    const T alpha = line_search(m, s, c, scale);
    if (alpha == 0) break;
    for (int d = c.tid; d < nv; d += c.nt) { s.qacc[d] += alpha * s.search[d]; s.Ma[d] += alpha * s.Mv[d]; }
    for (int r = c.tid; r < s.nefc; r += c.nt) s.Jaref[r] += alpha * s.Jv[r];
    bsync(c);
    const T old = cost;
    cost = solver_update(m, s, c, gn2);
    const T improvement = scale * (old - cost), gradient = scale * sqrt(gn2);
    if (improvement < m.tolerance || gradient < m.tolerance) { it++; break; }
    for (int d = c.tid; d < nv; d += c.nt) s.search[d] = -s.Mgrad[d];
    bsync(c);
  }
  if (c.tid == 0) { s.niter = it; if (it >= m.iterations) s.flags |= FLAG_SOLVER_MAXITER; }
  bsync(c);
}

// ------------------------------------------------------------------------------------------ integrator (mj_implicit / mj_Euler + mj_advance)
template <typename T>
B2P_HD void integrate(const DevModel<T>& m, Scratch<T>& s, const Ctx& c, bool freeze_tree) {
  const int n = m.ntd; const T h = m.timestep;
  for (int e = c.tid; e < n * n; e += c.nt) {
    const int i = e / n, j = e % n;
    T v = s.Marm[i * MAXTD + j];
    if (i == j) {
      v += h * m.dof_damping[i];
      if (m.integrator == B2P_INT_IMPLICITFAST)
        for (int a = 0; a < m.nu; a++) if (m.act_bias[a][2] != 0) for (int x = 0; x < m.act_trnnum[a]; x++) if (m.act_trndof[a][x] == i) v -= h * m.act_trncoef[a][x] * m.act_trncoef[a][x] * m.act_bias[a][2];
    } else if (m.integrator == B2P_INT_IMPLICITFAST) {
      // off-diagonal actuator terms only inside M's (ancestor) sparsity pattern
      const int hi = i > j ? i : j, lo = i > j ? j : i;
      if (m.anc_mask[hi + 1] & (1u << lo))
        for (int a = 0; a < m.nu; a++) if (m.act_bias[a][2] != 0 && m.act_trnnum[a] == 2) {
          const int d0 = m.act_trndof[a][0], d1 = m.act_trndof[a][1];
          if ((d0 == i && d1 == j) || (d0 == j && d1 == i)) v -= h * m.act_trncoef[a][0] * m.act_trncoef[a][1] * m.act_bias[a][2];
        }
    }
    s.H[i * MAXNV + j] = v;
  }
  for (int d = c.tid; d < m.nv; d += c.nt) {
    bool active = d < n || s.pactive[(d - n) / 6];
    const T f = active ? s.qfrc_smooth[d] + s.qfrc_con[d] : T(0);
    s.tmpv[d] = d < n ? f : f / s.Mdiag[d];
  }
  bsync(c);
  block_cholesky(s.H, n, MAXNV, c);
  warp_chol_solve(s.H, n, MAXNV, s.tmpv, c);
  bsync(c);
  for (int d = c.tid; d < m.nv; d += c.nt) { s.qvel[d] += h * s.tmpv[d]; s.qacc_ws[d] = s.qacc[d]; }
  bsync(c);
  for (int d = c.tid; d < n; d += c.nt) s.qpos[d] += h * s.qvel[d];
  for (int p = c.tid; p < m.nprop; p += c.nt) {
    T* q = s.qpos + n + 7 * p; const T* v = s.qvel + n + 6 * p;
    for (int k = 0; k < 3; k++) q[k] += h * v[k];
    T w[3] = {v[3], v[4], v[5]}; const T ang = h * normalize3(w), sn = sin(T(0.5) * ang);
    T qr[4] = {cos(T(0.5) * ang), sn * w[0], sn * w[1], sn * w[2]};
    mulquat(q + 3, q + 3, qr); normquat(q + 3);
  }
  if (c.tid == 0) s.time += h;
  bsync(c);
  if (freeze_tree) {   // dm_control JointStaticIsolator: non-prop joints restored after the step (App. B item 18)
    for (int d = c.tid; d < n; d += c.nt) { s.qpos[d] = s.hold_q[d]; s.qvel[d] = s.hold_v[d]; }
    bsync(c);
  }
}

// ------------------------------------------------------------------------------------------ OSC (mujoco_controllers/osc.py), executed by thread 0
template <typename T, int N>
B2P_HD void small_chol(T* A) {
  for (int j = 0; j < N; j++) {
    T d = A[j * N + j];
    for (int k = 0; k < j; k++) d -= A[j * N + k] * A[j * N + k];
    d = d > Num<T>::minval() ? sqrt(d) : T(0);
    A[j * N + j] = d;
    for (int i = j + 1; i < N; i++) { T v = A[i * N + j]; for (int k = 0; k < j; k++) v -= A[i * N + k] * A[j * N + k]; A[i * N + j] = d > 0 ? v / d : T(0); }
  }
}
template <typename T, int N>
B2P_HD void small_inv_from_chol(const T* L, T* Ainv) {
  for (int col = 0; col < N; col++) {
    T x[N];
    for (int i = 0; i < N; i++) { T v = (i == col) ? T(1) : T(0); for (int k = 0; k < i; k++) v -= L[i * N + k] * x[k]; x[i] = v / L[i * N + i]; }
    for (int i = N - 1; i >= 0; i--) { T v = x[i]; for (int k = i + 1; k < N; k++) v -= L[k * N + i] * x[k]; x[i] = v / L[i * N + i]; }
    for (int i = 0; i < N; i++) Ainv[i * N + col] = x[i];
  }
}
template <typename T>
B2P_HD void osc_ori_error(const T* q, const T* qd, T* e) {   // osc.py:132-150
  T qc[4] = {q[0], -q[1], -q[2], -q[3]}, qe[4]; normquat(qc);
  mulquat(qe, qd, qc);
  const T sg = qe[0] > 0 ? T(1) : (qe[0] < 0 ? T(-1) : T(0));
  e[0] = qe[1] * sg; e[1] = qe[2] * sg; e[2] = qe[3] * sg;
}
template <typename T>
B2P_HDN void osc_torque(const DevModel<T>& m, const Scratch<T>& s, const T* tg, T* tau) {
  T J[42], Ma[49], Minv[49], MiJt[42], Lxi[36], Lx[36], L[36];
  {  // osc.py:117-130 (mj_jacSite, arm columns)
    const T* pt = s.site_xpos[0]; T off[3] = {pt[0] - m.ref[0], pt[1] - m.ref[1], pt[2] - m.ref[2]};
    const unsigned mask = m.anc_mask[m.site_body[0]];
    for (int d = 0; d < 7; d++) {
      T cr[3] = {0, 0, 0}; const bool on = (mask >> d) & 1u;
      if (on) cross3(cr, s.cdof[d], off);
      for (int k = 0; k < 3; k++) { J[k * 7 + d] = on ? s.cdof[d][3 + k] + cr[k] : T(0); J[(3 + k) * 7 + d] = on ? s.cdof[d][k] : T(0); }
    }
  }
  for (int i = 0; i < 7; i++) for (int j = 0; j < 7; j++) Ma[i * 7 + j] = s.Marm[i * MAXTD + j];
  small_chol<T, 7>(Ma); small_inv_from_chol<T, 7>(Ma, Minv);                        // osc.py:106
  for (int i = 0; i < 7; i++) for (int r = 0; r < 6; r++) { T v = 0; for (int k = 0; k < 7; k++) v += Minv[i * 7 + k] * J[r * 7 + k]; MiJt[i * 6 + r] = v; }
  for (int r = 0; r < 6; r++) for (int q = 0; q < 6; q++) { T v = 0; for (int k = 0; k < 7; k++) v += J[r * 7 + k] * MiJt[k * 6 + q]; Lxi[r * 6 + q] = v; L[r * 6 + q] = v; }
  small_chol<T, 6>(L);
  T det = 1; for (int i = 0; i < 6; i++) det *= L[i * 6 + i] * L[i * 6 + i];
  if (fabs(det) >= T(1e-2)) small_inv_from_chol<T, 6>(L, Lx);                       // osc.py:109-112
  else {                                                                            // osc.py:113-115: pinv(rcond=1e-2) via Jacobi eigen-decomposition
    T A[36], V[36];
    for (int i = 0; i < 36; i++) { A[i] = Lxi[i]; V[i] = (i / 6 == i % 6) ? T(1) : T(0); }
    for (int sweep = 0; sweep < 60; sweep++) {
      T off = 0; for (int i = 0; i < 6; i++) for (int j = i + 1; j < 6; j++) off += A[i * 6 + j] * A[i * 6 + j];
      if (off < Num<T>::minval() * Num<T>::minval()) break;
      for (int p = 0; p < 6; p++) for (int q = p + 1; q < 6; q++) {
        if (fabs(A[p * 6 + q]) < Num<T>::minval() * Num<T>::minval()) continue;
        const T th = (A[q * 6 + q] - A[p * 6 + p]) / (2 * A[p * 6 + q]), t = (th >= 0 ? T(1) : T(-1)) / (fabs(th) + sqrt(th * th + 1)), cs = 1 / sqrt(t * t + 1), sn = t * cs;
        for (int k = 0; k < 6; k++) { const T a = A[k * 6 + p], b = A[k * 6 + q]; A[k * 6 + p] = cs * a - sn * b; A[k * 6 + q] = sn * a + cs * b; }
        for (int k = 0; k < 6; k++) { const T a = A[p * 6 + k], b = A[q * 6 + k]; A[p * 6 + k] = cs * a - sn * b; A[q * 6 + k] = sn * a + cs * b; }
        for (int k = 0; k < 6; k++) { const T a = V[k * 6 + p], b = V[k * 6 + q]; V[k * 6 + p] = cs * a - sn * b; V[k * 6 + q] = sn * a + cs * b; }
      }
    }
    T smax = 0; for (int i = 0; i < 6; i++) if (fabs(A[i * 6 + i]) > smax) smax = fabs(A[i * 6 + i]);
    for (int r = 0; r < 6; r++) for (int q = 0; q < 6; q++) { T v = 0; for (int k = 0; k < 6; k++) { const T ev = A[k * 6 + k]; if (fabs(ev) > T(1e-2) * smax) v += V[r * 6 + k] * V[q * 6 + k] / ev; } Lx[r * 6 + q] = v; }
  }
  T q[4], dx[6] = {0, 0, 0, 0, 0, 0}, pd[6], eo[3];
  mat2quat(q, s.site_xmat[0]);
  for (int r = 0; r < 6; r++) for (int i = 0; i < 7; i++) dx[r] += J[r * 7 + i] * s.qvel[i];
  const T* gn = m.osc_gains;
  for (int k = 0; k < 3; k++) { const T e = clampT(tg[k] - s.site_xpos[0][k], T(-0.05), T(0.05)); pd[k] = gn[0] * e + gn[1] * (tg[3 + k] - dx[k]); }   // osc.py:168-172
  T qd[4] = {tg[6], tg[7], tg[8], tg[9]}; normquat(qd);
  osc_ori_error(q, qd, eo);
  for (int k = 0; k < 3; k++) pd[3 + k] = gn[2] * eo[k] + gn[3] * (tg[10 + k] - dx[3 + k]);
  T Lpd[6], ne[7];
  for (int r = 0; r < 6; r++) { T v = 0; for (int k = 0; k < 6; k++) v += Lx[r * 6 + k] * pd[k]; Lpd[r] = v; }
  for (int i = 0; i < 7; i++) { T v = 0; for (int r = 0; r < 6; r++) v += J[r * 7 + i] * Lpd[r]; tau[i] = v; }            // osc.py:206-207
  for (int i = 0; i < 7; i++) ne[i] = gn[4] * (m.osc_null[i] - s.qpos[i]) + gn[5] * (0 - s.qvel[i]);
  T Nb[42];
  for (int i = 0; i < 7; i++) for (int k = 0; k < 6; k++) { T v = 0; for (int r = 0; r < 6; r++) v += MiJt[i * 6 + r] * Lx[r * 6 + k]; Nb[i * 6 + k] = v; }   // osc.py:218
  for (int i = 0; i < 7; i++) {
    T v = ne[i];
    for (int j = 0; j < 7; j++) { T jn = 0; for (int k = 0; k < 6; k++) jn += J[k * 7 + i] * Nb[j * 6 + k]; v -= jn * ne[j]; }
    tau[i] += v + s.qfrc_bias[i];                                                                                     // osc.py:219-222
  }
  // osc.py:224-236: tau @ pinv(actuator_moment)[arm, arm] is the identity for unit-gear joint motors
}
template <typename T>
B2P_HD bool osc_converged(const DevModel<T>& m, const Scratch<T>& s, const T* tg) {   // osc.py:240-247
  T q[4], e[3], d[3], qd[4] = {tg[6], tg[7], tg[8], tg[9]};
  normquat(qd); mat2quat(q, s.site_xmat[0]); osc_ori_error(q, qd, e);
  for (int k = 0; k < 3; k++) d[k] = s.site_xpos[0][k] - tg[k];
  T oe = fabs(e[0]); if (fabs(e[1]) > oe) oe = fabs(e[1]); if (fabs(e[2]) > oe) oe = fabs(e[2]);
  return norm3(d) < m.osc_pos_thresh && oe < m.osc_ori_thresh;
}

// ------------------------------------------------------------------------------------------ pipeline
template <typename T>
B2P_HD void prop_params(const DevModel<T>& m, Scratch<T>& s, const Ctx& c) {
  for (int p = c.tid; p < m.nprop; p += c.nt) {
    const T* z = s.psize[p]; const T mass = m.prop_mass[p];
    T I[3] = {mass / 3 * (z[1] * z[1] + z[2] * z[2]), mass / 3 * (z[0] * z[0] + z[2] * z[2]), mass / 3 * (z[0] * z[0] + z[1] * z[1])};
    for (int k = 0; k < 3; k++) s.pinertia[p][k] = I[k];
    s.pinvw[p][0] = 1 / mass; s.pinvw[p][1] = (1 / I[0] + 1 / I[1] + 1 / I[2]) / 3;
  }
  bsync(c);
}
template <typename T>
B2P_HD void step1(const DevModel<T>& m, const Globals<T>& g, Scratch<T>& s, const Ctx& c, bool full_contacts) {
  forward_sweep(m, s, c);
  backward_sweep(m, s, c);
  dynamics_terms(m, s, c);
  collide(m, g, s, c, full_contacts);
  make_rows(m, g, s, c);
}
template <typename T>
B2P_HD void step2(const DevModel<T>& m, Scratch<T>& s, const Ctx& c, bool do_integrate, bool freeze_tree) {
  actuation_and_smooth(m, s, c);
  solve_constraints(m, s, c);
  if (c.tid == 0) {
    for (int d = 0; d < m.nv; d++) { if (!(fabs(s.qacc[d]) < T(1e10))) s.flags |= FLAG_BADQACC; if (!(fabs(s.qvel[d]) < T(1e10))) s.flags |= FLAG_BADQVEL; }
    for (int d = 0; d < m.nq; d++) if (!(fabs(s.qpos[d]) < T(1e10))) s.flags |= FLAG_BADQPOS;
  }
  if (do_integrate) integrate(m, s, c, freeze_tree);
  else bsync(c);
}

// HBM state record of one env = the head of Scratch<T>: qpos[MAXNQ] | qvel[MAXNV] | qacc_warmstart[MAXNV] | time | pad
// launch parameters shared by kernel and host emulation
template <typename T>
struct StepArgs {
  int mode, n_env, n_steps, control_steps, state_stride, full_contacts;
  T* state;                  // [n_env][state_stride]: qpos | qvel | qacc_warmstart | time
  const T* ctrl;             // MODE_PHYSICS / SETTLE: [n_env][nu]
  const T* target;           // MODE_OSC: [n_env][13] pos3 vel3 quat4(wxyz) angvel3
  const unsigned char* grip; // MODE_OSC: [n_env] 0 = min, 1 = max
  const T* prop_size;        // [n_env][MAXP][3] half sizes
  const unsigned char* prop_active; // [n_env][MAXP]
  const unsigned char* env_mask;    // optional [n_env]: 0 = skip env
  unsigned char* converged;  // MODE_OSC out [n_env]
  int* status;               // out [n_env] OR-ed flags
  int* steps_done;           // MODE_SETTLE out [n_env]
  T settle_qvel_tol, settle_qacc_tol; int settle_min_steps;
  T* out;                    // MODE_FORWARD / any: optional derived-output record per env (see OutLayout)
  int out_stride;
};
// derived-output record (reals): site_xpos[2][3] site_xmat[2][9] qacc[MAXNV] qfrc_bias[MAXNV] ctrl[MAXU] ncon nefc niter |
//   contacts[MAXCON] x (dist, pos3, normal3, geom1, geom2, excluded) | Marm[MAXTD*MAXTD] | gpos[MAXG][3] | gmat[MAXG][9]
struct OutLayout {
  static constexpr int SITE_XPOS = 0, SITE_XMAT = 6, QACC = 24, QFRC_BIAS = QACC + MAXNV, CTRL = QFRC_BIAS + MAXNV, NCON = CTRL + MAXU, NEFC = NCON + 1,
                       NITER = NEFC + 1, CONTACT = NITER + 1, CONTACT_W = 10, MARM = CONTACT + MAXCON * CONTACT_W, GPOS = MARM + MAXTD * MAXTD,
                       GMAT = GPOS + MAXG * 3, QACC_SMOOTH = GMAT + MAXG * 9, EFC_FORCE = QACC_SMOOTH + MAXNV, SIZE = EFC_FORCE + MAXEFC;
};
template <typename T>
B2P_HD void write_out(const DevModel<T>& m, const Globals<T>& g, const Scratch<T>& s, const Ctx& c, T* o) {
  typedef OutLayout L;
  for (int i = c.tid; i < 6; i += c.nt) o[L::SITE_XPOS + i] = s.site_xpos[i / 3][i % 3];
  for (int i = c.tid; i < 18; i += c.nt) o[L::SITE_XMAT + i] = s.site_xmat[i / 9][i % 9];
  for (int i = c.tid; i < MAXNV; i += c.nt) { o[L::QACC + i] = i < m.nv ? s.qacc[i] : T(0); o[L::QFRC_BIAS + i] = i < m.nv ? s.qfrc_bias[i] : T(0); o[L::QACC_SMOOTH + i] = i < m.nv ? s.qacc_smooth[i] : T(0); }
  for (int i = c.tid; i < MAXU; i += c.nt) o[L::CTRL + i] = s.ctrl[i];
  if (c.tid == 0) { o[L::NCON] = T(s.ncon); o[L::NEFC] = T(s.nefc); o[L::NITER] = T(s.niter); }
  for (int k = c.tid; k < MAXCON; k += c.nt) {
    T* q = o + L::CONTACT + k * L::CONTACT_W;
    if (k < s.ncon) {
      const int p = s.c_pair[k];
      q[0] = s.c_dist[k]; for (int i = 0; i < 3; i++) { q[1 + i] = s.c_pos[k][i]; q[4 + i] = s.c_frame[k][i]; }
      q[7] = T(g.pair_g1[p]); q[8] = T(g.pair_g2[p]); q[9] = T(s.c_dist[k] < s.c_incl[k] ? 0 : 1);
    } else for (int i = 0; i < L::CONTACT_W; i++) q[i] = 0;
  }
  for (int i = c.tid; i < MAXTD * MAXTD; i += c.nt) o[L::MARM + i] = s.Marm[i];
  for (int i = c.tid; i < MAXG * 3; i += c.nt) o[L::GPOS + i] = i / 3 < m.ngeom ? s.gpos[i / 3][i % 3] : T(0);
  for (int i = c.tid; i < MAXG * 9; i += c.nt) o[L::GMAT + i] = i / 9 < m.ngeom ? s.gmat[i / 9][i % 9] : T(0);
  for (int i = c.tid; i < MAXEFC; i += c.nt) o[L::EFC_FORCE + i] = i < s.nefc ? s.force[i] : T(0);
}

// One scene, whole launch: load the state record, run the requested mode, store the record.
template <typename T>
B2P_HD void run_env(const DevModel<T>& m, const Globals<T>& g, Scratch<T>& s, const StepArgs<T>& a, int env, const Ctx& c, bool state_preloaded) {
  const int nv = m.nv;
  T* rec = a.state + (size_t)env * a.state_stride;
  if (!state_preloaded) {   // the CUDA kernel stages the record with one TMA bulk copy instead
    T* head = s.qpos;
    for (int i = c.tid; i < REC_SIZE; i += c.nt) head[i] = rec[i];
  }
  for (int i = c.tid; i < MAXP * 3; i += c.nt) s.psize[i / 3][i % 3] = a.prop_size ? a.prop_size[(size_t)env * MAXP * 3 + i] : m.geom_size[0][0];
  for (int i = c.tid; i < MAXP; i += c.nt) s.pactive[i] = i < m.nprop ? (a.prop_active ? a.prop_active[(size_t)env * MAXP + i] : 1) : 0;
  for (int i = c.tid; i < MAXU; i += c.nt) s.ctrl[i] = (a.ctrl && i < m.nu) ? a.ctrl[(size_t)env * m.nu + i] : T(0);
  if (c.tid == 0) { s.flags = 0; s.niter = 0; s.ncon = 0; s.nefc = 0; }
  bsync(c);
  if (!a.prop_size) {   // default: sizes from the model's prop geoms
    for (int g0 = c.tid; g0 < m.ngeom; g0 += c.nt) if (m.geom_prop[g0] >= 0) for (int k = 0; k < 3; k++) s.psize[m.geom_prop[g0]][k] = m.geom_size[g0][k];
    bsync(c);
  }
  prop_params(m, s, c);
  const bool full = a.full_contacts != 0;
  if (a.mode == MODE_FORWARD) {
    step1(m, g, s, c, full);
    step2(m, s, c, false, false);
  } else if (a.mode == MODE_PHYSICS) {
    for (int k = 0; k < a.n_steps; k++) { step1(m, g, s, c, full); step2(m, s, c, true, false); }
    step1(m, g, s, c, full);   // leave derived quantities consistent with the final state (dm_control legacy step order)
  } else if (a.mode == MODE_OSC) {
    // RobotArm.run_controller (robot_arm.py:61-94): per tick OSC torque + gripper command, control_steps x physics.step()
    T tg[13];
    for (int i = 0; i < 13; i++) tg[i] = a.target[(size_t)env * 13 + i];
    const bool grip = a.grip ? a.grip[env] != 0 : false;
    bool conv = false;
    step1(m, g, s, c, full);
    for (int t = 0; t < a.n_steps; t++) {
      if (c.tid == 0) { T tau[7]; osc_torque(m, s, tg, tau); for (int i = 0; i < 7; i++) s.ctrl[i] = tau[i]; s.ctrl[7] = grip ? T(255) : T(0); }   // min_max.py:14-19
      bsync(c);
      for (int k = 0; k < a.control_steps; k++) { step2(m, s, c, true, false); step1(m, g, s, c, full); }
      if (osc_converged(m, s, tg)) conv = true;
    }
    if (c.tid == 0 && a.converged) a.converged[env] = conv ? 1 : 0;
  } else if (a.mode == MODE_SETTLE) {
    // PropPlacer settle loop (prop_initializer.py:245-257): arm + gripper held static, per-env early exit
    for (int d = c.tid; d < m.ntd; d += c.nt) { s.hold_q[d] = s.qpos[d]; s.hold_v[d] = s.qvel[d]; }
    bsync(c);
    int done = 0;
    for (int k = 0; k < a.n_steps; k++) {
      step1(m, g, s, c, full); step2(m, s, c, true, true);
      done = k + 1;
      T mx[2] = {0, 0};   // max |qvel|, |qacc| over ACTIVE prop dofs: computed redundantly by every thread (nv is tiny)
      for (int d = m.ntd; d < nv; d++) if (s.pactive[(d - m.ntd) / 6]) { const T v = fabs(s.qvel[d]), ac = fabs(s.qacc[d]); if (v > mx[0]) mx[0] = v; if (ac > mx[1]) mx[1] = ac; }
      bsync(c);
      if (mx[0] < a.settle_qvel_tol && mx[1] < a.settle_qacc_tol && done > a.settle_min_steps) break;
    }
    step1(m, g, s, c, full);
    if (c.tid == 0 && a.steps_done) a.steps_done[env] = done;
  }
  bsync(c);
  if (!state_preloaded) { const T* head = s.qpos; for (int i = c.tid; i < REC_SIZE; i += c.nt) rec[i] = head[i]; }
  if (c.tid == 0 && a.status) a.status[env] |= s.flags;
  if (a.out) write_out(m, g, s, c, a.out + (size_t)env * a.out_stride);
}

}  // namespace b2p
