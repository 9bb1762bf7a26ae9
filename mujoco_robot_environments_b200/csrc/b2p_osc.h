#pragma once
#include "b2p_scratch.h"

namespace b2p {

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
template <typename T>
B2P_HD void osc_ori_error(const T* q, const T* qd, T* e) {   // osc.py:132-150
  T qc[4] = {q[0], -q[1], -q[2], -q[3]}, qe[4]; normquat(qc);
  mulquat(qe, qd, qc);
  const T sg = qe[0] > 0 ? T(1) : (qe[0] < 0 ? T(-1) : T(0));
  e[0] = qe[1] * sg; e[1] = qe[2] * sg; e[2] = qe[3] * sg;
}
// one column of (L L')^-1 from the Cholesky factor L (row-major N x N): the columns are independent, one lane each
template <typename T, int N>
B2P_HD void small_inv_col(const T* L, T* Ainv, int col) {
  T x[N];
#pragma unroll
  for (int i = 0; i < N; i++) { T v = (i == col) ? T(1) : T(0);
#pragma unroll
    for (int k = 0; k < i; k++) v -= L[i * N + k] * x[k];
    x[i] = v / L[i * N + i]; }
#pragma unroll
  for (int i = N - 1; i >= 0; i--) { T v = x[i];
#pragma unroll
    for (int k = i + 1; k < N; k++) v -= L[k * N + i] * x[k];
    x[i] = v / L[i * N + i]; }
#pragma unroll
  for (int i = 0; i < N; i++) Ainv[i * N + col] = x[i];
}
// OSC.compute_control_output (osc.py:180-238) for the scene's current state -> s.ctrl[0..6] (+ the gripper command in
// s.ctrl[7], min_max.py:14-19).  Warp-cooperative: the small matrices live in shared memory (the part of the H / FK-sweep
// union that is dead between step1 and step2) and their entries / columns are spread over the lanes; the two tiny Cholesky
// factorisations and the PD terms run on lane 0.
template <typename T, bool C>
B2P_HDN void osc_torque(const DevModel<T>& m, Scratch<T, C>& s, const T* tg, bool grip, const Ctx& c, const T* gn) {
  T* w = s.H;
  T *J = w, *Ma = w + 42, *Minv = w + 91, *MiJt = w + 140, *Lxi = w + 182, *L = w + 218, *Lx = w + 254, *Nb = w + 290;
  T *dx = w + 332, *pd = w + 338, *Lpd = w + 344, *ne = w + 350, *A = w + 360, *V = w + 396;
  static_assert(396 + 36 <= MAXTB * 28, "OSC scratch must stay inside the dead FK-sweep buffers (crb, cvel, cacc, cfrc)");
  {  // osc.py:117-130 (mj_jacSite, arm columns)
    const T* pt = s.site_xpos[0]; const T off[3] = {pt[0] - m.ref[0], pt[1] - m.ref[1], pt[2] - m.ref[2]};
    const unsigned mask = m.anc_mask[m.site_body[0]];
    for (int e = c.tid; e < 42; e += c.nt) {
      const int r = e / 7, d = e - 7 * r; const bool on = (mask >> d) & 1u; T v = 0;
      if (on) { if (r < 3) { T cr[3]; cross3(cr, s.cdof[d], off); v = s.cdof[d][3 + r] + (r == 0 ? cr[0] : (r == 1 ? cr[1] : cr[2])); } else v = s.cdof[d][r - 3]; }
      J[e] = v;
    }
  }
  for (int e = c.tid; e < 49; e += c.nt) { const int i = e / 7, j = e - 7 * i; Ma[e] = s.Mt[i >= j ? tri_idx(i, j) : tri_idx(j, i)]; }
  bsync(c);
  if (c.tid == 0) small_chol<T, 7>(Ma);                                                                             // osc.py:106
  bsync(c);
  for (int col = c.tid; col < 7; col += c.nt) small_inv_col<T, 7>(Ma, Minv, col);
  bsync(c);
  for (int e = c.tid; e < 42; e += c.nt) { const int i = e / 6, r = e - 6 * i; T v = 0; for (int k = 0; k < 7; k++) v += Minv[i * 7 + k] * J[r * 7 + k]; MiJt[e] = v; }
  bsync(c);
  for (int e = c.tid; e < 36; e += c.nt) { const int r = e / 6, q = e - 6 * r; T v = 0; for (int k = 0; k < 7; k++) v += J[r * 7 + k] * MiJt[k * 6 + q]; Lxi[e] = v; L[e] = v; }
  bsync(c);
  if (c.tid == 0) small_chol<T, 6>(L);
  bsync(c);
  T det = 1; for (int i = 0; i < 6; i++) det *= L[i * 6 + i] * L[i * 6 + i];
  if (fabs(det) >= T(1e-2)) { for (int col = c.tid; col < 6; col += c.nt) small_inv_col<T, 6>(L, Lx, col); }        // osc.py:109-112
  else if (c.tid == 0) {                                                            // osc.py:113-115: pinv(rcond=1e-2) via Jacobi eigen-decomposition
    for (int i = 0; i < 36; i++) { A[i] = Lxi[i]; V[i] = (i / 6 == i % 6) ? T(1) : T(0); }
    for (int sweep = 0; sweep < 60; sweep++) {
      T offd = 0; for (int i = 0; i < 6; i++) for (int j = i + 1; j < 6; j++) offd += A[i * 6 + j] * A[i * 6 + j];
      if (offd < Num<T>::minval() * Num<T>::minval()) break;
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
  for (int r = c.tid; r < 6; r += c.nt) { T v = 0; for (int i = 0; i < 7; i++) v += J[r * 7 + i] * s.qvel[i]; dx[r] = v; }
  for (int i = c.tid; i < 7; i += c.nt) ne[i] = gn[4] * (m.osc_null[i] - s.qpos[i]) + gn[5] * (0 - s.qvel[i]);
  bsync(c);
  if (c.tid == 0) {
    T q[4], eo[3];
    mat2quat(q, s.site_xmat[0]);
    for (int k = 0; k < 3; k++) { const T e = clampT(tg[k] - s.site_xpos[0][k], T(-0.05), T(0.05)); pd[k] = gn[0] * e + gn[1] * (tg[3 + k] - dx[k]); }   // osc.py:168-172
    T qd[4] = {tg[6], tg[7], tg[8], tg[9]}; normquat(qd);
    osc_ori_error(q, qd, eo);
    for (int k = 0; k < 3; k++) pd[3 + k] = gn[2] * eo[k] + gn[3] * (tg[10 + k] - dx[3 + k]);
  }
  bsync(c);
  for (int r = c.tid; r < 6; r += c.nt) { T v = 0; for (int k = 0; k < 6; k++) v += Lx[r * 6 + k] * pd[k]; Lpd[r] = v; }
  for (int e = c.tid; e < 42; e += c.nt) { const int i = e / 6, k = e - 6 * i; T v = 0; for (int r = 0; r < 6; r++) v += MiJt[i * 6 + r] * Lx[r * 6 + k]; Nb[e] = v; }   // osc.py:218
  bsync(c);
  for (int i = c.tid; i < 7; i += c.nt) {
    T t = 0; for (int r = 0; r < 6; r++) t += J[r * 7 + i] * Lpd[r];                                                  // osc.py:206-207
    T v = ne[i];
    for (int j = 0; j < 7; j++) { T jn = 0; for (int k = 0; k < 6; k++) jn += J[k * 7 + i] * Nb[j * 6 + k]; v -= jn * ne[j]; }
    s.ctrl[i] = t + (v + s.qfrc_bias[i]);                                                                             // osc.py:219-222
  }
  // osc.py:224-236: tau @ pinv(actuator_moment)[arm, arm] is the identity for unit-gear joint motors
  if (c.tid == 0) s.ctrl[7] = grip ? T(255) : T(0);                                                                  // min_max.py:14-19
  bsync(c);
}
template <typename T, bool C>
B2P_HD bool osc_converged(const DevModel<T>& m, const Scratch<T, C>& s, const T* tg) {   // osc.py:240-247
  T q[4], e[3], d[3], qd[4] = {tg[6], tg[7], tg[8], tg[9]};
  normquat(qd); mat2quat(q, s.site_xmat[0]); osc_ori_error(q, qd, e);
  for (int k = 0; k < 3; k++) d[k] = s.site_xpos[0][k] - tg[k];
  T oe = fabs(e[0]); if (fabs(e[1]) > oe) oe = fabs(e[1]); if (fabs(e[2]) > oe) oe = fabs(e[2]);
  return norm3(d) < m.osc_pos_thresh && oe < m.osc_ori_thresh;
}


}  // namespace b2p
