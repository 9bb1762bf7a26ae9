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
  for (int i = 0; i < 7; i++) for (int j = 0; j < 7; j++) Ma[i * 7 + j] = s.Mt[i >= j ? tri_idx(i, j) : tri_idx(j, i)];
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


}  // namespace b2p
