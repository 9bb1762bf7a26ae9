// Projected Gauss-Seidel on the DUAL of the problem the Newton solver minimises (MuJoCo's mj_solPGS; selected with b2p_set_solver).
//
//   minimise  1/2 f'(A + R) f + f'b   over f in Omega,    A = J M^-1 J',  b = J qacc_smooth - aref,
//   Omega: equality rows free, limit rows f >= 0, elliptic contacts f_n >= 0, |f_t / mu| <= f_n.
//
// A is nefc x nefc (up to 160 x 160 reals: more than a scene's whole shared-memory budget), so the sweep is matrix-free: the primal
// point a = qacc_smooth + M^-1 J' f is kept up to date and the residual of row r is J_r.a - aref_r + R_r f_r.  What is stored is
// X = M^-1 J' with J's own packed sparsity (M is block diagonal: the tree block and one diagonal 6-block per prop) in a per-warp
// global-memory pool, and the diagonal blocks of A + R (one real per scalar row, a symmetric 3 x 3 per contact).  Rows are visited
// in MuJoCo's order, one after the other (Gauss-Seidel), each with the warp's lanes across the row's columns.
//
// This solver is here as the independent check SURVEY 8(c) fixture 3 asks for (two different minimisers of one convex problem must
// agree: that pins R, D, contact.mu and the cone zones jointly) and because the north star names it; it is several times slower
// than Newton on this model class (tens of sweeps of sequential row updates per step), so Newton stays the default.
#pragma once
#include "b2p_solver.h"

namespace b2p {

template <typename T, bool C> B2P_HD T* xrow(const Scratch<T, C>& s, int r) { const int o = s.joff[r]; return s.xpool + (o >= 0 ? o : -o - 1); }

// min 1/2 x'Ax + x'b  s.t. (x0/d0)^2 + (x1/d1)^2 <= r^2  (mju_QCQP2): Newton on the multiplier of the scaled problem y = x / d
template <typename T>
B2P_HD void qcqp2(const T* A, const T* b, const T* d, T r, T* x) {
  const T a00 = A[0] * d[0] * d[0], a01 = A[1] * d[0] * d[1], a11 = A[3] * d[1] * d[1], b0 = b[0] * d[0], b1 = b[1] * d[1];
  T la = 0, y0 = 0, y1 = 0;
  for (int it = 0; it < 40; it++) {
    const T m00 = a00 + la, m11 = a11 + la, det = m00 * m11 - a01 * a01;
    if (det < Num<T>::minval() * Num<T>::minval()) break;
    y0 = -(m11 * b0 - a01 * b1) / det; y1 = -(-a01 * b0 + m00 * b1) / det;
    const T val = y0 * y0 + y1 * y1 - r * r;
    if (val < T(64) * Num<T>::eps() * (1 + r * r)) break;                      // inside (la = 0) or on the boundary
    const T z0 = (m11 * y0 - a01 * y1) / det, z1 = (-a01 * y0 + m00 * y1) / det;   // (As + la I)^-1 y
    const T deriv = -2 * (y0 * z0 + y1 * z1);
    if (deriv > -Num<T>::minval() * Num<T>::minval()) break;
    la -= val / deriv;
    if (la < 0) la = 0;
  }
  x[0] = y0 * d[0]; x[1] = y1 * d[1];
}

// J_r . x with the warp's lanes across the row's columns (deterministic butterfly sum: every lane ends with the same bits)
template <typename T, int K, bool C>
B2P_HD void rows_dot(const DevModel<T>& m, const Scratch<T, C>& s, const Ctx& c, int r, const T* x, T (&out)[K]) {
  const int a = s.row_a[r], b = s.row_b[r], wa = block_width(m, a), sa = block_start(m, a), w = wa + (b >= 0 ? 6 : 0), sb = b >= 0 ? block_start(m, b) : 0;
  for (int k = 0; k < K; k++) out[k] = 0;
  for (int e = c.tid; e < w; e += c.nt) {
    const T xv = x[e < wa ? sa + e : sb + (e - wa)];
    for (int k = 0; k < K; k++) out[k] += jrow(s, r + k)[e] * xv;
  }
  block_sum(c, out);
}
// a += sum_k X_{r+k} * df_k
template <typename T, int K, bool C>
B2P_HD void rows_axpy(const DevModel<T>& m, Scratch<T, C>& s, const Ctx& c, int r, const T (&df)[K], T* acc) {
  const int a = s.row_a[r], b = s.row_b[r], wa = block_width(m, a), sa = block_start(m, a), w = wa + (b >= 0 ? 6 : 0), sb = b >= 0 ? block_start(m, b) : 0;
  for (int e = c.tid; e < w; e += c.nt) {
    T v = 0;
    for (int k = 0; k < K; k++) v += xrow(s, r + k)[e] * df[k];
    acc[e < wa ? sa + e : sb + (e - wa)] += v;
  }
  bsync(c);
}

template <typename T, bool C>
B2P_HDN void solve_pgs(const DevModel<T>& m, Scratch<T, C>& s, const Ctx& c, bool active) {
  const int nv = m.nv;
  const bool has = active && s.nefc > 0;
  if (active && !has) {
    for (int d = c.tid; d < nv; d += c.nt) { s.qacc[d] = s.qacc_smooth[d]; s.qfrc_con[d] = 0; }
    if (c.tid == 0) s.niter = 0;
    bsync(c);
  }
  if (!has) return;
  const int n = s.nefc;
  const T scale = 1 / (m.meaninertia * (nv > 1 ? nv : 1));
  // ---- X = M^-1 J' (prop segments: diagonal; tree segments: one triangular solve with the factor of M per row)
  for (int r = c.tid; r < n; r += c.nt) {
    const int a = s.row_a[r], b = s.row_b[r], wa = block_width(m, a);
    const T* J = jrow(s, r); T* X = xrow(s, r);
    if (a != 0) for (int k = 0; k < 6; k++) X[k] = J[k] / s.Mdiag[block_start(m, a) + k];
    if (b >= 0) for (int k = 0; k < 6; k++) X[wa + k] = J[wa + k] / s.Mdiag[block_start(m, b) + k];
  }
  bsync(c);
  for (int r = 0; r < n; r++) {
    if (s.row_a[r] != 0) continue;
    const T* J = jrow(s, r); T* X = xrow(s, r);
    for (int k = c.tid; k < m.ntd; k += c.nt) s.tmpv[k] = J[k];
    bsync(c);
    sp_solve(s.Larm, m.sp_tree, s.tmpv, 0, m.ntd, c);
    for (int k = c.tid; k < m.ntd; k += c.nt) X[k] = s.tmpv[k];
    bsync(c);
  }
  // ---- warm start: the forces the primal map assigns to qacc_warmstart, if their dual cost beats f = 0
  cost_at(m, s, c, s.qacc_ws);            // s.force <- primal forces at the warm start (s.Jaref is scratch here)
  bsync(c);
  for (int d = c.tid; d < nv; d += c.nt) s.search[d] = 0;       // search <- M^-1 J' f
  bsync(c);
  for (int r = 0; r < n; r++) { const T df[1] = {s.force[r]}; if (df[0] != 0) rows_axpy<T, 1>(m, s, c, r, df, s.search); }
  {
    T v[1] = {0};
    for (int r = c.tid; r < n; r += c.nt) {
      const T f = s.force[r];
      v[0] += f * (T(0.5) * (row_dot(m, s, r, s.search) + f / s.D[r]) + row_dot(m, s, r, s.qacc_smooth) - s.aref[r]);
    }
    block_sum(c, v);
    const bool keep = v[0] < 0;
    for (int r = c.tid; r < n; r += c.nt) if (!keep) s.force[r] = 0;
    for (int d = c.tid; d < nv; d += c.nt) s.qacc[d] = s.qacc_smooth[d] + (keep ? s.search[d] : T(0));
    bsync(c);
  }
  // ---- diagonal blocks of A + R: scalar rows -> s.Jv[r]; contacts -> s.cW[k] = (00, 11, 22, 01, 02, 12)
  for (int r = c.tid; r < n; r += c.nt) {
    if (s.row_type[r] == ROW_CONTACT) continue;
    const int a = s.row_a[r], b = s.row_b[r], w = block_width(m, a) + (b >= 0 ? 6 : 0);
    const T* J = jrow(s, r); const T* X = xrow(s, r); T v = 0;
    for (int k = 0; k < w; k++) v += J[k] * X[k];
    s.Jv[r] = v + 1 / s.D[r];
  }
  for (int e = c.tid; e < 6 * s.ncon; e += c.nt) {
    const int k = e / 6, q = e - 6 * k, r = s.c_row[k];
    if (r < 0) continue;
    const int i = q < 3 ? q : (q == 5 ? 1 : 0), j = q < 3 ? q : (q == 3 ? 1 : 2);
    const int a = s.row_a[r], b = s.row_b[r], w = block_width(m, a) + (b >= 0 ? 6 : 0);
    const T* J = jrow(s, r + i); const T* X = xrow(s, r + j); T v = 0;
    for (int t = 0; t < w; t++) v += J[t] * X[t];
    s.cW[k][q] = v + (i == j ? 1 / s.D[r + i] : T(0));
  }
  bsync(c);
  // ---- sweeps
  int it = 0;
  for (; it < m.iterations; it++) {
    T improvement = 0;
    for (int r = 0; r < n; r++) {
      if (s.row_type[r] != ROW_CONTACT) {
        T rs[1]; rows_dot<T, 1>(m, s, c, r, s.qacc, rs);
        const T old = s.force[r], Arr = s.Jv[r];
        rs[0] += -s.aref[r] + old / s.D[r];
        T f = old - rs[0] / Arr;
        if (s.row_type[r] != ROW_EQ && f < 0) f = 0;
        const T d = f - old; T dc = d * rs[0] + T(0.5) * d * Arr * d;
        if (dc > T(1e-10)) { f = old; dc = 0; }
        improvement -= dc;
        if (f != old) { bsync(c); if (c.tid == 0) s.force[r] = f; const T df[1] = {f - old}; rows_axpy<T, 1>(m, s, c, r, df, s.qacc); }
        continue;
      }
      const int k = s.row_id[r]; const T* W = s.cW[k];
      const T Ab[9] = {W[0], W[3], W[4], W[3], W[1], W[5], W[4], W[5], W[2]};
      T rs[3]; rows_dot<T, 3>(m, s, c, r, s.qacc, rs);
      T old[3], f[3];
      for (int i = 0; i < 3; i++) { old[i] = s.force[r + i]; rs[i] += -s.aref[r + i] + old[i] / s.D[r + i]; f[i] = old[i]; }
      const T mv = Num<T>::minval();
      if (f[0] > mv) {                                            // (1) move along the ray through the current force
        T v1[3], den = 0, num = 0;
        for (int i = 0; i < 3; i++) { v1[i] = 0; for (int j = 0; j < 3; j++) v1[i] += Ab[i * 3 + j] * old[j]; }
        for (int i = 0; i < 3; i++) { den += old[i] * v1[i]; num += old[i] * rs[i]; }
        if (den >= mv) {
          T x = -num / den;
          if (f[0] + x * old[0] < 0) x = -1;                     // do not leave the cone through the apex
          for (int i = 0; i < 3; i++) f[i] += x * old[i];
        }
      }
      T bc[3];                                                    // gradient of the block cost with the rest fixed, at f = 0
      for (int i = 0; i < 3; i++) { bc[i] = rs[i]; for (int j = 0; j < 3; j++) bc[i] -= Ab[i * 3 + j] * old[j]; }
      if (f[0] < mv) {                                            // (2) at the apex: minimise along the projection of -bc onto the cone
        const T mu1 = s.c_fric[k], an = -bc[0], bt0 = -bc[1], bt1 = -bc[2], nb = sqrt(bt0 * bt0 + bt1 * bt1);
        T d[3] = {0, 0, 0};
        if (nb <= mu1 * an) { d[0] = an; d[1] = bt0; d[2] = bt1; }
        else if (mu1 * nb > -an) { d[0] = (an + mu1 * nb) / (1 + mu1 * mu1); d[1] = mu1 * d[0] * bt0 / nb; d[2] = mu1 * d[0] * bt1 / nb; }
        T den = 0, num = 0;
        for (int i = 0; i < 3; i++) { num += d[i] * bc[i]; for (int j = 0; j < 3; j++) den += d[i] * Ab[i * 3 + j] * d[j]; }
        T x = den > mv ? -num / den : T(0);
        if (x < 0) x = 0;
        for (int i = 0; i < 3; i++) f[i] = x * d[i];
      }
      if (f[0] >= mv) {                                           // (3) friction components with the normal held fixed
        const T A2[4] = {Ab[4], Ab[5], Ab[7], Ab[8]}, b2[2] = {bc[1] + Ab[3] * f[0], bc[2] + Ab[6] * f[0]}, d2[2] = {s.c_fric[k], s.c_fric[k]};
        T x2[2]; qcqp2(A2, b2, d2, f[0], x2);
        f[1] = x2[0]; f[2] = x2[1];
      }
      T dc = 0, df[3];
      for (int i = 0; i < 3; i++) df[i] = f[i] - old[i];
      for (int i = 0; i < 3; i++) { dc += df[i] * rs[i]; for (int j = 0; j < 3; j++) dc += T(0.5) * df[i] * Ab[i * 3 + j] * df[j]; }
      if (dc > T(1e-10)) { for (int i = 0; i < 3; i++) { f[i] = old[i]; df[i] = 0; } dc = 0; }   // never accept an increase
      improvement -= dc;
      if (df[0] != 0 || df[1] != 0 || df[2] != 0) {
        bsync(c);
        if (c.tid == 0) for (int i = 0; i < 3; i++) s.force[r + i] = f[i];
        rows_axpy<T, 3>(m, s, c, r, df, s.qacc);
      }
      r += 2;
    }
    if (improvement * scale < m.tolerance) { it++; break; }
  }
  bsync(c);
  // ---- finish: row states for the record, qfrc_constraint = J' f (qacc = qacc_smooth + M^-1 J' f is already in s.qacc)
  for (int r = c.tid; r < n; r += c.nt) s.row_state[r] = s.force[r] != 0 ? ST_QUADRATIC : ST_SATISFIED;
  bsync(c);
  mul_JT_force(m, s, c);
  if (c.tid == 0) { s.niter = it; if (it >= m.iterations) s.flags |= FLAG_SOLVER_MAXITER; }
  bsync(c);
}

}  // namespace b2p
