#pragma once
#include "b2p_linalg.h"
#include "b2p_rows.h"

#ifndef B2P_TOL_ULPS
#define B2P_TOL_ULPS 16
#endif

namespace b2p {

// ------------------------------------------------------------------------------------------ Newton solver
template <typename T, bool C>
B2P_HD void mul_M(const DevModel<T>& m, const Scratch<T, C>& s, T* r, const T* v, const Ctx& c) {
  for (int d = c.tid; d < m.nv; d += c.nt) {
    if (d < m.ntd) {   // fixed trip count and four partial sums (the packed triangle is padded to MAXTD rows; slots past ntd multiply zeros of v's padding away)
      T a0 = 0, a1 = 0, a2 = 0, a3 = 0; const int ntd = m.ntd;
#pragma unroll
      for (int j = 0; j < MAXTD; j += 4) {
        a0 += (j < ntd ? s.Mt[j <= d ? tri_idx(d, j) : tri_idx(j, d)] * v[j] : T(0));
        a1 += (j + 1 < ntd ? s.Mt[j + 1 <= d ? tri_idx(d, j + 1) : tri_idx(j + 1, d)] * v[j + 1] : T(0));
        a2 += (j + 2 < ntd ? s.Mt[j + 2 <= d ? tri_idx(d, j + 2) : tri_idx(j + 2, d)] * v[j + 2] : T(0));
        a3 += (j + 3 < ntd ? s.Mt[j + 3 <= d ? tri_idx(d, j + 3) : tri_idx(j + 3, d)] * v[j + 3] : T(0));
      }
      r[d] = (a0 + a1) + (a2 + a3);
    }
    else r[d] = s.Mdiag[d] * v[d];
  }
}
// ---- condim-6 contacts (the cylinder prop class: normal, 2 tangential, torsional, 2 rolling rows).  Rare, so they live in their own
// out-of-line functions with plain loops over the 6 rows, and their 6 x 6 cone Hessian sits in the warp's global w6 area; the
// condim-3 code above / below stays unrolled and in shared memory.
template <typename T> B2P_HD int tri6(int a, int b) { return a >= b ? (a * (a + 1)) / 2 + b : (b * (b + 1)) / 2 + a; }
template <typename T, bool C> B2P_HD void friction5(const Scratch<T, C>& s, int k, T* fr) { const T f0 = s.c_fric[k], ft = s.w6[k * W6_STRIDE + 21], fl = s.w6[k * W6_STRIDE + 22]; fr[0] = f0; fr[1] = f0; fr[2] = ft; fr[3] = fl; fr[4] = fl; }
template <typename T, bool C>
B2P_HDN T cone6_update(Scratch<T, C>& s, int k, int r, bool want_hess, unsigned& chg) {
  const T mu = s.c_mu[k]; T fr[5]; friction5(s, k, fr);
  T U[6], Tt = 0, cost = 0;
  U[0] = s.Jaref[r] * mu;
  for (int i = 1; i < 6; i++) { U[i] = s.Jaref[r + i] * fr[i - 1]; Tt += U[i] * U[i]; }
  Tt = sqrt(Tt); const T N = U[0];
  T* W = s.w6 + k * W6_STRIDE;
  const int old = s.row_state[r];
  if (N >= mu * Tt || (Tt <= 0 && N >= 0)) {
    for (int q = 0; q < 6; q++) { s.force[r + q] = 0; s.row_state[r + q] = ST_SATISFIED; }
    if (want_hess) for (int q = 0; q < 21; q++) W[q] = 0;
  } else if (mu * N + Tt <= 0 || (Tt <= 0 && N < 0)) {
    for (int q = 0; q < 6; q++) { const T jq = s.Jaref[r + q]; s.force[r + q] = -s.D[r + q] * jq; s.row_state[r + q] = ST_QUADRATIC; cost += T(0.5) * s.D[r + q] * jq * jq; }
    if (want_hess) { for (int q = 0; q < 21; q++) W[q] = 0; for (int q = 0; q < 6; q++) W[tri6<T>(q, q)] = s.D[r + q]; }
  } else {
    const T mv = Num<T>::minval(), dn = mu * mu * (1 + mu * mu), Dm = s.D[r] / (dn > mv ? dn : mv), NmT = N - mu * Tt;
    cost += T(0.5) * Dm * NmT * NmT;
    const T fn = -Dm * NmT * mu;
    s.force[r] = fn;
    for (int i = 1; i < 6; i++) s.force[r + i] = -fn / Tt * U[i] * fr[i - 1];
    for (int q = 0; q < 6; q++) s.row_state[r + q] = ST_CONE;
    if (want_hess) {
      T sc[6]; sc[0] = mu; for (int i = 1; i < 6; i++) sc[i] = fr[i - 1];
      const T iT = 1 / Tt, iT2 = iT * iT, iT3 = iT2 * iT;
      W[0] = Dm * sc[0] * sc[0];
      for (int a = 1; a < 6; a++) {
        W[tri6<T>(a, 0)] = -Dm * mu * U[a] * iT * sc[a] * sc[0];
        for (int b = 1; b <= a; b++) W[tri6<T>(a, b)] = (Dm * mu * mu * U[a] * U[b] * iT2 - Dm * mu * NmT * ((a == b ? iT : T(0)) - U[a] * U[b] * iT3)) * sc[a] * sc[b];
      }
    }
  }
  chg |= (unsigned)(old != s.row_state[r] || s.row_state[r] == ST_CONE);
  return cost;
}
// line-search terms of one condim-6 contact at step a: (quadratic q0 q1 q2 when the whole contact is quadratic) or (cone ec e1 e2)
template <typename T, bool C>
B2P_HDN void cone6_ls(const Scratch<T, C>& s, int k, int r, T a, T* q, T* e) {
  const T mu = s.c_mu[k]; T fr[5]; friction5(s, k, fr);
  T t0 = 0, t1 = 0, t2 = 0, UU = 0, UV = 0, VV = 0;
  for (int i = 0; i < 6; i++) { const T D = s.D[r + i], j0 = s.Jaref[r + i], jv = s.Jv[r + i]; t0 += T(0.5) * D * j0 * j0; t1 += D * j0 * jv; t2 += T(0.5) * D * jv * jv; }
  for (int i = 1; i < 6; i++) { const T u = s.Jaref[r + i] * fr[i - 1], v = s.Jv[r + i] * fr[i - 1]; UU += u * u; UV += u * v; VV += v * v; }
  const T U0 = s.Jaref[r] * mu, V0 = s.Jv[r] * mu, N = U0 + a * V0, Tsqr = UU + a * (2 * UV + a * VV);
  if (Tsqr <= 0) { if (N < 0) { q[0] += t0; q[1] += t1; q[2] += t2; } return; }
  const T Tt = sqrt(Tsqr);
  if (N >= mu * Tt) return;
  if (mu * N + Tt <= 0) { q[0] += t0; q[1] += t1; q[2] += t2; return; }
  const T mv = Num<T>::minval(), dn = mu * mu * (1 + mu * mu), Dm = s.D[r] / (dn > mv ? dn : mv);
  const T N1 = V0, T1 = (UV + a * VV) / Tt, T2 = VV / Tt - (UV + a * VV) * T1 / (Tt * Tt), NmT = N - mu * Tt;
  e[0] += T(0.5) * Dm * NmT * NmT; e[1] += Dm * NmT * (N1 - mu * T1); e[2] += Dm * ((N1 - mu * T1) * (N1 - mu * T1) + NmT * (-mu * T2));
}

// per-item (rows of eq/limit, whole contacts) force/state/cost from s.Jaref; returns this thread's cost part
template <typename T, bool C>
B2P_HDN T constraint_update(const DevModel<T>& m, Scratch<T, C>& s, const Ctx& c, bool want_hess, unsigned* dirty = nullptr) {
  // dirty: this thread saw a row change state, or a contact in the cone's middle zone (whose Hessian block moves with the point):
  // when no thread did, the Newton Hessian of the previous iteration -- and its factor in s.H -- is still exact
  T cost = 0; unsigned chg = 0;
  for (int r = c.tid; r < s.nrow_nc; r += c.nt) {   // equality and limit rows come first (MuJoCo's row order): no pass over the contact rows
    const int tp = s.row_type[r];
    const T j = s.Jaref[r];
    const int old = s.row_state[r];
    if (tp == ROW_EQ || j < 0) { s.force[r] = -s.D[r] * j; s.row_state[r] = ST_QUADRATIC; cost += T(0.5) * s.D[r] * j * j; }
    else { s.force[r] = 0; s.row_state[r] = ST_SATISFIED; }
    chg |= (unsigned)(old != s.row_state[r]);
  }
  for (int k = c.tid; k < s.ncon; k += c.nt) {
    const int r = s.c_row[k];
    if (r < 0) continue;
    if (C && s.c_dim[k] == 6) { cost += cone6_update(s, k, r, want_hess, chg); continue; }
    const T mu = s.c_mu[k], f1 = s.c_fric[k], f2 = f1;
    const T j0 = s.Jaref[r], j1 = s.Jaref[r + 1], j2 = s.Jaref[r + 2];
    const T N = j0 * mu, U1 = j1 * f1, U2 = j2 * f2, Tt = sqrt(U1 * U1 + U2 * U2);
    T* W = s.cW[k];
    const int old = s.row_state[r];
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
    chg |= (unsigned)(old != s.row_state[r] || s.row_state[r] == ST_CONE);
  }
  if (dirty) *dirty = chg;
  return cost;
}
// qfrc_con = J' force: one lane per dof, walking only the items of the dof's block
template <typename T, bool C>
B2P_HDN void mul_JT_force(const DevModel<T>& m, Scratch<T, C>& s, const Ctx& c) {
  for (int d = c.tid; d < m.nv; d += c.nt) {
    const int blk = dof_block(m, d), o = d - block_start(m, blk), n = s.nbrow[blk];
    const unsigned char* lst = block_rows(s, blk);
    T a = 0;
    for (int q = 0; q < n; q++) {
      const int r = lst[q], tp = s.row_type[r], nr = tp == ROW_CONTACT ? (C ? s.c_dim[s.row_id[r]] : 3) : ((tp == ROW_EQ && m.eq_type[s.row_id[r]] == B2P_EQ_CONNECT) ? 3 : 1);
      const T* seg = row_seg(m, s, r, blk);
      const int w = block_width(m, s.row_a[r]) + (s.row_b[r] >= 0 ? 6 : 0);
      for (int k = 0; k < nr; k++) { const T f = s.force[r + k]; if (f != 0) a += seg[k * w + o] * f; }
    }
    s.qfrc_con[d] = a;
  }
}
// H = M + J' W J on the packed lower triangle in solver order (props first, tree last), only inside the block
// structure s.sp.  A lane keeps up to EPL matrix elements of ONE block pair in registers and walks the row block's items once: a row's
// type / state / segment pointers are decoded once per lane and item, not once per element (same summation order per element).
template <typename T, bool C, int EPL>
B2P_HD void accum_items(const DevModel<T>& m, const Scratch<T, C>& s, int bi, int bj, const unsigned char* lst, int n, const int* oi, const int* oj, const bool* on, T* h) {
  for (int q = 0; q < n; q++) {
    const int r = lst[q], tp = s.row_type[r];
    const T* sj = bi == bj ? nullptr : row_seg(m, s, r, bj);
    if (bi != bj && !sj) continue;
    const T* si = row_seg(m, s, r, bi);
    if (bi == bj) sj = si;
    const int w = block_width(m, s.row_a[r]) + (s.row_b[r] >= 0 ? 6 : 0);
    if (C && tp == ROW_CONTACT && s.c_dim[s.row_id[r]] == 6) {
      if (s.row_state[r] != ST_SATISFIED) {
        const T* W = s.w6 + s.row_id[r] * W6_STRIDE;
#pragma unroll
        for (int k = 0; k < EPL; k++) if (on[k])
          for (int a = 0; a < 6; a++) { const T ai = si[a * w + oi[k]]; if (ai != 0) for (int b = 0; b < 6; b++) h[k] += W[tri6<T>(a, b)] * ai * sj[b * w + oj[k]]; }
      }
    } else if (tp == ROW_CONTACT) {
      if (s.row_state[r] != ST_SATISFIED) {
        const T* W = s.cW[s.row_id[r]];
        const T W0 = W[0], W1 = W[1], W2 = W[2], W3 = W[3], W4 = W[4], W5 = W[5];
#pragma unroll
        for (int k = 0; k < EPL; k++) if (on[k]) {
          const T a0 = si[oi[k]], a1 = si[w + oi[k]], a2 = si[2 * w + oi[k]], b0 = sj[oj[k]], b1 = sj[w + oj[k]], b2 = sj[2 * w + oj[k]];
          h[k] += W0 * a0 * b0 + W1 * a1 * b1 + W2 * a2 * b2 + W3 * (a0 * b1 + a1 * b0) + W4 * (a0 * b2 + a2 * b0) + W5 * (a1 * b2 + a2 * b1);
        }
      }
    } else {
      const int nr = (tp == ROW_EQ && m.eq_type[s.row_id[r]] == B2P_EQ_CONNECT) ? 3 : 1;
      for (int kk = 0; kk < nr; kk++) if (s.row_state[r + kk] == ST_QUADRATIC) {
        const T D = s.D[r + kk];
#pragma unroll
        for (int k = 0; k < EPL; k++) if (on[k]) h[k] += D * si[kk * w + oi[k]] * sj[kk * w + oj[k]];
      }
    }
  }
}
template <typename T, bool C>
B2P_HDN void assemble_H(const DevModel<T>& m, Scratch<T, C>& s, const Ctx& c) {
  const int np = m.nprop, np6 = 6 * np, ntd = m.ntd;
  // the props' diagonal 6 x 6 blocks (21 elements each) in ONE pass: 8 lanes per prop, 3 elements per lane -- instead of four passes
  // that keep 21 lanes busy each
  const bool merged = c.nt >= 8 * MAXP;
  if (merged) {
    const int p = c.tid >> 3, l8 = c.tid & 7;
    if (p < np) {
      const int bi = p + 1, n = s.nbrow[bi];
      int oi[3], oj[3]; T h[3]; bool on[3];
#pragma unroll
      for (int k = 0; k < 3; k++) {
        const int e = l8 + 8 * k; on[k] = e < 21; oi[k] = 0; oj[k] = 0; h[k] = 0;
        if (on[k]) { const unsigned t = m.tri[e]; oi[k] = t >> 8; oj[k] = t & 255u; h[k] = oi[k] == oj[k] ? s.Mdiag[ntd + 6 * p + oi[k]] : T(0); }
      }
      accum_items<T, C, 3>(m, s, bi, bi, block_rows(s, bi), n, oi, oj, on, h);
#pragma unroll
      for (int k = 0; k < 3; k++) if (on[k]) s.H[tri_idx(6 * p + oi[k], 6 * p + oj[k])] = h[k];
    }
  }
  for (int sbi = 0; sbi <= np; sbi++) for (int sbj = 0; sbj <= sbi; sbj++) {
    if (!sp_coupled(s.sp, sbi, sbj)) continue;
    if (merged && sbi == sbj && sbi < np) continue;
    const int bi = sbi == np ? 0 : sbi + 1, bj = sbj == np ? 0 : sbj + 1;      // dof blocks (0 = tree, 1+p = prop p)
    const int ni = block_width(m, bi), nj = block_width(m, bj), ui0 = sbi == np ? np6 : 6 * sbi, uj0 = sbj == np ? np6 : 6 * sbj;
    const int cnt = sbi == sbj ? (ni * (ni + 1)) >> 1 : ni * nj, n = s.nbrow[bi];
    const unsigned char* lst = block_rows(s, bi);
    constexpr int EPL = 4;
    for (int e0 = 0; e0 < cnt; e0 += EPL * c.nt) {
      int oi[EPL], oj[EPL]; T h[EPL]; bool on[EPL];
#pragma unroll
      for (int k = 0; k < EPL; k++) {
        const int e = e0 + k * c.nt + c.tid; on[k] = e < cnt; oi[k] = 0; oj[k] = 0; h[k] = 0;
        if (on[k]) {
          if (sbi == sbj) { const unsigned t = m.tri[e]; oi[k] = t >> 8; oj[k] = t & 255u; } else { oi[k] = e / nj; oj[k] = e - oi[k] * nj; }
          if (sbi == sbj) h[k] = bi == 0 ? s.Mt[tri_idx(oi[k], oj[k])] : (oi[k] == oj[k] ? s.Mdiag[ntd + 6 * (bi - 1) + oi[k]] : T(0));
        }
      }
      accum_items<T, C, EPL>(m, s, bi, bj, lst, n, oi, oj, on, h);
#pragma unroll
      for (int k = 0; k < EPL; k++) if (on[k]) s.H[tri_idx(ui0 + oi[k], uj0 + oj[k])] = h[k];
    }
  }
}
template <typename T> struct LsPt { T alpha, cost, d1, d2; };
template <typename T, bool C>
B2P_HDN LsPt<T> ls_eval(const DevModel<T>& m, Scratch<T, C>& s, const Ctx& c, T a, const T* qg) {
  T q0 = 0, q1 = 0, q2 = 0, ec = 0, e1 = 0, e2 = 0;
  for (int r = c.tid; r < s.nrow_nc; r += c.nt) {
    const int tp = s.row_type[r];
    const T D = s.D[r], j0 = s.Jaref[r], jv = s.Jv[r];
    if (tp == ROW_EQ || j0 + a * jv < 0) { q0 += T(0.5) * D * j0 * j0; q1 += D * j0 * jv; q2 += T(0.5) * D * jv * jv; }
  }
  for (int k = c.tid; k < s.ncon; k += c.nt) {
    const int r = s.c_row[k];
    if (r < 0) continue;
    if (C && s.c_dim[k] == 6) { T qq[3] = {0, 0, 0}, ee[3] = {0, 0, 0}; cone6_ls(s, k, r, a, qq, ee); q0 += qq[0]; q1 += qq[1]; q2 += qq[2]; ec += ee[0]; e1 += ee[1]; e2 += ee[2]; continue; }
    const T mu = s.c_mu[k], f1 = s.c_fric[k], f2 = f1;
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
  block_sum(c, v);
  LsPt<T> p; p.alpha = a; p.cost = v[0] + a * a * qg[2] + a * qg[1] + qg[0]; p.d1 = v[1] + 2 * a * qg[2] + qg[1]; p.d2 = v[2] + 2 * qg[2];
  if (p.d2 < Num<T>::minval()) p.d2 = Num<T>::minval();
  return p;
}
// Jaref = J*x - aref, Ma = M*x, returns total cost (all threads)
template <typename T, bool C>
B2P_HDN T cost_at(const DevModel<T>& m, Scratch<T, C>& s, const Ctx& c, const T* x) {
  for (int r = c.tid; r < s.nefc; r += c.nt) s.Jaref[r] = row_dot(m, s, r, x) - s.aref[r];
  mul_M(m, s, s.Ma, x, c);
  bsync(c);
  T v[1] = {constraint_update(m, s, c, false)};
  for (int d = c.tid; d < m.nv; d += c.nt) v[0] += T(0.5) * (s.Ma[d] - s.qfrc_smooth[d]) * (x[d] - s.qacc_smooth[d]);
  block_sum(c, v);
  return v[0];
}
// constraint cost alone of the residuals jar (the cost part of constraint_update, same terms in the same order; no force / state writes)
template <typename T, bool C>
B2P_HD T constraint_cost(const Scratch<T, C>& s, const Ctx& c, const T* jar) {
  T cost = 0;
  for (int r = c.tid; r < s.nrow_nc; r += c.nt) { const T j = jar[r]; if (s.row_type[r] == ROW_EQ || j < 0) cost += T(0.5) * s.D[r] * j * j; }
  for (int k = c.tid; k < s.ncon; k += c.nt) {
    const int r = s.c_row[k];
    if (r < 0) continue;
    const T mu = s.c_mu[k], f1 = s.c_fric[k];
    const T N = jar[r] * mu, U1 = jar[r + 1] * f1, U2 = jar[r + 2] * f1, Tt = sqrt(U1 * U1 + U2 * U2);
    if (N >= mu * Tt || (Tt <= 0 && N >= 0)) { }
    else if (mu * N + Tt <= 0 || (Tt <= 0 && N < 0)) { for (int q = 0; q < 3; q++) { const T jq = jar[r + q]; cost += T(0.5) * s.D[r + q] * jq * jq; } }
    else { const T mv = Num<T>::minval(), dn = mu * mu * (1 + mu * mu), Dm = s.D[r] / (dn > mv ? dn : mv), NmT = N - mu * Tt; cost += T(0.5) * Dm * NmT * NmT; }
  }
  return cost;
}
// The warm-start choice in one pass: costs at qacc_smooth and at qacc_warmstart with every row decoded once for both points
// (box-prop instantiation; the class-capable one keeps two cost_at calls).  Leaves s.Jaref / s.Ma describing the chosen point.
template <typename T, bool C>
B2P_HDN bool choose_start(const DevModel<T>& m, Scratch<T, C>& s, const Ctx& c) {
  for (int r = c.tid; r < s.nefc; r += c.nt) { s.Jv[r] = row_dot(m, s, r, s.qacc_smooth) - s.aref[r]; s.Jaref[r] = row_dot(m, s, r, s.qacc_ws) - s.aref[r]; }
  mul_M(m, s, s.Ma, s.qacc_ws, c);       // (the Gauss term at qacc_smooth is zero whatever M qacc_smooth is: its factor x - qacc_smooth is)
  bsync(c);
  T v[2] = {constraint_cost(s, c, s.Jv), constraint_cost(s, c, s.Jaref)};
  for (int d = c.tid; d < m.nv; d += c.nt) v[1] += T(0.5) * (s.Ma[d] - s.qfrc_smooth[d]) * (s.qacc_ws[d] - s.qacc_smooth[d]);
  block_sum(c, v);
  const bool use_ws = v[1] < v[0];
  if (!use_ws) {
    bsync(c);
    for (int r = c.tid; r < s.nefc; r += c.nt) s.Jaref[r] = s.Jv[r];
    mul_M(m, s, s.Ma, s.qacc_smooth, c);
  }
  bsync(c);
  return use_ws;
}
// cost and gradient at s.qacc (s.Jaref / s.Ma consistent with it); also the per-contact cone Hessian weights
template <typename T, bool C>
B2P_HDN T solver_update(const DevModel<T>& m, Scratch<T, C>& s, const Ctx& c, T& gradnorm2, bool& hess_dirty, T& gradscale2) {   // gradscale2: squared size of the terms the gradient is the difference of
  unsigned chg = 0;
  T v[3] = {constraint_update(m, s, c, true, &chg), 0, 0};
  hess_dirty = block_or(c, chg) != 0;
  bsync(c);
  mul_JT_force(m, s, c);
  bsync(c);
  for (int d = c.tid; d < m.nv; d += c.nt) {
    const T gd = s.Ma[d] - s.qfrc_smooth[d] - s.qfrc_con[d];
    s.grad[d] = gd; s.Mgrad[d] = gd; v[1] += gd * gd;
    { const T a1 = s.Ma[d] - s.qfrc_smooth[d], a2 = s.qfrc_con[d]; v[2] += a1 * a1 + a2 * a2; }
    v[0] += T(0.5) * (s.Ma[d] - s.qfrc_smooth[d]) * (s.qacc[d] - s.qacc_smooth[d]);
  }
  block_sum(c, v);
  gradnorm2 = v[1]; gradscale2 = v[2];
  return v[0];
}
// Newton direction: search = -H^-1 grad with H = M + J' W J factored on its block structure
template <typename T, bool C>
B2P_HDN void newton_direction(const DevModel<T>& m, Scratch<T, C>& s, const Ctx& c, bool refactor) {
  PROF_T0
  if (refactor) {   // otherwise s.H still holds the factor of the identical Hessian of the previous iteration
    assemble_H(m, s, c);
    bsync(c);
    for (int u = c.tid; u < m.nv; u += c.nt) s.tmpv[u] = s.H[tri_idx(u, u)];   // original diagonal, by solver index: the pivot floor's reference
    bsync(c);
    PROF_ADD(PF_ASSEMBLE)
    sp_cholesky(s.H, s.sp, m.tri, c, s.tmpv);
    PROF_ADD(PF_CHOL)
  }
  sp_solve(s.H, s.sp, s.Mgrad, 6 * m.nprop, m.ntd, c);
  for (int d = c.tid; d < m.nv; d += c.nt) s.search[d] = -s.Mgrad[d];
  bsync(c);
  PROF_ADD(PF_SOLVE)
}
template <typename T, bool C>
B2P_HDN T line_search(const DevModel<T>& m, Scratch<T, C>& s, const Ctx& c, T scale) {
  T v[4] = {0, 0, 0, 0};
  mul_M(m, s, s.Mv, s.search, c);
  for (int r = c.tid; r < s.nefc; r += c.nt) s.Jv[r] = row_dot(m, s, r, s.search);
  bsync(c);
  for (int d = c.tid; d < m.nv; d += c.nt) {
    const T sd = s.search[d], res = s.Ma[d] - s.qfrc_smooth[d];
    v[0] += sd * sd; v[1] += sd * res; v[2] += T(0.5) * sd * s.Mv[d]; v[3] += T(0.5) * res * (s.qacc[d] - s.qacc_smooth[d]);
  }
  block_sum(c, v);
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
template <typename T, bool C>
B2P_HDN void solve_constraints(const DevModel<T>& m, Scratch<T, C>& s, const Ctx& c, bool active) {
  const int nv = m.nv;
  const bool has = active && s.nefc > 0;
  if (active && !has) {
    for (int d = c.tid; d < nv; d += c.nt) { s.qacc[d] = s.qacc_smooth[d]; s.qfrc_con[d] = 0; }
    if (c.tid == 0) s.niter = 0;
    bsync(c);
  }
  const T scale = 1 / (m.meaninertia * (nv > 1 ? nv : 1));
  T gn2 = 0, gs2 = 0, cost = 0; bool dirty = true;
  PROF_T0
  if (has) {
    // warm start: the cheaper of qacc_warmstart and qacc_smooth.  The warm start wins in all but the first step of a contact, so it is
    // evaluated LAST: s.Jaref / s.Ma then already describe the starting point and the third evaluation is only needed otherwise.
    bool use_ws;
    if constexpr (C) {
      const T cs = cost_at(m, s, c, s.qacc_smooth), cw = cost_at(m, s, c, s.qacc_ws);
      use_ws = cw < cs;
    } else use_ws = choose_start(m, s, c);
    for (int d = c.tid; d < nv; d += c.nt) s.qacc[d] = use_ws ? s.qacc_ws[d] : s.qacc_smooth[d];
    bsync(c);
    if (C && !use_ws) cost_at(m, s, c, s.qacc);
    PROF_ADD(PF_SUB3)
    NANCHK(2, s.Jaref, s.nefc)
    cost = solver_update(m, s, c, gn2, dirty, gs2);
    NANCHK(3, s.grad, nv) NANCHK(4, &s.cW[0][0], 6 * s.ncon)
    dirty = true;   // nothing factored yet in this step
  }
  PROF_ADD(PF_SOLVE_INIT)
  int it = 0;
  // MuJoCo factors the Hessian before testing for termination; the factor at the converged point is never used, so the
  // test comes first here (identical iterates, one factorisation less per step).  In lockstep mode the CTA's warps start
  // every Newton iteration together (a scene that has converged idles through the iterations its neighbours still need).
  bool run = has && it < m.iterations;
  while (csync_or(c, LK_NEWTON, run)) {
    if (!run) continue;
    newton_direction(m, s, c, dirty);
    NANCHK(5 + 16 * it, s.search, nv)
    PROF_RESET
    const T alpha = line_search(m, s, c, scale);
    { T al_[1] = {alpha}; NANCHK(6 + 16 * it, al_, 1) }
    PROF_ADD(PF_LS)
    if (alpha == 0) { run = false; continue; }
    for (int d = c.tid; d < nv; d += c.nt) { s.qacc[d] += alpha * s.search[d]; s.Ma[d] += alpha * s.Mv[d]; }
    for (int r = c.tid; r < s.nefc; r += c.nt) s.Jaref[r] += alpha * s.Jv[r];
    bsync(c);
    const T old = cost;
    cost = solver_update(m, s, c, gn2, dirty, gs2);
    NANCHK(7 + 16 * it, s.grad, nv) NANCHK(8 + 16 * it, &s.cW[0][0], 6 * s.ncon)
    PROF_ADD(PF_UPDATE)
    const T improvement = scale * (old - cost), gradient = scale * sqrt(gn2);
    it++;
    // MuJoCo's test, with the tolerance floored at what the arithmetic can resolve: the cost is a sum of O(|cost|) terms and the
    // gradient a difference of O(|M a - f_smooth| + |J'f|) terms, so anything below a few ulp of those is rounding noise.  In fp64
    // the floors (~1e-14) are far below the tolerance (1e-8) and nothing changes; in fp32 (~1e-6) the gradient after an exact Newton
    // step could never pass the 1e-8 test, and the solver spent a second iteration on almost every step only to see no improvement
    // (2.05 iterations per step against 1.02 in fp64 on the same scenes; B2P_TOL_ULPS = 0 restores the plain test).
    const T fl_c = T(B2P_TOL_ULPS) * Num<T>::eps() * scale * fabs(cost), fl_g = T(B2P_TOL_ULPS) * Num<T>::eps() * scale * sqrt(gs2);
    const T tol_eff = m.tolerance > fl_c ? m.tolerance : fl_c, tol_g = m.tolerance > fl_g ? m.tolerance : fl_g;
    if (improvement < tol_eff || gradient < tol_g || it >= m.iterations) run = false;
  }
  if (has) {
#ifdef B2P_HOST_COUNTERS
    g_newton_iters += it; g_newton_solves++;
#endif
    PROF_COUNT(PF_NITER, it)
    if (c.tid == 0) { s.niter = it; if (it >= m.iterations) s.flags |= FLAG_SOLVER_MAXITER; }
    bsync(c);
  }
}


}  // namespace b2p
