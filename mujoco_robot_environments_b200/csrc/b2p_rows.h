#pragma once
#include "b2p_linalg.h"
#include "b2p_collide.h"

namespace b2p {

// ------------------------------------------------------------------------------------------ constraints
template <typename T>
B2P_HDN T impedance_pow(T x, T mid, T pw) { return x <= mid ? pow(x, pw) / pow(mid, pw - 1) : 1 - pow(1 - x, pw) / pow(1 - mid, pw - 1); }
template <typename T>
B2P_HD T impedance(const T* si, T pos, T margin) {
  const T d0 = clampT(si[0], T(0.0001), T(0.9999)), d1 = clampT(si[1], T(0.0001), T(0.9999)), w = si[2] > 0 ? si[2] : T(0);
  const T mid = clampT(si[3], T(0.0001), T(0.9999)), pw = si[4] > 1 ? si[4] : T(1);
  if (d0 == d1 || w <= Num<T>::minval()) return T(0.5) * (d0 + d1);
  const T x = fabs((pos - margin) / w); T y;
  if (x >= 1) return d1;
  if (x <= 0) return d0;
  if (pw == 1) y = x;
  else if (pw == 2) y = x <= mid ? x * x / mid : 1 - (1 - x) * (1 - x) / (1 - mid);   // MuJoCo's default power: no pow() (x * x IS the correctly rounded pow(x, 2))
  else y = impedance_pow(x, mid, pw);
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
template <typename T, bool C>
B2P_HD void add_jac(const DevModel<T>& m, const Scratch<T, C>& s, T* row, int seg0, int body, const T* pt, const T* ax, T sign) {
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
// ... and of its angular velocity (rows 3..5 of a condim-6 contact: torsional and rolling friction)
template <typename T, bool C>
B2P_HD void add_jac_rot(const DevModel<T>& m, const Scratch<T, C>& s, T* row, int seg0, int body, const T* ax, T sign) {
  if (body == 0) return;
  if (body >= PROP_BODY0) {
    const int p = body - PROP_BODY0;
    for (int k = 0; k < 3; k++) row[seg0 + 3 + k] += sign * (ax[0] * s.pmat[p][k] + ax[1] * s.pmat[p][3 + k] + ax[2] * s.pmat[p][6 + k]);
  } else {
    const unsigned mask = m.anc_mask[body];
    for (int d = 0; d < m.ntd; d++) if (mask & (1u << d)) row[seg0 + d] += sign * dot3(ax, s.cdof[d]);
  }
}
template <typename T, bool C>
B2P_HD T row_dot(const DevModel<T>& m, const Scratch<T, C>& s, int r, const T* x) {
  const int a = s.row_a[r], b = s.row_b[r], wa = block_width(m, a), sa = block_start(m, a);
  const T* j = jrow(s, r);
  // fixed trip counts (16 tree slots with a guard, 6 prop slots) and four partial sums: the loop unrolls and the FMA chain is 4x shorter
  T v0 = 0, v1 = 0, v2 = 0, v3 = 0;
  if (a == 0) {
#pragma unroll
    for (int k = 0; k < MAXTD; k += 4) {
      // both factors are guarded, not just the Jacobian entry: with fewer than 16 dofs in the model (the drawing scene has 7) x[k] past
      // the row is whatever the shared memory held, and 0 * NaN is NaN.  (Selecting each factor keeps the fused multiply-add, i.e. the
      // bits, of the one-guard form; selecting the product would round it separately.)
      v0 += (k < wa ? j[k] : T(0)) * (k < wa ? x[k] : T(0)); v1 += (k + 1 < wa ? j[k + 1] : T(0)) * (k + 1 < wa ? x[k + 1] : T(0));
      v2 += (k + 2 < wa ? j[k + 2] : T(0)) * (k + 2 < wa ? x[k + 2] : T(0)); v3 += (k + 3 < wa ? j[k + 3] : T(0)) * (k + 3 < wa ? x[k + 3] : T(0));
    }
  } else {
    v0 = j[0] * x[sa] + j[4] * x[sa + 4]; v1 = j[1] * x[sa + 1] + j[5] * x[sa + 5]; v2 = j[2] * x[sa + 2]; v3 = j[3] * x[sa + 3];
  }
  if (b >= 0) { const int sb = block_start(m, b); const T* jb = j + wa; v0 += jb[0] * x[sb] + jb[4] * x[sb + 4]; v1 += jb[1] * x[sb + 1] + jb[5] * x[sb + 5]; v2 += jb[2] * x[sb + 2]; v3 += jb[3] * x[sb + 3]; }
  return (v0 + v1) + (v2 + v3);
}
template <typename T, bool C> B2P_HD T row_vel(const DevModel<T>& m, const Scratch<T, C>& s, int r) { return row_dot(m, s, r, s.qvel); }

// mj_makeConstraint + mj_referenceConstraint: equality -> limits -> contacts, rows complete incl. aref.
// Pass 1, one lane per ITEM (an equality, a joint limit side, or a contact = 3 rows): which items exist, their row numbers and
// packed-Jacobian offsets from warp prefix sums (so the row order is deterministic and identical to MuJoCo's), the per-block item
// lists (brow) and which blocks share a row (cpl_pairs) -- the sparsity the Newton solver exploits.
// Pass 2, one lane per ROW: the Jacobian row, impedance / regularisation and the reference acceleration.  The rows of one item are
// independent of each other, and a row on the tree is a 15-dof loop per body: one lane per item made the two connect equalities
// (3 rows x 2 tree bodies each) the critical path of the whole stage.
template <typename T, bool C>
B2P_HDN void make_rows(const DevModel<T>& m, const Globals<T>& g, Scratch<T, C>& s, const Ctx& c) {
  PROF_T0
  // joint-limit sides that are inside their margin (rarely more than one of the 30): found up front, so that the item list below is
  // equalities + ACTIVE limit sides + contacts and usually fits one round of 32 lanes
  unsigned limmask = 0;
  for (int i = c.tid; i < 2 * m.ntd; i += c.nt) {
    const int j = i >> 1, side = (i & 1) ? 1 : -1;
    if (m.jnt_limited[j] && side * (m.jnt_range[j][(side + 1) / 2] - s.qpos[j]) < m.jnt_margin[j]) limmask |= 1u << i;
  }
  limmask = block_or(c, limmask);
  int nlim = 0; for (unsigned t = limmask; t; t &= t - 1) nlim++;
  const int nitem = m.neq + nlim + s.ncon, nblk = m.nprop + 1;
  int base = 0, jbase = 0, first_drop = Pool<T>::EFC + 1; unsigned pairs = 0;
  int nb[SP_MAXBLK];
  for (int b = 0; b < SP_MAXBLK; b++) nb[b] = 0;
  for (int i0 = 0; i0 < nitem; i0 += c.nt) {
    const int it = i0 + c.tid; int nr = 0, kind = -1, id = 0, ba = 0, bb = -1;
    if (it < nitem) {
      if (it < m.neq) { kind = ROW_EQ; id = it; nr = m.eq_type[it] == B2P_EQ_CONNECT ? 3 : 1; }
      else if (it < m.neq + nlim) {
        unsigned t = limmask; for (int k = it - m.neq; k > 0; k--) t &= t - 1;      // the (it - neq)-th set bit, ascending: MuJoCo's order
        int i = 0; while (!((t >> i) & 1u)) i++;
        kind = ROW_LIMIT; id = (i >> 1) * 2 + (i & 1); nr = 1;
      } else {
        id = it - m.neq - nlim;
        if (s.c_dist[id] < s.c_incl[id]) {
          kind = ROW_CONTACT; nr = C ? s.c_dim[id] : 3;
          if (C && m.cone == B2P_CONE_PYRAMIDAL) { kind = ROW_PYR; nr = 4; }      // 2 (condim - 1) pyramid edges, condim 3
          // segment assignment: tree -> A; a prop -> B if A is taken (by the tree or by the other prop)
          const int p = s.c_pair[id], b1 = m.geom_body[g.pair_g1[p]], b2 = m.geom_body[g.pair_g2[p]];
          const bool t1 = b1 > 0 && b1 < PROP_BODY0, t2 = b2 > 0 && b2 < PROP_BODY0;
          ba = (t1 || t2) ? 0 : -1;
          if (b1 >= PROP_BODY0) { if (ba < 0) ba = 1 + b1 - PROP_BODY0; else bb = 1 + b1 - PROP_BODY0; }
          if (b2 >= PROP_BODY0) { if (ba < 0) ba = 1 + b2 - PROP_BODY0; else bb = 1 + b2 - PROP_BODY0; }
          if (ba < 0) { kind = -1; nr = 0; }     // static-static pairs never reach here (filtered at compile time)
        }
      }
    }
    const int w = nr > 0 ? block_width(m, ba) + (bb >= 0 ? 6 : 0) : 0;
    int tot, jtot; const int ofs = block_scan(c, nr, tot), jofs = block_scan(c, nr * w, jtot);
    const int r0 = base + ofs, j0 = jbase + jofs;
    if (nr > 0 && (r0 + nr > Pool<T>::EFC || j0 + nr * w > Pool<T>::JP)) {   // dropped items always form a suffix (offsets include dropped ones)
      if (r0 < first_drop) first_drop = r0;
      if (kind == ROW_CONTACT || kind == ROW_PYR) s.c_row[id] = -1;
      s.flags |= FLAG_EFCFULL; nr = 0;
    }
    if (nr > 0) {
      const bool in_smem = j0 + nr * w <= Pool<T>::JPS;   // an item's rows never straddle the two pools
      for (int q = 0; q < nr; q++) {
        s.row_type[r0 + q] = (signed char)kind; s.row_id[r0 + q] = (short)((C && kind == ROW_PYR) ? 4 * id + q : id); s.row_a[r0 + q] = (signed char)ba; s.row_b[r0 + q] = (signed char)bb;
        s.joff[r0 + q] = (short)(in_smem ? j0 + q * w : -(j0 + q * w) - 1);
      }
      if (kind == ROW_CONTACT) s.c_row[id] = (short)r0;
      else if (C && kind == ROW_PYR) s.c_row[id] = -1;      // no cone block: the edges are scalar rows of their own
      else if (kind == ROW_EQ) s.eq_row0[id] = (short)r0;
      if (bb >= 0) {   // solver blocks: prop p -> p, tree -> nprop
        const int sa = ba == 0 ? m.nprop : ba - 1, sb = bb - 1, hi = sa > sb ? sa : sb, lo = sa > sb ? sb : sa;
        pairs |= 1u << (hi * SP_MAXBLK + lo);
      }
    }
    if (C && m.cone == B2P_CONE_PYRAMIDAL) {   // a pyramid edge is an item of its own in the lists: an item's walkers take a non-contact row for one row
      for (int b = 0; b < nblk; b++) {
        const bool f = nr > 0 && (ba == b || bb == b); const int cnt = f ? (kind == ROW_PYR ? nr : 1) : 0;
        int nt; const int rk = block_scan(c, cnt, nt);
        const int cap = b == 0 ? Pool<T>::EFC : MAXCON;
        for (int q = 0; q < cnt; q++) {
          const int at = nb[b] + rk + q;
          if (at >= cap) { s.flags |= FLAG_EFCFULL; continue; }
          if (b == 0) s.brow0[at] = (unsigned char)(r0 + q); else s.browp[b - 1][at] = (unsigned char)(r0 + q);
        }
        nb[b] += nt; if (nb[b] > cap) nb[b] = cap;
      }
    } else
    for (int b = 0; b < nblk; b++) {   // per-block item lists, in row order
      int nt; const bool f = nr > 0 && (ba == b || bb == b); const int rk = block_rank(c, f, nt);
      if (f) { if (b == 0) s.brow0[nb[0] + rk] = (unsigned char)r0; else s.browp[b - 1][nb[b] + rk] = (unsigned char)r0; }
      nb[b] += nt;
    }
    base += tot; jbase += jtot;
  }
  first_drop = block_min(c, first_drop);
  pairs = block_or(c, pairs);
  // the block structure depends on the coupling bits only: it is kept from step to step and rebuilt (by one lane, serially) only
  // when a contact starts or stops coupling two blocks
  const bool rebuild = s.sp.nblk != nblk || s.cpl_pairs != pairs;
  const int nefc = base < first_drop ? base : first_drop;
  bsync(c);
  if (c.tid == 0) {
    s.nefc = nefc; s.jused = jbase; s.cpl_pairs = pairs;
    { int nc = nefc; for (int r = 0; r < nefc; r++) if (s.row_type[r] == ROW_CONTACT) { nc = r; break; } s.nrow_nc = nc; }
    for (int b = 0; b < SP_MAXBLK; b++) s.nbrow[b] = nb[b];
  }
  PROF_ADD(PF_SUB0)
  // Every row kind is brought to ONE shape first -- (body, point, sign) twice, an axis, impedance parameters -- so that the warp runs the
  // expensive part (two point Jacobians, the row's velocity) once for all its lanes instead of once per kind: a round used to cost
  // connect-equality path + contact path + ..., one after the other, because divergent branches serialise.
  for (int r = c.tid; r < nefc; r += c.nt) {
    const int kind = s.row_type[r], id = s.row_id[r], ba = s.row_a[r], bb = s.row_b[r];
    const int w = block_width(m, ba) + (bb >= 0 ? 6 : 0);
    const int jo = s.joff[r];
    T* J0 = jo >= 0 ? s.Jp + jo : s.jovf + (-jo - 1);
    for (int k = 0; k < w; k++) J0[k] = 0;
    // body_p / pt_p / seg_p get +ax, body_n / pt_n / seg_n get -ax (both skipped for the scalar rows, which write their entries directly)
    int body_p = 0, body_n = 0, seg_p = 0, seg_n = 0; bool rot = false;
    T pt_p[3] = {0, 0, 0}, pt_n[3] = {0, 0, 0}, ax[3] = {0, 0, 0};
    // per-kind inputs of the ONE mj_makeImpedance evaluation below (the kinds only differ in which parameters they hand it)
    T posterm = 0, pos_imp = 0, margin = 0, diag = 1, rscale = 1; bool fric = false; int mu_of = -1;
    const T* solref = m.eq_solref[0]; const T* solimp = m.eq_solimp[0]; PairMix<T> mx;
    if (kind == ROW_EQ && m.eq_type[id] == B2P_EQ_CONNECT) {
      const int q = r - s.eq_row0[id];
      const int b1 = m.eq_obj1[id], b2 = m.eq_obj2[id]; T res[3];
      mulmv(pt_p, s.xmat[b1], m.eq_data[id]); mulmv(pt_n, s.xmat[b2], m.eq_data[id] + 3);
      for (int k = 0; k < 3; k++) { pt_p[k] += s.xpos[b1][k]; pt_n[k] += s.xpos[b2][k]; res[k] = pt_p[k] - pt_n[k]; }
      diag = m.tb_invweight0[b1][0] + m.tb_invweight0[b2][0]; pos_imp = norm3(res);
      ax[0] = T(q == 0); ax[1] = T(q == 1); ax[2] = T(q == 2);
      body_p = b1; body_n = b2;
      solref = m.eq_solref[id]; solimp = m.eq_solimp[id];
      posterm = q == 0 ? res[0] : (q == 1 ? res[1] : res[2]);
    } else if (kind == ROW_EQ) {
      const int j1 = m.eq_obj1[id], j2 = m.eq_obj2[id]; const T* cf = m.eq_data[id];
      const T x = s.qpos[j2] - m.qpos0[j2];
      posterm = s.qpos[j1] - m.qpos0[j1] - (cf[0] + cf[1] * x);
      J0[j1] = 1; J0[j2] = -cf[1];
      solref = m.eq_solref[id]; solimp = m.eq_solimp[id]; pos_imp = posterm; diag = m.dof_invweight0[j1] + m.dof_invweight0[j2];
    } else if (kind == ROW_LIMIT) {
      const int j = id >> 1, side = (id & 1) ? 1 : -1;
      const T dist = side * (m.jnt_range[j][(side + 1) / 2] - s.qpos[j]);
      J0[j] = T(-side);
      solref = m.jnt_solref[j]; solimp = m.jnt_solimp[j]; pos_imp = dist; margin = m.jnt_margin[j]; diag = m.dof_invweight0[j];
      posterm = dist - margin;
    } else if (C && kind == ROW_PYR) {
      // One edge of MuJoCo's pyramidal cone (mj_instantiateContact / mj_makeImpedance): J = J_normal +- mu J_tangent_k -- the point
      // Jacobian is linear in the axis, so the edge is a "normal row" along n +- mu t_k; it carries the contact's distance, so K, B and
      // the impedance are the normal's; diagApprox = tran (1 + mu^2); every edge then gets R_py = 2 mu_reg^2 R_first (below).
      const int cid = id >> 2, q = id & 3, k = 1 + (q >> 1);
      const int p = s.c_pair[cid], g1 = g.pair_g1[p], g2 = g.pair_g2[p], b1 = m.geom_body[g1], b2 = m.geom_body[g2];
      if (!s.reclassed) mx.pp = g.pair_param[p]; else pair_mix(m, g, s, p, g1, g2, mx);
      seg_n = (b1 >= PROP_BODY0 && 1 + b1 - PROP_BODY0 == bb) ? block_width(m, ba) : 0; seg_p = (b2 >= PROP_BODY0 && 1 + b2 - PROP_BODY0 == bb) ? block_width(m, ba) : 0;
      const T iw1 = b1 >= PROP_BODY0 ? s.pinvw[b1 - PROP_BODY0][0] : m.geom_invweight0[g1][0], iw2 = b2 >= PROP_BODY0 ? s.pinvw[b2 - PROP_BODY0][0] : m.geom_invweight0[g2][0];
      const T mu = s.c_fric[cid], sg = (q & 1) ? T(-1) : T(1);
      solref = mx.pp.solref; solimp = mx.pp.solimp; pos_imp = s.c_dist[cid]; margin = s.c_incl[cid];
      diag = (iw1 + iw2) + mu * mu * (iw1 + iw2);                                   // of the FIRST edge (all edges of a condim-3 contact have this value)
      const T* fr = s.c_frame[cid];
      for (int e = 0; e < 3; e++) ax[e] = fr[e] + sg * mu * fr[3 * k + e];
      cp3(pt_p, s.c_pos[cid]); cp3(pt_n, s.c_pos[cid]);
      body_p = b2; body_n = b1;
      posterm = s.c_dist[cid] - s.c_incl[cid];
      rscale = T(-1);                                                               // marks the edge for the R_py step below
      mu_of = -1 - cid;
    } else {
      const int q = r - s.c_row[id];
      const int p = s.c_pair[id], g1 = g.pair_g1[p], g2 = g.pair_g2[p], b1 = m.geom_body[g1], b2 = m.geom_body[g2];
      if (!C || !s.reclassed) mx.pp = g.pair_param[p]; else pair_mix(m, g, s, p, g1, g2, mx);
      // segment of each body inside the packed row: the tree (or the first prop) heads it, a second block follows
      seg_n = (b1 >= PROP_BODY0 && 1 + b1 - PROP_BODY0 == bb) ? block_width(m, ba) : 0; seg_p = (b2 >= PROP_BODY0 && 1 + b2 - PROP_BODY0 == bb) ? block_width(m, ba) : 0;
      const T iw1 = b1 >= PROP_BODY0 ? s.pinvw[b1 - PROP_BODY0][0] : m.geom_invweight0[g1][0], iw2 = b2 >= PROP_BODY0 ? s.pinvw[b2 - PROP_BODY0][0] : m.geom_invweight0[g2][0];
      solref = mx.pp.solref; solimp = mx.pp.solimp; pos_imp = s.c_dist[id]; margin = s.c_incl[id]; diag = iw1 + iw2;
      cp3(ax, s.c_frame[id] + 3 * (q < 3 ? q : q - 3)); cp3(pt_p, s.c_pos[id]); cp3(pt_n, s.c_pos[id]);
      body_p = b2; body_n = b1; rot = C && q >= 3;
      if (q == 0) { posterm = s.c_dist[id] - s.c_incl[id]; mu_of = id; }
      else {
        fric = true;
        if (C && q >= 3) {   // App. B item 9: R_k = R_1 mu_0^2 / mu_k^2 for the torsional (k = 3) and rolling (k = 4, 5) rows
          const T fk = s.w6[id * W6_STRIDE + (q == 3 ? 21 : 22)], f0 = s.c_fric[id];
          rscale = f0 * f0 / (fk * fk);
        }
      }
    }
    T K, B, imp, R;
    row_params(m, solref, solimp, pos_imp, margin, diag, false, K, B, imp, R);
    {
      const T mv = Num<T>::minval(), R1 = R / (m.impratio > mv ? m.impratio : mv);
      if (mu_of >= 0) s.c_mu[mu_of] = s.c_fric[mu_of] * sqrt(R1 / R);
      if (fric) { K = 0; R = (C && rot) ? R1 * rscale : R1; }
      if (C && rscale < 0) { const T mur = s.c_fric[-1 - mu_of] * sqrt(R1 / R); R = 2 * mur * mur * R; }      // pyramid edge
    }
    if (C && rot) { add_jac_rot(m, s, J0, seg_p, body_p, ax, T(1)); add_jac_rot(m, s, J0, seg_n, body_n, ax, T(-1)); }
    else { add_jac(m, s, J0, seg_p, body_p, pt_p, ax, T(1)); add_jac(m, s, J0, seg_n, body_n, pt_n, ax, T(-1)); }   // body 0: nothing to add
    s.D[r] = 1 / R; s.aref[r] = -B * row_vel(m, s, r) - K * imp * posterm;
  }
  bsync(c);
  PROF_ADD(PF_SUB1)
  if (rebuild) sp_build(s.sp, m.nprop, m.ntd, pairs, c);
  PROF_ADD(PF_SUB2)
}

}  // namespace b2p
