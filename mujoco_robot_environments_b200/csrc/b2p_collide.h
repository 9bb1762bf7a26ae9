#pragma once
#include "b2p_scratch.h"
#ifndef B2P_BB_SMEM
#define B2P_BB_SMEM 1
#endif
#ifndef B2P_BB_SMEM_F32
#define B2P_BB_SMEM_F32 1
#endif
#ifndef B2P_RAW_SMEM
#define B2P_RAW_SMEM 0
#endif
#ifndef B2P_BP_PREFETCH
#define B2P_BP_PREFETCH 1
#endif
#ifndef B2P_NBR_BATCH_F32
#define B2P_NBR_BATCH_F32 4
#endif

namespace b2p {

// ------------------------------------------------------------------------------------------ collision
template <typename T>
struct CObj { const T* pos; const T* mat; const T* size; const T* vert; int type, vadr, vnum, cur; T margin; const int* start; };

// support point in world direction dir (|dir| = 1), inflated by margin/2.  Meshes: hill climbing over the
// hull graph with warm start inside one query (what MuJoCo does for meshes with a graph).  The FIRST call of a query starts
// from the hull's extreme vertex along the nearest of 26 tabulated directions (Globals::geom_start) instead of vertex 0:
// the climb is a chain of dependent global loads, and this cuts it from ~sqrt(nvert) steps to one or two.  On a convex
// hull the climb ends at the same maximum wherever it starts.
// neighbours fetched per batch in the hill climb
template <typename T> struct NbrBatch { static constexpr int value = B2P_NBR_BATCH_F32; };
template <> struct NbrBatch<double> { static constexpr int value = 8; };
// slot of a (unit) direction in the 3x3x3 start table: each component quantised to {-1, 0, +1} at +-0.4
template <typename T> B2P_HD int start_slot(const T* d) {
  return ((d[0] > T(0.4)) - (d[0] < T(-0.4)) + 1) * 9 + ((d[1] > T(0.4)) - (d[1] < T(-0.4)) + 1) * 3 + ((d[2] > T(0.4)) - (d[2] < T(-0.4)) + 1);
}
#ifdef B2P_HOST_COUNTERS
static long g_support_calls = 0, g_climb_steps = 0, g_mpr_calls = 0, g_cand = 0, g_bb_calls = 0, g_steps = 0, g_pair_mpr[1024] = {0}, g_newton_iters = 0, g_newton_solves = 0;
#endif
template <typename T>
B2P_HDN void support(CObj<T>& o, const Globals<T>& g, const T* dir, T* res) {
  T ld[3], lp[3];
#ifdef B2P_HOST_COUNTERS
  g_support_calls++;
#endif
  mulmtv(ld, o.mat, dir);
  if (o.type == B2P_GEOM_BOX) {
    for (int k = 0; k < 3; k++) lp[k] = ld[k] >= 0 ? o.size[k] : -o.size[k];
  } else if (o.type == B2P_GEOM_SPHERE) {
    for (int k = 0; k < 3; k++) lp[k] = o.size[0] * ld[k];
  } else if (o.type == B2P_GEOM_CYLINDER) {   // disc of radius size[0] swept along local z by +- size[1]
    const T rho = sqrt(ld[0] * ld[0] + ld[1] * ld[1]), sc = rho > Num<T>::minval() ? o.size[0] / rho : T(0);
    lp[0] = sc * ld[0]; lp[1] = sc * ld[1]; lp[2] = ld[2] >= 0 ? o.size[1] : -o.size[1];
  } else {
    int v = o.cur;
    if (v < 0) v = o.start[start_slot(ld)];
    T best = dot3(g.mesh_vert + 3 * v, ld);
    // every climb step is two dependent global loads: the vertex's neighbour span, then up to 8 neighbours at once (coordinates
    // inline, independent of each other) -- not a load -> index -> load chain per neighbour
    for (int guard = 0; guard < 4096; guard++) {
      int nb = v; const NbrSpan sp = g.nbr_span[v];
      constexpr int NB = NbrBatch<T>::value;
      for (int q0 = 0; q0 < sp.num; q0 += NB) {
        T dq[NB]; int uq[NB];
#pragma unroll
        for (int j = 0; j < NB; j++) {
          if (q0 + j < sp.num) { const NbrPt<T> pt = g.nbr_pt[sp.adr + q0 + j]; dq[j] = pt.x * ld[0] + pt.y * ld[1] + pt.z * ld[2]; uq[j] = pt.idx; }
          else { dq[j] = best; uq[j] = v; }
        }
#pragma unroll
        for (int j = 0; j < NB; j++) if (dq[j] > best) { best = dq[j]; nb = uq[j]; }
      }
      if (nb == v) break;
      v = nb;
#ifdef B2P_HOST_COUNTERS
      g_climb_steps++;
#endif
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
B2P_HDN bool mpr_penetration(CObj<T>& o1, CObj<T>& o2, const Globals<T>& g, int maxiter, T tol, T& depth, T* dir, T* pos, T* wit = nullptr) {   // wit: separation witness of a query that ends at its first test (SepCache)
  Sup<T> p[4], v4; T d[3], va[3], vb[3];
#ifdef B2P_HOST_COUNTERS
  g_mpr_calls++;
#endif
  for (int k = 0; k < 3; k++) { p[0].v1[k] = o1.pos[k]; p[0].v2[k] = o2.pos[k]; p[0].v[k] = p[0].v1[k] - p[0].v2[k]; }
  if (is_zero(p[0].v[0]) && is_zero(p[0].v[1]) && is_zero(p[0].v[2])) p[0].v[0] += T(1e-5);
  for (int k = 0; k < 3; k++) d[k] = -p[0].v[k];
  normalize3(d);
  mink_support(o1, o2, g, d, p[1]);
  T dt = dot3(p[1].v, d);
  if (is_zero(dt) || dt < 0) {
    if (wit) {   // the objects (inflated by margin / 2 each) are -dt apart along d: d and the two exact support heights (margin taken out)
      T a[3], b[3];
      for (int k = 0; k < 3; k++) { a[k] = p[1].v1[k] - o1.pos[k]; b[k] = p[1].v2[k] - o2.pos[k]; }
      cp3(wit, d); wit[3] = (dot3(a, d) - T(0.5) * o1.margin) + (-dot3(b, d) - T(0.5) * o2.margin); wit[4] = 1;
    }
    return false;
  }
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

// ---- analytic pairs of the sphere / cylinder prop classes (oracle: plane_sphere, plane_cylinder, sphere_sphere, sphere_solid).  The
// normal points from the pair's first geom to its second in the PAIR TABLE's order (MuJoCo re-sorts a pair by geom type: same
// contact with normal and geom1 / geom2 swapped).
template <typename T>
B2P_HD int plane_sphere(const T* ppos, const T* pmat, const T* c, T r, T margin, RawCon<T>& out) {
  const T n[3] = {pmat[2], pmat[5], pmat[8]}, dif[3] = {c[0] - ppos[0], c[1] - ppos[1], c[2] - ppos[2]};
  const T dist = dot3(dif, n) - r;
  if (dist > margin) return 0;
  out.dist[0] = dist; cp3(out.normal, n);
  for (int k = 0; k < 3; k++) out.pos[0][k] = c[k] - n[k] * (r + T(0.5) * dist);
  return 1;
}
template <typename T>
B2P_HD int plane_cylinder(const T* ppos, const T* pmat, const T* c, const T* mat, T r, T h, T margin, RawCon<T>& out) {
  const T n[3] = {pmat[2], pmat[5], pmat[8]}, a[3] = {mat[2], mat[5], mat[8]};
  const T na = dot3(n, a), sg = na > 0 ? T(-1) : T(1);
  T v[3] = {-n[0] + na * a[0], -n[1] + na * a[1], -n[2] + na * a[2]};
  if (normalize3(v) < T(1e-9)) { v[0] = mat[0]; v[1] = mat[3]; v[2] = mat[6]; }
  T u[3]; cross3(u, a, v);
  int cnt = 0;
  cp3(out.normal, n);
  for (int q = 0; q < 4; q++) {
    const T cap = q == 1 ? -sg : sg, ca = q < 2 ? T(1) : T(-0.5), sa = q < 2 ? T(0) : (q == 2 ? T(0.8660254037844387) : T(-0.8660254037844387));
    T pt[3];
    for (int k = 0; k < 3; k++) pt[k] = c[k] + cap * h * a[k] + r * (ca * v[k] + sa * u[k]);
    const T dif[3] = {pt[0] - ppos[0], pt[1] - ppos[1], pt[2] - ppos[2]}, dist = dot3(dif, n);
    if (dist > margin) continue;
    out.dist[cnt] = dist;
    for (int k = 0; k < 3; k++) out.pos[cnt][k] = pt[k] - T(0.5) * dist * n[k];
    cnt++;
  }
  return cnt;
}
template <typename T>
B2P_HD int sphere_sphere(const T* c1, T r1, const T* c2, T r2, T margin, RawCon<T>& out) {
  T d[3] = {c2[0] - c1[0], c2[1] - c1[1], c2[2] - c1[2]};
  const T L = normalize3(d), dist = L - r1 - r2;
  if (dist > margin) return 0;
  out.dist[0] = dist; cp3(out.normal, d);
  for (int k = 0; k < 3; k++) out.pos[0][k] = c1[k] + d[k] * (r1 + T(0.5) * dist);
  return 1;
}
// closest point of a solid (box / cylinder) to the sphere centre; normal from the SOLID to the sphere
template <typename T>
B2P_HDN int sphere_solid(int solid_type, const T* cs, T rs, const T* cb, const T* mb, const T* hs, T margin, RawCon<T>& out) {
  const T dif[3] = {cs[0] - cb[0], cs[1] - cb[1], cs[2] - cb[2]};
  T lc[3], q[3], nl[3] = {0, 0, 0}, depth = 0;
  mulmtv(lc, mb, dif);
  bool inside;
  if (solid_type == B2P_GEOM_BOX) {
    inside = true;
    for (int k = 0; k < 3; k++) { q[k] = lc[k] < -hs[k] ? -hs[k] : (lc[k] > hs[k] ? hs[k] : lc[k]); if (q[k] != lc[k]) inside = false; }
    if (inside) {
      int best = 0; T bd = Num<T>::big();
      for (int k = 0; k < 3; k++) { const T d = hs[k] - fabs(lc[k]); if (d < bd) { bd = d; best = k; } }
      for (int k = 0; k < 3; k++) if (k == best) { nl[k] = lc[k] >= 0 ? T(1) : T(-1); q[k] = nl[k] * hs[k]; }
      depth = -bd;
    }
  } else {
    const T rho = sqrt(lc[0] * lc[0] + lc[1] * lc[1]), r = hs[0], h = hs[1];
    inside = rho < r && fabs(lc[2]) < h;
    if (inside) {
      if (r - rho < h - fabs(lc[2])) { const T ir = rho > Num<T>::minval() ? 1 / rho : T(0); nl[0] = rho > Num<T>::minval() ? lc[0] * ir : T(1); nl[1] = lc[1] * ir; nl[2] = 0; q[0] = nl[0] * r; q[1] = nl[1] * r; q[2] = lc[2]; depth = -(r - rho); }
      else { nl[0] = nl[1] = 0; nl[2] = lc[2] >= 0 ? T(1) : T(-1); q[0] = lc[0]; q[1] = lc[1]; q[2] = nl[2] * h; depth = -(h - fabs(lc[2])); }
    } else {
      const T sc = rho > r ? r / rho : T(1);
      q[0] = lc[0] * sc; q[1] = lc[1] * sc; q[2] = lc[2] < -h ? -h : (lc[2] > h ? h : lc[2]);
    }
  }
  if (!inside) { for (int k = 0; k < 3; k++) nl[k] = lc[k] - q[k]; depth = normalize3(nl); }
  const T dist = depth - rs;
  if (dist > margin) return 0;
  T nw[3], qw[3];
  mulmv(nw, mb, nl); mulmv(qw, mb, q);
  out.dist[0] = dist; cp3(out.normal, nw);
  for (int k = 0; k < 3; k++) out.pos[0][k] = cb[k] + qw[k] + T(0.5) * dist * nw[k];
  return 1;
}

// Contact parameters of pair p.  The host mixed them per pair for the compiled (box) props; a pair that involves a re-classed prop
// (sphere / cylinder: per-scene data) is mixed here by the same rule (App. B item 10: priority wins, else max friction,
// solmix-weighted solref / solimp, max condim, max margin / gap).
template <typename T> struct PairMix { PairParam<T> pp; int condim; };
template <typename T, bool C>
B2P_HDN void pair_mix(const DevModel<T>& m, const Globals<T>& g, const Scratch<T, C>& s, int p, int g1, int g2, PairMix<T>& o) {
  const int pr1 = m.geom_prop[g1], pr2 = m.geom_prop[g2];
  const bool re1 = pr1 >= 0 && s.ptype[pr1] != B2P_GEOM_BOX, re2 = pr2 >= 0 && s.ptype[pr2] != B2P_GEOM_BOX;
  if (!re1 && !re2) { o.pp = g.pair_param[p]; o.condim = 3; return; }
  const GeomRaw<T>& a = re1 ? m.pclass[s.ptype[pr1] == B2P_GEOM_SPHERE ? 0 : 1] : g.geom_raw[g1];
  const GeomRaw<T>& b = re2 ? m.pclass[s.ptype[pr2] == B2P_GEOM_SPHERE ? 0 : 1] : g.geom_raw[g2];
  if (a.priority != b.priority) {
    const GeomRaw<T>& w = a.priority > b.priority ? a : b;
    o.condim = w.condim;
    for (int k = 0; k < 3; k++) o.pp.friction[k] = w.friction[k];
    for (int k = 0; k < 2; k++) o.pp.solref[k] = w.solref[k];
    for (int k = 0; k < 5; k++) o.pp.solimp[k] = w.solimp[k];
  } else {
    o.condim = a.condim > b.condim ? a.condim : b.condim;
    for (int k = 0; k < 3; k++) o.pp.friction[k] = a.friction[k] > b.friction[k] ? a.friction[k] : b.friction[k];
    T mix;
    if (a.solmix >= T(1e-15) && b.solmix >= T(1e-15)) mix = a.solmix / (a.solmix + b.solmix);
    else if (a.solmix < T(1e-15) && b.solmix < T(1e-15)) mix = T(0.5); else mix = a.solmix < T(1e-15) ? T(0) : T(1);
    if (a.solref[0] > 0 && b.solref[0] > 0) for (int k = 0; k < 2; k++) o.pp.solref[k] = mix * a.solref[k] + (1 - mix) * b.solref[k];
    else for (int k = 0; k < 2; k++) o.pp.solref[k] = a.solref[k] < b.solref[k] ? a.solref[k] : b.solref[k];
    for (int k = 0; k < 5; k++) o.pp.solimp[k] = mix * a.solimp[k] + (1 - mix) * b.solimp[k];
  }
  o.pp.margin = a.margin > b.margin ? a.margin : b.margin; o.pp.gap = a.gap > b.gap ? a.gap : b.gap;
}
// detection distance of pair p in this scene
template <typename T, bool C>
B2P_HD T pair_md(const DevModel<T>& m, const Globals<T>& g, const Scratch<T, C>& s, int p, int g1, int g2, bool full_contacts) {
  if (!C || !s.reclassed) {   // max(margin) - max(gap) from the per-geom values in the model constants (shared memory): the same reals the host-mixed table holds
    const T mg1 = m.geom_margin[g1], mg2 = m.geom_margin[g2], gp1 = m.geom_gap[g1], gp2 = m.geom_gap[g2];
    const T mg = mg1 > mg2 ? mg1 : mg2;
    return full_contacts ? mg : mg - (gp1 > gp2 ? gp1 : gp2);
  }
  PairMix<T> mx; pair_mix(m, g, s, p, g1, g2, mx);
  return full_contacts ? mx.pp.margin : mx.pp.margin - mx.pp.gap;
}

// mjc_PlaneConvex for a hull: the support vertex along -normal, then the support vertices of three directions tilted off -normal
// (120 degrees apart), kept when they are new vertices inside the margin: at most 4 contacts (oracle: plane_convex, same constants).
#define B2P_PLANE_TILT 0.05
template <typename T>
B2P_HDN int plane_convex(const T* ppos, const T* pmat, CObj<T>& hull, const Globals<T>& g, T margin, RawCon<T>& out) {
  const T n[3] = {pmat[2], pmat[5], pmat[8]}, t1[3] = {pmat[0], pmat[3], pmat[6]}, t2[3] = {pmat[1], pmat[4], pmat[7]};
  int cnt = 0, seen[4] = {-1, -1, -1, -1};
  cp3(out.normal, n);
  for (int q = 0; q < 4; q++) {
    T dir[3];
    if (q == 0) { for (int k = 0; k < 3; k++) dir[k] = -n[k]; }
    else {
      const T ca = q == 1 ? T(1) : T(-0.5), sa = q == 1 ? T(0) : (q == 2 ? T(0.8660254037844387) : T(-0.8660254037844387));   // cos / sin of 0, 120, 240 degrees
      for (int k = 0; k < 3; k++) dir[k] = -n[k] + T(B2P_PLANE_TILT) * (ca * t1[k] + sa * t2[k]);
      normalize3(dir);
    }
    T w[3];
    support(hull, g, dir, w);
    bool dup = false;
    for (int i = 0; i < 4; i++) if (i < cnt && seen[i] == hull.cur) dup = true;
    if (dup) continue;
    const T dif[3] = {w[0] - ppos[0], w[1] - ppos[1], w[2] - ppos[2]};
    const T dist = dot3(dif, n);
    if (dist > margin) continue;
    for (int i = 0; i < 4; i++) if (i == cnt) seen[i] = hull.cur;
    out.dist[cnt] = dist;
    for (int k = 0; k < 3; k++) out.pos[cnt][k] = w[k] - T(0.5) * dist * n[k];
    cnt++;
  }
  return cnt;
}

// SAT (15 axes) + reference-face clipping / edge-edge point.  Returns #contacts, or -1 when the boxes are
// separated by 0 < s <= margin (caller then runs the margin-inflated convex query: one closest-feature contact).
#define BB_TIE 1e-6
template <typename T>
B2P_HDN int box_box(const T* p1, const T* m1, const T* s1, const T* p2, const T* m2, const T* s2, T margin, RawCon<T>& out, T* clip_scratch = nullptr) {
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
    if (s > best + T(BB_TIE)) { best = s; code = 3 + j; }   // parallel faces tie up to rounding: box 1 keeps the reference face
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
  // clip polygon (a quad cut by 4 planes: at most 8 vertices), ping-pong.  The dynamically indexed arrays would live in local
  // memory; the caller lends 48 reals of dead shared memory per lane when it can (BB_CLIP_REALS).
  T lpoly[8][3], ltmp[8][3]; int np = 4;
  T (*poly)[3] = clip_scratch ? reinterpret_cast<T (*)[3]>(clip_scratch) : lpoly;
  T (*tmp)[3] = clip_scratch ? reinterpret_cast<T (*)[3]>(clip_scratch + 24) : ltmp;
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
      T da = (A[0] - pa[0]) * ax[0] + (A[1] - pa[1]) * ax[1] + (A[2] - pa[2]) * ax[2] - lim;
      T db = (Bp[0] - pa[0]) * ax[0] + (Bp[1] - pa[1]) * ax[1] + (Bp[2] - pa[2]) * ax[2] - lim;
      if (fabs(da) < T(BB_TIE)) da = 0;   // a vertex within a micron of the clip plane lies ON it (aligned boxes: gripper pads, cube on cube)
      if (fabs(db) < T(BB_TIE)) db = 0;
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

// oriented-bounding-box test along the 6 face normals: true when the boxes (centre c, half extents h in the geom frame)
// are more than md apart along one of them.  Hull vertices lie inside their box, so this never drops a real contact.
template <typename T>
B2P_HD bool obb_separated(const T* p1, const T* m1, const T* c1, const T* h1, const T* p2, const T* m2, const T* c2, const T* h2, T md) {
  T w1[3], w2[3], t[3], R[9];
  mulmv(w1, m1, c1); mulmv(w2, m2, c2);
  for (int k = 0; k < 3; k++) t[k] = (p2[k] + w2[k]) - (p1[k] + w1[k]);
  for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) R[3 * i + j] = fabs(m1[i] * m2[j] + m1[3 + i] * m2[3 + j] + m1[6 + i] * m2[6 + j]);
  const T slack = md + T(1e-6);
  for (int i = 0; i < 3; i++) {
    const T ti = t[0] * m1[i] + t[1] * m1[3 + i] + t[2] * m1[6 + i];
    if (fabs(ti) > h1[i] + h2[0] * R[3 * i] + h2[1] * R[3 * i + 1] + h2[2] * R[3 * i + 2] + slack) return true;
  }
  for (int j = 0; j < 3; j++) {
    const T tj = t[0] * m2[j] + t[1] * m2[3 + j] + t[2] * m2[6 + j];
    if (fabs(tj) > h2[j] + h1[0] * R[j] + h1[1] * R[3 + j] + h1[2] * R[6 + j] + slack) return true;
  }
  return false;
}

// Upper bound of a geom's support height (about its frame origin, no margin) along the unit world direction d.  Boxes: exact.  Hulls:
// the 26-DOP bound -- d in the geom frame is split into non-negative multiples of one axis, one face diagonal and one body diagonal of
// its octant, and the support function is sublinear, so the same combination of the tabulated support values is >= the true support.
template <typename T>
B2P_HD T support_bound(const T* mat, const T* half, const T* dop, int type, const T* d) {
  T l[3]; mulmtv(l, mat, d);
  const T a0 = fabs(l[0]), a1 = fabs(l[1]), a2 = fabs(l[2]);
  if (type == B2P_GEOM_BOX) return a0 * half[0] + a1 * half[1] + a2 * half[2];
  if (type == B2P_GEOM_SPHERE) return half[0];
  if (type == B2P_GEOM_CYLINDER) return half[0] * sqrt(l[0] * l[0] + l[1] * l[1]) + a2 * half[2];   // half = bounding half extents (r, r, h)
  const int s0 = l[0] < 0 ? -1 : 1, s1 = l[1] < 0 ? -1 : 1, s2 = l[2] < 0 ? -1 : 1;
  // order the components: hi >= mid >= lo
  int ih = 0, im = 1, il = 2; T hi = a0, mid = a1, lo = a2;
  if (mid > hi) { T t = hi; hi = mid; mid = t; int k = ih; ih = im; im = k; }
  if (lo > mid) { T t = mid; mid = lo; lo = t; int k = im; im = il; il = k; }
  if (mid > hi) { T t = hi; hi = mid; mid = t; int k = ih; ih = im; im = k; }
  const int w[3] = {9, 3, 1}, sg[3] = {s0, s1, s2};
  const int q1 = 13 + sg[ih] * w[ih], q2 = q1 + sg[im] * w[im], q3 = q2 + sg[il] * w[il];
  return (hi - mid) * dop[q1] + (mid - lo) * dop[q2] + lo * dop[q3];
}
// true when the geoms are certainly more than md apart along the line of centres -- exactly the first test of the MPR query
// (mpr_penetration: dt <= 0 after the first Minkowski support), decided here from bounds without walking a hull graph
template <typename T>
B2P_HD bool centre_axis_separated(const T* p1, const T* m1, const T* h1, const T* dop1, int t1, const T* p2, const T* m2, const T* h2, const T* dop2, int t2, T md) {
  T d[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
  const T len = normalize3(d);
  if (len < T(1e-6)) return false;
  const T nd[3] = {-d[0], -d[1], -d[2]};
  const T gap = len - support_bound(m1, h1, dop1, t1, d) - support_bound(m2, h2, dop2, t2, nd) - md;
  return gap > T(256) * Num<T>::eps() * (1 + len);
}

// Separation witnesses (Scratch::sep_*): a hull pair whose MPR query ended at its first test was `gap` apart along the line of
// centres d.  The witness keeps d in BOTH geom frames (dA = RA' d, eB = RB' (-d)) and H = hA(d) + hB(-d), the exact support heights.
// At a later step d' = RA dA is the same direction as seen from geom 1, so hA(d') = hA(d) exactly, and hB(-d') <= hB(-d) +
// rB |d' + RB eB| (a hull of radius rB, the support function is rB-Lipschitz).  If the bound still leaves the pair more than the
// margin apart, the query is skipped: the links of the gripper's four-bar linkage sit millimetres apart for ever, every query on
// them used to walk two hull graphs (chains of dependent L2 loads) only to find that out.  Conservative, so contact sets are unchanged.
template <typename T>
B2P_HD bool witness_separated(const T* w, const T* p1, const T* m1, const T* p2, const T* m2, T rb, T md) {
  T d[3], e[3];
  mulmv(d, m1, w); mulmv(e, m2, w + 3);
  const T err[3] = {d[0] + e[0], d[1] + e[1], d[2] + e[2]};
  const T dif[3] = {p2[0] - p1[0], p2[1] - p1[1], p2[2] - p1[2]};
  const T gap = dot3(dif, d) - w[6] - rb * norm3(err) - md;
  return gap > T(1024) * Num<T>::eps() * (1 + fabs(w[6]));
}

constexpr int BB_CLIP_REALS = 48;   // box_box clip scratch per lane: 2 polygons x 8 vertices x 3
// broad phase (bounding spheres over the static pair table) + narrow phase, deterministic contact order
template <typename T, bool C>
B2P_HDN void collide(const DevModel<T>& m, const Globals<T>& g, Scratch<T, C>& s, const Ctx& c, bool full_contacts, bool active) {
  PROF_T0
  if (active) {
  if (c.tid == 0) { s.ncand = 0; s.ncon = 0; }
  for (int q = c.tid; q < m.ngeom; q += c.nt) {
    const int p = m.geom_prop[q];
    s.grbound[q] = p >= 0 ? ((C && s.ptype[p] == B2P_GEOM_SPHERE) ? s.psize[p][0] : norm3(s.phalf[p])) : m.geom_rbound[q];
  }
  bsync(c);
  // Pass 1 over the static pair table (a packed copy in the model constants, i.e. shared memory: no global loads in this pass):
  // bounding spheres / plane distance only -- a handful of instructions per pair, every lane busy -- compacted into s.cand by ballot.
  int base = 0;
  for (int p0 = 0; p0 < m.npair; p0 += c.nt) {
    const int p = p0 + c.tid; bool flag = false;
    const unsigned pk = p < m.npair ? m.pair_pk[p] : 0u;
    const int g1 = pk & 255u, g2 = pk >> 8;
    // detection distance from the geoms' own margins (the pair's mixed value is max(margin) - max(gap), App. B item 10)
    const T mg1 = m.geom_margin[g1], mg2 = m.geom_margin[g2], gp1 = m.geom_gap[g1], gp2 = m.geom_gap[g2];
    const T md = (mg1 > mg2 ? mg1 : mg2) - (full_contacts ? T(0) : (gp1 > gp2 ? gp1 : gp2));
    if (p < m.npair) {
      const int pr1 = m.geom_prop[g1], pr2 = m.geom_prop[g2];
      if (!((pr1 >= 0 && !s.pactive[pr1]) || (pr2 >= 0 && !s.pactive[pr2]))) {
        T dif[3] = {s.gpos[g2][0] - s.gpos[g1][0], s.gpos[g2][1] - s.gpos[g1][1], s.gpos[g2][2] - s.gpos[g1][2]};
        if (m.geom_type[g1] == B2P_GEOM_PLANE) {   // (md here is the compiled pair's: the box class carries the largest margin of the prop classes, so this stays conservative)
          T nrm[3] = {s.gmat[g1][2], s.gmat[g1][5], s.gmat[g1][8]};
          flag = !(dot3(dif, nrm) - s.grbound[g2] > md);
        } else {
          const T rr = s.grbound[g1] + s.grbound[g2] + md;
          flag = !(dot3(dif, dif) > rr * rr);
        }
      }
    }
    int tot; const int ofs = block_rank(c, flag, tot);
    if (flag && base + ofs < MAXCAND) s.cand[base + ofs] = (short)p;
    base += tot;
  }
  unsigned full = base > MAXCAND ? FLAG_CANDFULL : 0;
  const int n1 = base < MAXCAND ? base : MAXCAND;
  bsync(c);
  // Pass 2 over the survivors, lanes full again: oriented bounding boxes along their 6 face normals, then -- hull pairs only, they
  // would go to MPR -- the query's own first test on support bounds and the separation witnesses of earlier steps.  All three are
  // conservative, so the contact set is exactly that of the unfiltered table.  Compaction is in place (a round reads its 32 entries
  // before it writes, and never writes past what it has read).
  base = 0;
  for (int i0 = 0; i0 < n1; i0 += c.nt) {
    const int i = i0 + c.tid; bool flag = false; int p = 0;
    if (i < n1) {
      p = s.cand[i];
      const unsigned pk = m.pair_pk[p]; const int g1 = pk & 255u, g2 = pk >> 8; const T md = pair_md(m, g, s, p, g1, g2, full_contacts);
      if (s.gtype[g1] == B2P_GEOM_PLANE) flag = true;
      else {
        const int pr1 = m.geom_prop[g1], pr2 = m.geom_prop[g2];
        const T* h1 = pr1 >= 0 ? s.phalf[pr1] : m.geom_size[g1]; const T* h2 = pr2 >= 0 ? s.phalf[pr2] : m.geom_size[g2];
        flag = !obb_separated(s.gpos[g1], s.gmat[g1], m.geom_aabb_c[g1], h1, s.gpos[g2], s.gmat[g2], m.geom_aabb_c[g2], h2, md);
        if (flag && (s.gtype[g1] == B2P_GEOM_MESH || s.gtype[g2] == B2P_GEOM_MESH)) {
          flag = !centre_axis_separated(s.gpos[g1], s.gmat[g1], h1, g.geom_dop + 27 * g1, (int)s.gtype[g1], s.gpos[g2], s.gmat[g2], h2, g.geom_dop + 27 * g2, (int)s.gtype[g2], md);
          if (flag) for (int k = 0; k < SEP_N; k++) if (s.sep_pair[k] == p) { if (witness_separated(s.sep_w[k], s.gpos[g1], s.gmat[g1], s.gpos[g2], s.gmat[g2], s.grbound[g2], md)) flag = false; }
        }
      }
    }
    int tot; const int ofs = block_rank(c, flag, tot);
    bsync(c);
    if (flag) s.cand[base + ofs] = (short)p;
    base += tot;
  }
  if (c.tid == 0 && full) s.flags |= full;
  if (c.tid == 0) s.ncand = base;
  bsync(c);
  }
  PROF_ADD(PF_BROAD)
  csync(c, LK_NARROW);
  if (!active) return;
  const int ncand = s.ncand;
#ifdef B2P_HOST_COUNTERS
  g_cand += ncand; g_steps++;
#endif
  const T tol = m.ccd_tolerance > 16 * Num<T>::eps() ? m.ccd_tolerance : 16 * Num<T>::eps();
  int cbase = 0;
  for (int c0 = 0; c0 < ncand; c0 += c.nt) {
    const int ci = c0 + c.tid; int n = 0; RawCon<T> raw_local; int p = -1; T md = 0; T wit[5]; wit[4] = 0;
    // lanes that will run box_box borrow their clip polygons from the FK / RNE sweep buffers (crb, cvel, cacc, cfrc), which are dead
    // once dynamics_terms has run: slot = rank among this round's box-box lanes, while slots last.  The polygons are dynamically
    // indexed arrays, i.e. local memory otherwise.  Same-box A/B on B200 (tools/gpu_ab.sh, c3, 12 scenes per SM): local 8.06 M, lent
    // 8.27 M env-steps/s; lending the raw contact record as well (B2P_RAW_SMEM) 8.20 M -- it turns every access to it into a generic load.
    int bb_flag = 0;
    if (ci < ncand) { const unsigned pk = m.pair_pk[s.cand[ci]]; bb_flag = s.gtype[pk & 255u] == B2P_GEOM_BOX && s.gtype[pk >> 8] == B2P_GEOM_BOX; }
    int bb_tot; const int bb_slot = block_rank(c, bb_flag != 0, bb_tot);
    constexpr int RAW_REALS = (int)((sizeof(RawCon<T>) + sizeof(T) - 1) / sizeof(T)), SLOT_REALS = BB_CLIP_REALS + RAW_REALS;
    const bool lend = B2P_BB_SMEM && (sizeof(T) == 8 || B2P_BB_SMEM_F32) && bb_flag && bb_slot < (MAXTB * 28) / SLOT_REALS;
    T* clip = lend ? &s.crb[0][0] + bb_slot * SLOT_REALS : nullptr;
    RawCon<T>& raw = (B2P_RAW_SMEM && lend) ? *reinterpret_cast<RawCon<T>*>(clip + BB_CLIP_REALS) : raw_local;
    if (ci < ncand) {
      p = s.cand[ci];
      const unsigned pk = m.pair_pk[p]; const int g1 = pk & 255u, g2 = pk >> 8, t1 = s.gtype[g1], t2 = s.gtype[g2];
      md = pair_md(m, g, s, p, g1, g2, full_contacts);
      const int pr1 = m.geom_prop[g1], pr2 = m.geom_prop[g2];
      const T* sz1 = pr1 >= 0 ? s.psize[pr1] : m.geom_size[g1] + (C && t1 == B2P_GEOM_CYLINDER);    // tree cylinder: (r, r, h) in the model, (r, h) from + 1
      const T* sz2 = pr2 >= 0 ? s.psize[pr2] : m.geom_size[g2] + (C && t2 == B2P_GEOM_CYLINDER);
#if defined(B2P_PROF_NARROW) && defined(__CUDA_ARCH__)
      const long long nt0_ = clock64(); int nkind_ = 0;
#endif
      if (t1 == B2P_GEOM_PLANE) {
        if (t2 == B2P_GEOM_BOX) n = plane_box(s.gpos[g1], s.gmat[g1], s.gpos[g2], s.gmat[g2], sz2, md, raw);
        else if (C && t2 == B2P_GEOM_SPHERE) n = plane_sphere(s.gpos[g1], s.gmat[g1], s.gpos[g2], sz2[0], md, raw);
        else if (C && t2 == B2P_GEOM_CYLINDER) n = plane_cylinder(s.gpos[g1], s.gmat[g1], s.gpos[g2], s.gmat[g2], sz2[0], sz2[1], md, raw);
        else {   // an arm link that reaches the ground behind / beside the table
          CObj<T> h = {s.gpos[g2], s.gmat[g2], sz2, nullptr, t2, m.geom_vertadr[g2], m.geom_vertnum[g2], -1, T(0), g.geom_start + 27 * g2};
          n = plane_convex(s.gpos[g1], s.gmat[g1], h, g, md, raw);
        }
      }
      else {
        bool convex = true;
#if defined(B2P_PROF_NARROW) && defined(__CUDA_ARCH__)
        nkind_ = (t1 == B2P_GEOM_BOX && t2 == B2P_GEOM_BOX) ? 1 : 2;
#endif
#ifdef B2P_HOST_COUNTERS
        if (t1 == B2P_GEOM_BOX && t2 == B2P_GEOM_BOX) g_bb_calls++;
#endif
        if (t1 == B2P_GEOM_BOX && t2 == B2P_GEOM_BOX) { n = box_box(s.gpos[g1], s.gmat[g1], sz1, s.gpos[g2], s.gmat[g2], sz2, md, raw, clip); convex = n < 0; if (n < 0) n = 0; }
        else if (C && (t1 == B2P_GEOM_SPHERE || t2 == B2P_GEOM_SPHERE)) {   // the sphere class against sphere / box / cylinder is analytic; against a hull it goes to MPR
          if (t1 == B2P_GEOM_SPHERE && t2 == B2P_GEOM_SPHERE) { n = sphere_sphere(s.gpos[g1], sz1[0], s.gpos[g2], sz2[0], md, raw); convex = false; }
          else if (t1 == B2P_GEOM_SPHERE && (t2 == B2P_GEOM_BOX || t2 == B2P_GEOM_CYLINDER)) {
            n = sphere_solid(t2, s.gpos[g1], sz1[0], s.gpos[g2], s.gmat[g2], sz2, md, raw); convex = false;
            for (int k = 0; k < 3; k++) raw.normal[k] = -raw.normal[k];      // solid -> sphere is second -> first here
          } else if (t2 == B2P_GEOM_SPHERE && (t1 == B2P_GEOM_BOX || t1 == B2P_GEOM_CYLINDER)) { n = sphere_solid(t1, s.gpos[g2], sz2[0], s.gpos[g1], s.gmat[g1], sz1, md, raw); convex = false; }
        }
#ifdef B2P_HOST_COUNTERS
        if (convex) g_pair_mpr[p & 1023]++;
#endif
        if (convex) {
          CObj<T> a = {s.gpos[g1], s.gmat[g1], sz1, nullptr, t1, m.geom_vertadr[g1], m.geom_vertnum[g1], -1, md, g.geom_start + 27 * g1};
          CObj<T> b = {s.gpos[g2], s.gmat[g2], sz2, nullptr, t2, m.geom_vertadr[g2], m.geom_vertnum[g2], -1, md, g.geom_start + 27 * g2};
          T depth, dir[3], pos[3];
          wit[4] = 0;
#ifdef B2P_SKIP_MPR
          if (false) {
#else
          if (mpr_penetration(a, b, g, m.ccd_iterations, tol, depth, dir, pos, wit) && depth >= 0) {
#endif
            raw.dist[0] = md - depth; cp3(raw.normal, dir); cp3(raw.pos[0], pos); n = 1;
          }
        }
      }
#if defined(B2P_PROF_NARROW) && defined(__CUDA_ARCH__)
      { const unsigned long long dt_ = (unsigned long long)(clock64() - nt0_); atomicAdd(&g_prof[nkind_ == 1 ? PF_OSC : (nkind_ == 2 ? PF_MISC : PF_SMOOTH)], dt_); atomicAdd(&g_prof[nkind_ == 1 ? PF_INTEGRATE : (nkind_ == 2 ? PF_UPDATE : PF_LS)], 1ULL); }
#endif
      int keep = 0;   // drop points not strictly inside the detection margin
      for (int i = 0; i < n; i++) if (raw.dist[i] < md) { if (keep != i) { raw.dist[keep] = raw.dist[i]; cp3(raw.pos[keep], raw.pos[i]); } keep++; }
      n = keep;
    }
    {   // new separation witnesses -> the scene's cache: the pair's own slot if it has one, else the next slots of the ring, in lane order
      bool fresh = wit[4] != 0; int slot = -1;
      if (fresh) for (int k = 0; k < SEP_N; k++) if (s.sep_pair[k] == p) slot = k;
      int nnew; const int rk = block_rank(c, fresh && slot < 0, nnew);
      const bool own = fresh && slot >= 0;
      if (fresh && slot < 0) slot = (s.sep_next + rk) % SEP_N;
      bsync(c);
      for (int phase = 0; phase < 2; phase++) {   // refreshed entries first, then the ring (which may evict any entry): no two lanes write one slot at once
        if (fresh && own == (phase == 0) && (own || rk < SEP_N)) {
          const int g1 = g.pair_g1[p], g2 = g.pair_g2[p];
          const T nd[3] = {-wit[0], -wit[1], -wit[2]};
          s.sep_pair[slot] = (short)p; mulmtv(s.sep_w[slot], s.gmat[g1], wit); mulmtv(s.sep_w[slot] + 3, s.gmat[g2], nd); s.sep_w[slot][6] = wit[3];
        }
        bsync(c);
      }
      if (c.tid == 0 && nnew) s.sep_next = (s.sep_next + (nnew < SEP_N ? nnew : SEP_N)) % SEP_N;
      bsync(c);
    }
    int tot; const int ofs = block_scan(c, n, tot);
    for (int i = 0; i < n; i++) {
      const int k = cbase + ofs + i;
      if (k >= MAXCON) break;
      s.c_dist[k] = raw.dist[i]; cp3(s.c_pos[k], raw.pos[i]); cp3(s.c_frame[k], raw.normal); make_frame(s.c_frame[k]);
      if (!C || !s.reclassed) { const PairParam<T>& pp = g.pair_param[p]; s.c_incl[k] = pp.margin - pp.gap; s.c_fric[k] = pp.friction[0]; s.c_dim[k] = 3; }
      else {
        PairMix<T> mx; pair_mix(m, g, s, p, g.pair_g1[p], g.pair_g2[p], mx);
        s.c_incl[k] = mx.pp.margin - mx.pp.gap; s.c_fric[k] = mx.pp.friction[0]; s.c_dim[k] = (unsigned char)(mx.condim == 6 ? 6 : 3);
        if (mx.condim == 6) { s.w6[k * W6_STRIDE + 21] = mx.pp.friction[1]; s.w6[k * W6_STRIDE + 22] = mx.pp.friction[2]; }
      }
      s.c_pair[k] = (short)p; s.c_row[k] = -1;
    }
    cbase += tot;
  }
  if (c.tid == 0) { s.ncon = cbase < MAXCON ? cbase : MAXCON; if (cbase > MAXCON) s.flags |= FLAG_CONTACTFULL; }
  bsync(c);
  PROF_ADD(PF_NARROW)
}


}  // namespace b2p
