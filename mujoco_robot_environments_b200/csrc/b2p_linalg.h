// Warp-level block-sparse Cholesky on packed lower-triangular storage.
//
// The Newton Hessian H = M + J' W J of a rearrangement scene is block structured: one dense tree block (arm +
// gripper, <= 15 dofs) and one 6x6 block per free prop, coupled only where a constraint row touches two blocks
// (gripper pad <-> prop, prop <-> prop).  At rest on the table every prop block is decoupled, so factoring the
// dense 39x39 matrix would waste ~10x the work.  Solver order puts the props FIRST and the tree LAST, so
// eliminating a prop never creates fill between the tree and an unrelated prop.
//
// SpStruct lists, per block, the rows that are structurally non-zero below its diagonal (own rows, then the rows of
// later coupled blocks) and the columns of earlier coupled blocks; the factorisation, the triangular solves and
// the Hessian assembly all walk those lists, one lane per matrix element, with a __syncwarp() between phases.
#pragma once
#include "b2p_scratch.h"

namespace b2p {

B2P_HD int tri_idx(int i, int j) { return ((i * (i + 1)) >> 1) + j; }   // i >= j

// dof index of solver index u (props first, tree last); np6 = 6 * nprop for the full system, 0 for tree-only
B2P_HD int sp_dof(int u, int np6, int ntd) { return u < np6 ? ntd + u : u - np6; }

// fill-in closure + list construction from the direct coupling bits (cpl[b] = mask of LATER blocks coupled to b).
// Block ids here are solver blocks: 0..np-1 props, np = tree.  Executed redundantly by every lane, result in smem.
B2P_HD void sp_build(SpStruct& st, int nprop, int ntd, unsigned direct_pairs, const Ctx& c) {
  // direct_pairs bit (i * SP_MAXBLK + j), i > j  <=>  blocks i and j share a constraint row
  if (c.tid == 0) {
    const int nb = nprop + 1;
    unsigned char later[SP_MAXBLK];
    for (int b = 0; b < nb; b++) { later[b] = 0; for (int i = b + 1; i < nb; i++) if (direct_pairs & (1u << (i * SP_MAXBLK + b))) later[b] |= (unsigned char)(1u << i); }
    for (int b = 0; b < nb; b++)       // eliminating b couples every pair of its later neighbours
      for (int i = b + 1; i < nb; i++) if (later[b] & (1u << i)) later[i] |= (unsigned char)(later[b] & ~((2u << i) - 1u));
    st.nblk = nb;
    for (int b = 0; b < nb; b++) {
      const int start = b < nprop ? 6 * b : 6 * nprop, len = b < nprop ? 6 : ntd;
      int n = 0;
      for (int k = 0; k < len; k++) st.rows[b][n++] = (unsigned char)(start + k);
      for (int i = b + 1; i < nb; i++) if (later[b] & (1u << i)) {
        const int s2 = i < nprop ? 6 * i : 6 * nprop, l2 = i < nprop ? 6 : ntd;
        for (int k = 0; k < l2; k++) st.rows[b][n++] = (unsigned char)(s2 + k);
      }
      int np = 0;
      for (int j = 0; j < b; j++) if (later[j] & (1u << b)) for (int k = 0; k < 6; k++) st.prev[b][np++] = (unsigned char)(6 * j + k);
      st.blk[b].start = (unsigned char)start; st.blk[b].len = (unsigned char)len; st.blk[b].nrows = (unsigned char)n; st.blk[b].nprev = (unsigned char)np;
      st.later[b] = later[b];
    }
  }
  bsync(c);
}
// dense single-block structure (tree-only systems: M and M - h*dF/dv)
B2P_HD void sp_build_dense(SpStruct& st, int n, const Ctx& c) {
  if (c.tid == 0) {
    st.nblk = 1; st.later[0] = 0;
    for (int k = 0; k < n; k++) st.rows[0][k] = (unsigned char)k;
    st.blk[0].start = 0; st.blk[0].len = (unsigned char)n; st.blk[0].nrows = (unsigned char)n; st.blk[0].nprev = 0;
  }
  bsync(c);
}
B2P_HD bool sp_coupled(const SpStruct& st, int bi, int bj) { return bi == bj || ((st.later[bj] >> bi) & 1u); }   // bi >= bj

// a block that is coupled to nothing (a prop resting on the table) is factored / solved by ONE lane in registers
// Pivot floor of the Cholesky factorisations.  The Newton Hessian M + J'WJ carries cone blocks whose entries grow like 1/|tangential
// residual|; in fp32 the rounding noise of such entries (1e-7 relative) can exceed the small pivots the light gripper links and
// props contribute, and a pivot that comes out <= 0 used to be replaced by 1e-15 -- whose inverse square root (3e7) sent the Newton
// direction to inf (2 of 17 000 fp32 data-generation episodes).  A pivot below 64 eps of its ORIGINAL diagonal entry is noise: it is
// replaced by that bound, which keeps the factor finite and the direction a descent direction (the exact line search does the rest).
template <typename T> B2P_HD T pivot_floor(T diag0) { const T f = 64 * Num<T>::eps() * fabs(diag0); return f > Num<T>::minval() ? f : Num<T>::minval(); }
B2P_HD bool sp_local(const SpStruct& st, int b) { return st.blk[b].len == 6 && st.blk[b].nrows == 6 && st.blk[b].nprev == 0; }
template <typename T>
B2P_HD void chol6_local(T* A, int s0) {
  T a[21];
#pragma unroll
  for (int i = 0; i < 6; i++)
#pragma unroll
    for (int j = 0; j <= i; j++) a[(i * (i + 1)) / 2 + j] = A[tri_idx(s0 + i, s0 + j)];
  T d0[6];
#pragma unroll
  for (int k = 0; k < 6; k++) d0[k] = a[(k * (k + 1)) / 2 + k];
#pragma unroll
  for (int k = 0; k < 6; k++) {
    T d = a[(k * (k + 1)) / 2 + k];
    const T fl = pivot_floor(d0[k]);
    d = d > fl ? sqrt(d) : sqrt(fl);
    const T id = 1 / d;
    a[(k * (k + 1)) / 2 + k] = d;
#pragma unroll
    for (int i = k + 1; i < 6; i++) a[(i * (i + 1)) / 2 + k] *= id;
#pragma unroll
    for (int i = k + 1; i < 6; i++)
#pragma unroll
      for (int j = k + 1; j <= i; j++) a[(i * (i + 1)) / 2 + j] -= a[(i * (i + 1)) / 2 + k] * a[(j * (j + 1)) / 2 + k];
  }
#pragma unroll
  for (int i = 0; i < 6; i++)
#pragma unroll
    for (int j = 0; j <= i; j++) A[tri_idx(s0 + i, s0 + j)] = a[(i * (i + 1)) / 2 + j];
}
template <typename T>
B2P_HD void solve6_local(const T* L, int s0, T* x) {   // x = 6 consecutive reals
  T l[21], y[6];
#pragma unroll
  for (int i = 0; i < 6; i++)
#pragma unroll
    for (int j = 0; j <= i; j++) l[(i * (i + 1)) / 2 + j] = L[tri_idx(s0 + i, s0 + j)];
#pragma unroll
  for (int i = 0; i < 6; i++) { T v = x[i];
#pragma unroll
    for (int j = 0; j < i; j++) v -= l[(i * (i + 1)) / 2 + j] * y[j];
    y[i] = v / l[(i * (i + 1)) / 2 + i]; }
#pragma unroll
  for (int i = 5; i >= 0; i--) { T v = y[i];
#pragma unroll
    for (int j = i + 1; j < 6; j++) v -= l[(j * (j + 1)) / 2 + i] * y[j];
    y[i] = v / l[(i * (i + 1)) / 2 + i]; }
#pragma unroll
  for (int i = 0; i < 6; i++) x[i] = y[i];
}

// ---------------------------------------------------------------------------------------- dense tail block in registers
// The tree block (arm + gripper, <= 16 dofs) is always the LAST solver block and is dense.  On the device it is factored
// with lane i owning row i in registers and the pivot column broadcast by shuffles: no shared-memory round trips and no
// barriers (136 SHFL + 120 FMA, fully unrolled) instead of 15 x (3 barriers + table-driven element loops).  The
// triangular solves keep x one element per lane and read L from shared memory (row/column reads of the packed triangle
// are bank-conflict free for 16 rows).  The host build of these templates (tests/_hostemu) keeps the generic path.
B2P_HD bool sp_dense_tail(const SpStruct& st, int b) { return b == st.nblk - 1 && st.blk[b].nrows == st.blk[b].len && st.blk[b].len > 6 && st.blk[b].len <= 16; }
#ifdef __CUDA_ARCH__
// Branch-free on purpose: entries above the diagonal (j > i) hold don't-care values that are never read, so every lane
// executes the same FMA stream and the row stays in registers (predicated updates made ptxas demote it to local memory).
// Two lanes per row: lane = 16 h + i owns row i's columns j = 2 q + h (q = 0..7), so a column's rank-1 update is 8 - k/2 shuffle +
// FMA pairs instead of 15 - k, and a row is loaded / stored in 8 steps instead of 16.  Every element sees the same operations in the
// same order as with one lane per row (a_ij -= L_ik L_jk for k ascending, then the scaling by the pivot): the same bits.
template <typename T>
__device__ __forceinline__ void chol16_warp(T* A, int s0, int n, int lane) {
  T a[8];
  const int i = lane & 15, h = lane >> 4, rb = tri_idx(s0 + i, s0);
#pragma unroll
  for (int q = 0; q < 8; q++) { const int j = 2 * q + h; const bool in = i < n && j <= i; const T v = A[in ? rb + j : 0]; a[q] = in ? v : (j == i ? T(1) : T(0)); }
  const T fl = pivot_floor(i < n ? A[rb + i] : T(1));               // relative to the row's ORIGINAL diagonal entry
#pragma unroll
  for (int k = 0; k < 16; k++) {
    const int hk = k & 1, qk = k >> 1;
    const T akk = __shfl_sync(0xffffffffu, a[qk] > fl ? a[qk] : fl, k + 16 * hk);   // lane (k, hk): its pivot, floored
    const T d = sqrt(akk);
    const T id = 1 / d;
    T lik = __shfl_sync(0xffffffffu, a[qk], i + 16 * hk);                            // a_ik of my row, from the half that owns column k
    lik = i == k ? d : lik * id;
    a[qk] = h == hk ? lik : a[qk];
    if (hk == 0) { const T ljk = __shfl_sync(0xffffffffu, lik, 2 * qk + h); a[qk] = h == 1 ? a[qk] - lik * ljk : a[qk]; }   // column k + 1 (the odd half's)
#pragma unroll
    for (int q = qk + 1; q < 8; q++) a[q] -= lik * __shfl_sync(0xffffffffu, lik, 2 * q + h);
  }
#pragma unroll
  for (int q = 0; q < 8; q++) { const int j = 2 * q + h; if (i < n && j <= i) A[rb + j] = a[q]; }
}
// x_tail <- (L L')^-1 x_tail for the dense tail block; also applies the backward updates to the np coupled earlier columns
template <typename T>
__device__ __forceinline__ void solve16_warp(const T* L, const SpStruct& st, int b, T* x, int np6, int ntd, int lane) {
  const int s0 = st.blk[b].start, n = st.blk[b].len, np = st.blk[b].nprev;
  const bool mine = lane < n;
  const int me = mine ? lane : 0;
  T xv = mine ? x[sp_dof(s0 + me, np6, ntd)] : T(0);
  const T rd = 1 / L[tri_idx(s0 + me, s0 + me)];
  const int rowbase = tri_idx(s0 + me, s0);
  for (int k = 0; k < n; k++) {                // forward: L y = b
    const T l = (mine && lane > k) ? L[rowbase + k] : T(0);
    const T xk = __shfl_sync(0xffffffffu, xv * rd, k);
    xv = lane == k ? xk : xv - l * xk;
  }
  T acc = 0;                                   // backward: L' x = y; lane a < np accumulates sum_k L[k][prev a] * x_k
  const int pcol = lane < np ? st.prev[b][lane] : 0;
  for (int k = n - 1; k >= 0; k--) {
    const int kb = tri_idx(s0 + k, 0);
    const T l = lane < k ? L[kb + s0 + lane] : T(0);
    const T lp = lane < np ? L[kb + pcol] : T(0);
    const T xk = __shfl_sync(0xffffffffu, xv * rd, k);
    xv = lane == k ? xk : xv - l * xk;
    acc += lp * xk;
  }
  if (mine) x[sp_dof(s0 + lane, np6, ntd)] = xv;
  if (lane < np) x[sp_dof(pcol, np6, ntd)] -= acc;
}
#endif

// in-place A = L L' on the packed lower triangle, restricted to the structure.  tri[e] = (a << 8) | c enumerates
// the pairs a >= c of a triangular index e (host-built table in DevModel).
template <typename T>
B2P_HDN void sp_cholesky(T* A, const SpStruct& st, const unsigned short* tri, const Ctx& c, const T* diag0 = nullptr) {   // diag0: original diagonal by solver index (pivot floor), or none
  for (int b = c.tid; b < st.nblk; b += c.nt) if (sp_local(st, b)) chol6_local(A, st.blk[b].start);
  for (int b = 0; b < st.nblk; b++) {
    if (sp_local(st, b)) continue;
#ifdef __CUDA_ARCH__
    if (sp_dense_tail(st, b)) { bsync(c); chol16_warp(A, st.blk[b].start, st.blk[b].len, c.tid); bsync(c); continue; }
#endif
    const int start = st.blk[b].start, len = st.blk[b].len, nrows = st.blk[b].nrows;
    for (int kk = 0; kk < len; kk++) {
      const int k = start + kk, R = nrows - kk - 1, kb = tri_idx(k, 0);
      const unsigned char* lst = st.rows[b] + kk + 1;
      T d = A[kb + k];
      const T fl = diag0 ? pivot_floor(diag0[k]) : Num<T>::minval();
      d = d > fl ? sqrt(d) : sqrt(fl);
      const T id = 1 / d;
      bsync(c);                                     // every lane has read the pivot
      if (c.tid == 0) A[kb + k] = d;
      for (int a = c.tid; a < R; a += c.nt) { const int i = lst[a]; A[tri_idx(i, k)] *= id; }
      bsync(c);
      const int cnt = (R * (R + 1)) >> 1;
      for (int e = c.tid; e < cnt; e += c.nt) {
        const unsigned t = tri[e]; const int i = lst[t >> 8], j = lst[t & 255u];
        A[tri_idx(i, j)] -= A[tri_idx(i, k)] * A[tri_idx(j, k)];
      }
      bsync(c);
    }
  }
}
// x <- (L L')^-1 x ; x is indexed by dof through sp_dof(u, np6, ntd)
template <typename T>
B2P_HDN void sp_solve(const T* L, const SpStruct& st, T* x, int np6, int ntd, const Ctx& c) {
  for (int b = c.tid; b < st.nblk; b += c.nt) if (sp_local(st, b)) solve6_local(L, st.blk[b].start, x + sp_dof(st.blk[b].start, np6, ntd));
  for (int b = 0; b < st.nblk; b++) {
    if (sp_local(st, b)) continue;
#ifdef __CUDA_ARCH__
    if (sp_dense_tail(st, b)) { bsync(c); solve16_warp(L, st, b, x, np6, ntd, c.tid); bsync(c); continue; }   // forward + backward of the tail in one go
#endif
    const int start = st.blk[b].start, len = st.blk[b].len, nrows = st.blk[b].nrows;
    for (int kk = 0; kk < len; kk++) {
      const int k = start + kk, R = nrows - kk - 1, dk = sp_dof(k, np6, ntd);
      const unsigned char* lst = st.rows[b] + kk + 1;
      const T xk = x[dk] / L[tri_idx(k, k)];
      bsync(c);
      if (c.tid == 0) x[dk] = xk;
      for (int a = c.tid; a < R; a += c.nt) { const int i = lst[a]; x[sp_dof(i, np6, ntd)] -= L[tri_idx(i, k)] * xk; }
      bsync(c);
    }
  }
  for (int b = st.nblk - 1; b >= 0; b--) {
    if (sp_local(st, b)) continue;
#ifdef __CUDA_ARCH__
    if (sp_dense_tail(st, b)) continue;
#endif
    const int start = st.blk[b].start, len = st.blk[b].len, np = st.blk[b].nprev;
    for (int kk = len - 1; kk >= 0; kk--) {
      const int k = start + kk, dk = sp_dof(k, np6, ntd), kb = tri_idx(k, 0);
      const T xk = x[dk] / L[kb + k];
      bsync(c);
      if (c.tid == 0) x[dk] = xk;
      for (int a = c.tid; a < np + kk; a += c.nt) { const int j = a < np ? st.prev[b][a] : start + (a - np); x[sp_dof(j, np6, ntd)] -= L[kb + j] * xk; }
      bsync(c);
    }
  }
}

}  // namespace b2p
