#pragma once
#include "b2p_scratch.h"

namespace b2p {

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


}  // namespace b2p
