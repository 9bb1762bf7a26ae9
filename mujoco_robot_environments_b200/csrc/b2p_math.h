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

enum { ROW_EQ = 0, ROW_LIMIT = 1, ROW_CONTACT = 2, ROW_PYR = 3 };   // ROW_PYR: one edge of a pyramidal friction cone -- a scalar inequality row like a limit (the solver only tells EQ and CONTACT apart)
enum { ST_SATISFIED = 0, ST_QUADRATIC = 1, ST_CONE = 2 };
enum { FLAG_BADQPOS = 1, FLAG_BADQVEL = 2, FLAG_BADQACC = 4, FLAG_CONTACTFULL = 8, FLAG_EFCFULL = 16, FLAG_CANDFULL = 32, FLAG_SOLVER_MAXITER = 64 };
enum { MODE_PHYSICS = 0, MODE_OSC = 1, MODE_SETTLE = 2, MODE_FORWARD = 3, MODE_PLACE = 4 };
// counter-based uniform in [0, 1): placement proposals are a function of (seed, scene, attempt, coordinate) only
B2P_HD unsigned hash32(unsigned x) { x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16; return x; }
B2P_HD double uniform01(unsigned seed, unsigned env, unsigned k, unsigned d) { return (hash32(hash32(hash32(seed ^ (0x9e3779b9U * (env + 1U))) + k) + d) >> 8) * (1.0 / 16777216.0); }

struct Ctx { int tid, nt; int lock; int bar_id, bar_n; };   // bar_n > 0: the lockstep group is bar_n threads on named barrier bar_id (a sub-group of the CTA's warps)   // lock: bit mask of the CTA-wide stage barriers in use (LK_*), 0 = warps run free

// Optional per-stage cycle counters (build with -DB2P_PROF; read back through b2p_prof_read).  Off in the product build.
enum { PF_FK = 0, PF_DYN, PF_COLLIDE, PF_ROWS, PF_OSC, PF_SMOOTH, PF_SOLVE_INIT, PF_ASSEMBLE, PF_CHOL, PF_SOLVE, PF_LS, PF_UPDATE, PF_INTEGRATE, PF_BARRIER, PF_MISC, PF_NITER, PF_NSTEP, PF_BROAD, PF_NARROW, PF_TOTAL, PF_SUB0, PF_SUB1, PF_SUB2, PF_SUB3, PF_N };   // PF_SUBk: scratch counters for whatever stage is under the microscope
#if defined(B2P_PROF) && defined(__CUDACC__)
static __device__ unsigned long long g_prof[PF_N];
#endif
#if defined(B2P_PROF) && defined(__CUDA_ARCH__)
#define PROF_T0 long long pf_t0_ = clock64();
#define PROF_ADD(i) { const long long pf_t1_ = clock64(); if (c.tid == 0) atomicAdd(&g_prof[i], (unsigned long long)(pf_t1_ - pf_t0_)); pf_t0_ = clock64(); }
#define PROF_COUNT(i, n) { if (c.tid == 0) atomicAdd(&g_prof[i], (unsigned long long)(n)); }
#define PROF_RESET pf_t0_ = clock64();
#else
#define PROF_T0
#define PROF_ADD(i)
#define PROF_COUNT(i, n)
#define PROF_RESET
#endif   // lock: the CTA's warps advance stage by stage together (shared instruction stream)

template <typename T> struct Num;
template <> struct Num<double> { static B2P_HD double eps() { return 2.220446049250313e-16; } static B2P_HD double minval() { return 1e-15; } static B2P_HD double big() { return 1e300; } };
template <> struct Num<float> { static B2P_HD float eps() { return 1.1920929e-7f; } static B2P_HD float minval() { return 1e-15f; } static B2P_HD float big() { return 1e30f; } };

// ------------------------------------------------------------------------------------------ warp primitives
// One WARP steps one scene: Ctx = {lane, 32} on the device, {0, 1} in the host build of the templates.
// Parallel loops are `for (i = c.tid; i < n; i += c.nt)`, phases are separated by bsync() (= __syncwarp()).
B2P_HD void bsync(const Ctx&) {
#ifdef __CUDA_ARCH__
  __syncwarp();
#endif
}
// CTA-wide stage barrier, only in lockstep mode: the warps of a CTA (one scene each) enter every pipeline stage
// together so that they share instruction-cache lines -- the per-step code is ~10x the SM's 32 KB L1.5 I-cache.
// Every warp of the CTA must reach every csync(), whether or not it holds a scene.
enum { LK_FK = 1, LK_COLLIDE = 2, LK_ROWS = 4, LK_STEP2 = 8, LK_INTEGRATE = 16, LK_TICK = 32, LK_NEWTON = 64, LK_NARROW = 128, LK_SOLVE = 256, LK_ALL = 511, LK_DEFAULT = 63 + 128 + 256 };
B2P_HD void csync(const Ctx& c, int site) {
#ifdef __CUDA_ARCH__
  if (c.lock & site) { if (c.bar_n) asm volatile("bar.sync %0, %1;" ::"r"(c.bar_id), "r"(c.bar_n) : "memory"); else __syncthreads(); }
#else
  (void)c; (void)site;
#endif
}
// CTA-wide OR of a per-warp predicate at barrier `site`; the warp's own predicate when that barrier is not in use
B2P_HD bool csync_or(const Ctx& c, int site, bool p) {
#ifdef __CUDA_ARCH__
  if (!(c.lock & site)) return p;
  if (!c.bar_n) return __syncthreads_or(p) != 0;
  unsigned r;
  asm volatile("{ .reg .pred p, q; setp.ne.u32 q, %3, 0; bar.red.or.pred p, %1, %2, q; selp.u32 %0, 1, 0, p; }" : "=r"(r) : "r"(c.bar_id), "r"(c.bar_n), "r"((unsigned)p) : "memory");
  return r != 0;
#else
  (void)c; (void)site; return p;
#endif
}
B2P_HD bool csync_and(const Ctx& c, bool p) {
#ifdef __CUDA_ARCH__
  if (!c.lock) return p;
  if (!c.bar_n) return __syncthreads_and(p) != 0;
  unsigned r;
  asm volatile("{ .reg .pred p, q; setp.ne.u32 q, %3, 0; bar.red.and.pred p, %1, %2, q; selp.u32 %0, 1, 0, p; }" : "=r"(r) : "r"(c.bar_id), "r"(c.bar_n), "r"((unsigned)p) : "memory");
  return r != 0;
#else
  (void)c; return p;
#endif
}
// deterministic warp-wide sum of K values (xor butterfly: every lane ends with the same bits)
template <typename T, int K>
B2P_HD void block_sum(const Ctx& c, T (&v)[K], T* /*unused*/ = nullptr) {
#ifdef __CUDA_ARCH__
#pragma unroll
  for (int k = 0; k < K; k++) {
    T x = v[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    v[k] = x;
  }
#else
  (void)c; (void)v;
#endif
}
// exclusive prefix sum of a small non-negative int over the warp (lane order); total returned to all lanes
B2P_HD int block_scan(const Ctx& c, int v, int& total, int* /*unused*/ = nullptr) {
#ifdef __CUDA_ARCH__
  int x = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(0xffffffffu, x, o); if (c.tid >= o) x += y; }
  total = __shfl_sync(0xffffffffu, x, 31);
  return x - v;
#else
  (void)c; total = v; return 0;
#endif
}
// minimum of an int over the warp
B2P_HD int block_min(const Ctx& c, int v) {
#ifdef __CUDA_ARCH__
  (void)c;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { const int y = __shfl_xor_sync(0xffffffffu, v, o); v = y < v ? y : v; }
#else
  (void)c;
#endif
  return v;
}
// bitwise OR of an unsigned over the warp
B2P_HD unsigned block_or(const Ctx& c, unsigned v) {
#ifdef __CUDA_ARCH__
  (void)c;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v |= __shfl_xor_sync(0xffffffffu, v, o);
#else
  (void)c;
#endif
  return v;
}
// rank of this lane among the lanes whose flag is set (lane order); total = number of set flags
B2P_HD int block_rank(const Ctx& c, bool flag, int& total) {
#ifdef __CUDA_ARCH__
  const unsigned b = __ballot_sync(0xffffffffu, flag);
  total = __popc(b);
  return __popc(b & ((1u << c.tid) - 1u));
#else
  (void)c; total = flag ? 1 : 0; return 0;
#endif
}
// OR of a per-lane predicate over the warp
B2P_HD bool warp_any(const Ctx& c, bool p) {
#ifdef __CUDA_ARCH__
  (void)c; return __any_sync(0xffffffffu, p);
#else
  (void)c; return p;
#endif
}

// Debug build only (-DB2P_NANTRACE, tools/catch_bad_state.py): the first probe that sees a non-finite value leaves its code in bits 8.. of the scene's status
#ifdef B2P_NANTRACE
#define NANCHK(code, ptr, n) { unsigned nb_ = 0; for (int i_ = c.tid; i_ < (n); i_ += c.nt) if (!(fabs((ptr)[i_]) < 1e30)) nb_ = 1; nb_ = block_or(c, nb_); if (nb_ && c.tid == 0 && !(s.flags >> 8)) s.flags |= (code) << 8; bsync(c); }
#else
#define NANCHK(code, ptr, n)
#endif

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
template <typename T> B2P_HD T clampT(T x, T lo, T hi) { return x < lo ? lo : (x > hi ? hi : x); }
template <typename T> B2P_HD void cp3(T* d, const T* s) { d[0] = s[0]; d[1] = s[1]; d[2] = s[2]; }
template <typename T> B2P_HD void mulquat(T* r, const T* a, const T* b) {
  T w = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3], x = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
  T y = a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1], z = a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0];
  r[0] = w; r[1] = x; r[2] = y; r[3] = z;
}
// v rotated by the unit quaternion q: v + 2 w (u x v) + 2 u x (u x v)
template <typename T> B2P_HD void rotquat(T* r, const T* q, const T* v) {
  T uv[3], uuv[3];
  cross3(uv, q + 1, v); cross3(uuv, q + 1, uv);
  for (int k = 0; k < 3; k++) r[k] = v[k] + 2 * (q[0] * uv[k] + uuv[k]);
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


}  // namespace b2p
