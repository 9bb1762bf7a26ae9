// Ray against a convex polytope (the arm / gripper hulls of the ray-cast observations) and the per-tile plane lists that keep the
// clip loop short.  Shared by csrc/b2p_render.cu (device) and tests/hostemu/b2p_render_cull_test.cpp (host: the listed loop must
// give the very bits of the full loop on every pixel of a tile -- tests/test_render_cull.py).
//
// A ray enters the polytope at the LARGEST entry parameter over the planes it crosses inwards and leaves at the SMALLEST exit
// parameter over the planes it crosses outwards; it misses when the former exceeds the latter.  For the 32 x 8 pixel rays of one
// tile both extrema are taken by a handful of the <= 128 planes: the facets that project into the tile and their neighbours.  The
// rays' directions are affine in the pixel coordinates, so every plane's denominator (normal . direction) ranges over an interval
// given by the tile's corners, and with it the plane's parameter t = dist / den (dist does not depend on the pixel: all rays start at
// the camera).  A plane that is an entry for every ray of the tile and whose largest t stays below the tile-wide LOWER bound of the
// entry maximum can never be that maximum; likewise an exit plane whose smallest t stays above the upper bound of the exit minimum.
// Such planes are left out of the tile's list.  Every bound carries a slack far above the rounding of the per-ray arithmetic, so the
// listed loop visits (in the same order) every plane that can update the running extrema of the full loop: same hit, same t, same
// facet, bit for bit.
#pragma once
#include <math.h>

#if defined(__CUDACC__)
#define RHD __host__ __device__ __forceinline__
#else
#define RHD inline
#endif

namespace b2p_rcull {

struct Plane4 { float x, y, z, w; };     // outward unit normal (geom frame), offset: inside is n . p <= w

#if defined(__CUDA_ARCH__)
__device__ __forceinline__ Plane4 ld_plane(const Plane4* p) { const float4 v = __ldg(reinterpret_cast<const float4*>(p)); Plane4 r = {v.x, v.y, v.z, v.w}; return r; }
__device__ __forceinline__ float fma_rn(float a, float b, float c) { return __fmaf_rn(a, b, c); }
#else
inline Plane4 ld_plane(const Plane4* p) { return *p; }
inline float fma_rn(float a, float b, float c) { return fmaf(a, b, c); }
#endif

// The clip loop over `cnt` planes: planes[list[q]] when a list is given, planes[q] otherwise.  ol / dl: ray origin and direction in
// the geom frame.  Entry / exit parameters are kept as fractions num / den (entry: den < 0, exit: den > 0) and compared by cross
// multiplication, so the loop has no division and no data-dependent branch besides its exit; one division at the end.
// Explicit fused multiply-adds in the two dot products (the oracle calls C99 fmaf at the same places; everything else is unfused).
RHD bool hull_clip(const float* ol, const float* dl, const Plane4* planes, const unsigned char* list, int cnt, float* t_out, int* best_out) {
  float tn_n = 1e30f, tn_d = -1.0f, tf_n = 1e30f, tf_d = 1.0f; int best = -1;
  for (int q = 0; q < cnt; q++) {
    const int i = list ? (int)list[q] : q;
    const Plane4 P = ld_plane(planes + i);
    const float den = fma_rn(P.z, dl[2], fma_rn(P.y, dl[1], P.x * dl[0]));
    const float dist = P.w - fma_rn(P.z, ol[2], fma_rn(P.y, ol[1], P.x * ol[0]));
    if (den == 0.0f) { if (dist < 0.0f) return false; continue; }
    if (den < 0.0f) { if (dist * tn_d > tn_n * den) { tn_n = dist; tn_d = den; best = i; } }      // dist / den > tn
    else if (dist * tf_d < tf_n * den) { tf_n = dist; tf_d = den; }                              // dist / den < tf
    if (tn_n * tf_d < tf_n * tn_d) return false;                                                 // tn > tf
  }
  if (best < 0) return false;
  const float tn = tn_n / tn_d;
  if (tn <= 0.0f) return false;
  *t_out = tn; *best_out = best;
  return true;
}

// ---------------------------------------------------------------------------------------- the rays of one tile, in a geom's frame
// pixel (u, v) -> camera-frame direction (dc0, dc1, -1), dc0 = (u - cx) / f, dc1 = (cy - v) / f; in the geom frame the direction is
// A dc0 + B dc1 + C0 and the origin is ol, for every pixel alike.
struct TileRays { float A[3], B[3], C0[3], ol[3]; float dc0_lo, dc0_hi, dc1_lo, dc1_hi; };

RHD void tile_rays(const float* cam_pos, const float* cam_mat, float f, float cx, float cy, int tx0, int ty0, int tw, int th,
                   const float* pos, const float* mat, TileRays& R) {
  const float w0 = cam_pos[0] - pos[0], w1 = cam_pos[1] - pos[1], w2 = cam_pos[2] - pos[2];
  for (int k = 0; k < 3; k++) {
    R.ol[k] = mat[k] * w0 + mat[3 + k] * w1 + mat[6 + k] * w2;                         // as in the per-ray code: the same bits
    R.A[k] = mat[k] * cam_mat[0] + mat[3 + k] * cam_mat[3] + mat[6 + k] * cam_mat[6];
    R.B[k] = mat[k] * cam_mat[1] + mat[3 + k] * cam_mat[4] + mat[6 + k] * cam_mat[7];
    R.C0[k] = -(mat[k] * cam_mat[2] + mat[3 + k] * cam_mat[5] + mat[6 + k] * cam_mat[8]);
  }
  R.dc0_lo = ((float)tx0 - cx) / f; R.dc0_hi = ((float)(tx0 + tw - 1) - cx) / f;
  R.dc1_lo = (cy - (float)(ty0 + th - 1)) / f; R.dc1_hi = (cy - (float)ty0) / f;
}

enum { PL_MIXED = 0, PL_ENTRY = 1, PL_EXIT = 2 };

// class of one plane over the tile and the range of its parameter t (entry / exit classes only)
RHD int plane_class(const Plane4& P, const TileRays& R, float* tmin, float* tmax) {
  const float al = P.x * R.A[0] + P.y * R.A[1] + P.z * R.A[2], be = P.x * R.B[0] + P.y * R.B[1] + P.z * R.B[2], ga = P.x * R.C0[0] + P.y * R.C0[1] + P.z * R.C0[2];
  const float a0 = al * R.dc0_lo, a1 = al * R.dc0_hi, b0 = be * R.dc1_lo, b1 = be * R.dc1_hi;
  // |normal| = 1 and |direction| < 2 inside any field of view below 120 degrees: 1e-5 absolute is ~50 x the rounding of either form
  const float slack = 1e-5f;
  const float den_lo = ga + fminf(a0, a1) + fminf(b0, b1) - slack, den_hi = ga + fmaxf(a0, a1) + fmaxf(b0, b1) + slack;
  if (!(den_hi < 0.0f) && !(den_lo > 0.0f)) return PL_MIXED;          // also catches NaN
  const float dist = P.w - fma_rn(P.z, R.ol[2], fma_rn(P.y, R.ol[1], P.x * R.ol[0]));
  const float t0 = dist / den_lo, t1 = dist / den_hi;
  *tmin = fminf(t0, t1); *tmax = fmaxf(t0, t1);
  return den_hi < 0.0f ? PL_ENTRY : PL_EXIT;
}
RHD float cull_margin(float bound) { return 1e-4f * (1.0f + fabsf(bound)); }
// every ray of the tile enters later than it leaves: the hull is not in this tile (the oriented-box rectangle that listed it is loose
// for elongated links seen at an angle)
RHD bool tile_misses(float entry_lo, float exit_hi) { return entry_lo > exit_hi + cull_margin(entry_lo) + cull_margin(exit_hi); }
// entry_lo: max over the tile's entry planes of tmin (lower bound of every ray's entry parameter); exit_hi: min over its exit
// planes of tmax (upper bound of every ray's exit parameter)
RHD bool plane_keep(int cls, float tmin, float tmax, float entry_lo, float exit_hi) {
  if (cls == PL_ENTRY) return !(tmax < entry_lo - cull_margin(entry_lo));
  if (cls == PL_EXIT) return !(tmin > exit_hi + cull_margin(exit_hi));
  return true;
}

}  // namespace b2p_rcull
