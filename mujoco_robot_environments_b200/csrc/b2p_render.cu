// CUDA ray-cast observations for the rearrangement scenes (C-ABI: include/b2p_render.h).
//
// What it replaces: the reference's mujoco.Renderer rgb / depth / segmentation of the overhead camera
// (tasks/rearrangement.py:182-186, :460-478, :273-275).  B200-first design:
//   * The camera, ground, table, arm base and target zones never move, so their image (rgb + depth + segmentation, 3.4 MB)
//     is ray-cast ONCE at handle creation and stays resident in the 126 MB L2.
//   * Per frame, a tiny bounds kernel projects the oriented bounding box of every moving geom (arm links, gripper, props) of
//     every scene to a pixel rectangle.  The persistent frame kernel (148 x k CTAs, work items = (scene, 32x8 tile) pulled
//     by one warp each from an atomic queue) then either COPIES the background tile with 16-byte vector loads / stores -- ~85 % of the tiles
//     of a typical frame: pure HBM write streaming -- or ray-casts the tile against the moving geoms whose rectangle overlaps
//     it; what the ray finds among the static geoms and the zones comes from per-pixel background layers written with the
//     static image (see cast_pixel), and every listed hull gets a per-tile list of the planes that can matter (b2p_render_cull.h).
//   * The render model (<= 32 geoms, 6 zones, camera, lights: 5 kB) is staged once per CTA in shared memory; hull planes
//     stay in global memory (L2).  RGB is staged per tile in shared memory so every global store is a full 16-byte vector.
// The algorithmic HBM traffic is the output itself: 3 + 4 + 4 bytes per pixel.
//
// This translation unit is compiled with -fmad=false and uses only IEEE +,-,*,/,sqrt so that every pixel is bit-identical
// to the plain-C restatement in oracle/b2p_render_oracle.c (tests/test_gpu_render.py).
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstring>
#include <string>
#include <vector>

#include "../../include/b2p.h"
#include "../../include/b2p_render.h"
#include "b2p_render_cull.h"

namespace b2p { int api_fail(int code, const char* msg); }
#define RCK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return b2p::api_fail(B2P_ERR_CUDA, (std::string(#x) + ": " + cudaGetErrorString(e_)).c_str()); } while (0)

namespace {

constexpr int TW = 32, TH = 8, NT = 256, NW = NT / 32, RMAXG = 32, RMAXZ = 8;
constexpr int GEOM_PLANE = 0, GEOM_BOX = 6, GEOM_MESH = 7;
constexpr unsigned MAGIC = 0x52503242u;
constexpr int PL_CAP = 1024, PL_ROUNDS = 4;      // per-warp pool of listed plane indices; hulls of up to 32 * PL_ROUNDS planes get a list
using b2p_rcull::Plane4;

struct RGeom { int type, padr, pnum, prop, is_static, pad_i[3]; float rgba[4], size[3], cen[3], rbound, spos[3], smat[9], pad_f; };
struct RZone { float pos[3], half[3], rgba[4], pad[2]; };
struct RCam { float pos[3], mat[9], f, cx, cy, zfar, hl_amb, hl_diff, lpos[3], ldiff, rgb2[3], check, pad[6]; };
static_assert(sizeof(RGeom) == 128 && sizeof(RZone) == 48 && sizeof(RCam) == 128, "B2PR record sizes");
struct RModel { int ngeom, nplane, nzone, W, H, nprop, n_static, pad; RCam cam; RGeom geom[RMAXG]; RZone zone[RMAXZ]; int static_ids[RMAXG]; };

struct FrameArgs {
  int n_env, pose_stride, dtype, background;
  const void* xpos; const void* xmat; const float* prop_rgba; const void* prop_size; const unsigned char* prop_active;
  const short4* rects;
  const uint8_t* bg_rgb; const float* bg_depth; const int* bg_seg; float* layers;      // layers: [NLAYER][H * W], see cast_pixel
  uint8_t* rgb; float* depth; int* seg;
  const float4* planes;
  unsigned long long* queue; unsigned long long* tile_stats;
};

__device__ __forceinline__ float ld_real(const void* p, size_t i, int dtype) { return dtype == 0 ? (float)((const double*)p)[i] : ((const float*)p)[i]; }

// ---------------------------------------------------------------------------------------- moving-geom pixel rectangles
// one thread per (scene, geom): conservative pixel rectangle of the geom's oriented bounding box (empty for static geoms,
// inactive props and boxes entirely behind / beside the view)
__global__ void render_bounds_kernel(const RModel* __restrict__ gm, FrameArgs a, short4* __restrict__ rects) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.n_env * RMAXG) return;
  const int env = i / RMAXG, g = i % RMAXG;
  short4 r = make_short4(1, 1, 0, 0);
  if (g < gm->ngeom) {
    const RGeom& G = gm->geom[g];
    bool on = !G.is_static;
    float hs[3] = {G.size[0], G.size[1], G.size[2]};       // box half sizes / padded hull AABB half sizes, centred at G.cen
    if (on && G.prop >= 0) {
      if (a.prop_active && !a.prop_active[(size_t)env * gm->nprop + G.prop]) on = false;
      if (a.prop_size) for (int k = 0; k < 3; k++) hs[k] = ld_real(a.prop_size, ((size_t)env * gm->nprop + G.prop) * 3 + k, a.dtype);
    }
    if (on) {
      const size_t pb = (size_t)env * a.pose_stride;
      float p[3], m[9];
      for (int k = 0; k < 3; k++) p[k] = ld_real(a.xpos, pb + 3 * g + k, a.dtype);
      for (int k = 0; k < 9; k++) m[k] = ld_real(a.xmat, pb + 9 * g + k, a.dtype);
      const RCam& C = gm->cam;
      // the 8 corners of the oriented box in the camera frame (x right, y up, z back): their pixel bounding box, +-2 px slack
      float u0 = 1e9f, u1 = -1e9f, v0 = 1e9f, v1 = -1e9f; int behind = 0;
      for (int c8 = 0; c8 < 8; c8++) {
        const float l0 = G.cen[0] + ((c8 & 1) ? hs[0] : -hs[0]) * 1.01f, l1 = G.cen[1] + ((c8 & 2) ? hs[1] : -hs[1]) * 1.01f, l2 = G.cen[2] + ((c8 & 4) ? hs[2] : -hs[2]) * 1.01f;
        float w[3];
        for (int k = 0; k < 3; k++) w[k] = p[k] + m[3 * k] * l0 + m[3 * k + 1] * l1 + m[3 * k + 2] * l2 - C.pos[k];
        const float x = C.mat[0] * w[0] + C.mat[3] * w[1] + C.mat[6] * w[2], y = C.mat[1] * w[0] + C.mat[4] * w[1] + C.mat[7] * w[2],
                    dz = -(C.mat[2] * w[0] + C.mat[5] * w[1] + C.mat[8] * w[2]);
        if (dz < 0.02f) { behind++; continue; }
        const float u = C.cx + C.f * x / dz, v = C.cy - C.f * y / dz;
        u0 = fminf(u0, u); u1 = fmaxf(u1, u); v0 = fminf(v0, v); v1 = fmaxf(v1, v);
      }
      if (behind > 0 && behind < 8) { u0 = v0 = -1e9f; u1 = v1 = 1e9f; }       // straddles the camera plane: do not cull
      if (behind < 8) {
        u0 = fmaxf(u0 - 2.0f, 0.0f); v0 = fmaxf(v0 - 2.0f, 0.0f); u1 = fminf(u1 + 2.0f, (float)(gm->W - 1)); v1 = fminf(v1 + 2.0f, (float)(gm->H - 1));
        if (u0 <= u1 && v0 <= v1) r = make_short4((short)u0, (short)v0, (short)u1, (short)v1);
      }
    }
  }
  rects[i] = r;
}

// ---------------------------------------------------------------------------------------- ray maths (mirrored in the C oracle)
struct Hit { float t; int id; float n[3]; float rgb[3]; };

// ray (o, d) against an oriented box; entering face only (the camera is never inside a geom)
__device__ __forceinline__ bool ray_box(const float* o, const float* d, const float* pos, const float* mat, const float* half, float* t_out, float* n_out) {
  float ol[3], dl[3];
  const float w0 = o[0] - pos[0], w1 = o[1] - pos[1], w2 = o[2] - pos[2];
  for (int k = 0; k < 3; k++) { ol[k] = mat[k] * w0 + mat[3 + k] * w1 + mat[6 + k] * w2; dl[k] = mat[k] * d[0] + mat[3 + k] * d[1] + mat[6 + k] * d[2]; }
  float tn = -1e30f, tf = 1e30f; int ax = 0; float sg = 0.0f;
  for (int k = 0; k < 3; k++) {
    if (dl[k] != 0.0f) {
      const float inv = 1.0f / dl[k];
      float t1 = (-half[k] - ol[k]) * inv, t2 = (half[k] - ol[k]) * inv;
      const float s = dl[k] > 0.0f ? -1.0f : 1.0f;
      if (t1 > t2) { const float q = t1; t1 = t2; t2 = q; }
      if (t1 > tn) { tn = t1; ax = k; sg = s; }
      if (t2 < tf) tf = t2;
    } else if (fabsf(ol[k]) > half[k]) return false;
  }
  if (tn > tf || tn <= 0.0f) return false;
  *t_out = tn;
  n_out[0] = sg * mat[ax]; n_out[1] = sg * mat[3 + ax]; n_out[2] = sg * mat[6 + ax];
  return true;
}

__device__ __forceinline__ bool ray_hull(const float* o, const float* d, const float* pos, const float* mat, const float* cen, const float* half,
                                         const Plane4* __restrict__ planes, int np, const unsigned char* list, int nlist, float t_max, float* t_out, float* n_out) {
  float ol[3], dl[3];
  const float w0 = o[0] - pos[0], w1 = o[1] - pos[1], w2 = o[2] - pos[2];
  for (int k = 0; k < 3; k++) { ol[k] = mat[k] * w0 + mat[3 + k] * w1 + mat[6 + k] * w2; dl[k] = mat[k] * d[0] + mat[3 + k] * d[1] + mat[6 + k] * d[2]; }
  // padded AABB pre-test (conservative: can only reject rays that miss the hull)
  {
    float tn = -1e30f, tf = 1e30f;
    for (int k = 0; k < 3; k++) {
      const float c = ol[k] - cen[k];
      if (dl[k] != 0.0f) {
        const float inv = 1.0f / dl[k];
        float t1 = (-half[k] - c) * inv, t2 = (half[k] - c) * inv;
        if (t1 > t2) { const float q = t1; t1 = t2; t2 = q; }
        if (t1 > tn) tn = t1;
        if (t2 < tf) tf = t2;
      } else if (fabsf(c) > half[k]) return false;
    }
    if (tn > tf || tf <= 0.0f) return false;
    if (tn > t_max) return false;      // the hull lies inside the box: it cannot beat (or tie) a hit nearer than the box's entry
  }
  // the clip loop (b2p_render_cull.h): over the tile's plane list when there is one, over all np planes otherwise
  float tn; int best;
  if (!b2p_rcull::hull_clip(ol, dl, planes, list, list ? nlist : np, &tn, &best)) return false;
  const Plane4 P = b2p_rcull::ld_plane(planes + best);
  *t_out = tn;
  for (int k = 0; k < 3; k++) n_out[k] = mat[3 * k] * P.x + mat[3 * k + 1] * P.y + mat[3 * k + 2] * P.z;
  return true;
}

__device__ __forceinline__ bool ray_ground(const float* o, const float* d, const float* pos, const float* mat, const float* size, float check, const float* rgb1, const float* rgb2,
                                           float* t_out, float* n_out, float* rgb_out) {
  float ol[3], dl[3];
  const float w0 = o[0] - pos[0], w1 = o[1] - pos[1], w2 = o[2] - pos[2];
  for (int k = 0; k < 3; k++) { ol[k] = mat[k] * w0 + mat[3 + k] * w1 + mat[6 + k] * w2; dl[k] = mat[k] * d[0] + mat[3 + k] * d[1] + mat[6 + k] * d[2]; }
  if (!(dl[2] < 0.0f) || !(ol[2] > 0.0f)) return false;
  const float t = -ol[2] / dl[2];
  const float x = ol[0] + t * dl[0], y = ol[1] + t * dl[1];
  if (fabsf(x) > size[0] || fabsf(y) > size[1]) return false;
  const int cell = (int)floorf(x / check) + (int)floorf(y / check);
  const float* c = (cell & 1) ? rgb2 : rgb1;
  *t_out = t; n_out[0] = mat[2]; n_out[1] = mat[5]; n_out[2] = mat[8];
  rgb_out[0] = c[0]; rgb_out[1] = c[1]; rgb_out[2] = c[2];
  return true;
}

__device__ __forceinline__ float shade_factor(const RCam& C, const float* p, const float* n) {
  const float hl = fmaxf(0.0f, n[0] * C.mat[2] + n[1] * C.mat[5] + n[2] * C.mat[8]);        // headlight: directional along the view axis
  const float L0 = C.lpos[0] - p[0], L1 = C.lpos[1] - p[1], L2 = C.lpos[2] - p[2];
  const float inv = 1.0f / sqrtf(L0 * L0 + L1 * L1 + L2 * L2);
  const float sl = fmaxf(0.0f, (n[0] * L0 + n[1] * L1 + n[2] * L2) * inv);
  return C.hl_amb + C.hl_diff * hl + C.ldiff * sl;
}
__device__ __forceinline__ float clamp01(float x) { return fminf(1.0f, fmaxf(0.0f, x)); }

struct TileGeoms {
  float pos[RMAXG][3], mat[RMAXG][9], size[RMAXG][3], rgb[RMAXG][3]; unsigned char order[RMAXG]; int n_order;
  short pl_off[RMAXG], pl_cnt[RMAXG];            // per hull geom: its plane list inside plist (cnt < 0: none, use every plane)
  unsigned char plist[PL_CAP];
};

// Background layers (written once with the static image, read by every ray-cast tile): what a pixel's ray finds among the things
// that never move.  [0] depth of the nearest static opaque hit (zfar: none), [1] its geom id, [2..4] its shaded colour,
// [5] depth of the nearest zone in front of it, [6] that zone's index (-1: none), [7..9] alpha * the zone's shaded colour.
// A frame's ray then only has to be tested against the MOVING geoms listed for its tile: the nearest opaque hit is the nearer of
// that and layer 0 (ties: smaller geom id), and the zone of layer 6 is blended in when it lies in front of the winner -- the
// floats the one-pass formulation computes, in the same operations, so every pixel keeps its bits.
constexpr int NLAYER = 10;

// one pixel.  BG: the static geoms and the zones (writes the layers); otherwise the listed moving geoms against the layers
template <bool BG>
__device__ void cast_pixel(const RModel& M, const TileGeoms& tg, const Plane4* __restrict__ planes, int u, int v, float* __restrict__ lay, size_t lstride,
                           unsigned char* rgb_out, float* depth_out, int* seg_out) {
  const RCam& C = M.cam;
  const float dc0 = ((float)u - C.cx) / C.f, dc1 = (C.cy - (float)v) / C.f;             // camera-frame direction (x, y, -1): t is the axial depth
  float d[3], o[3] = {C.pos[0], C.pos[1], C.pos[2]};
  for (int k = 0; k < 3; k++) d[k] = C.mat[3 * k] * dc0 + C.mat[3 * k + 1] * dc1 - C.mat[3 * k + 2];
  float best_t = C.zfar; int best_id = -1; float bn[3] = {0, 0, 0}, brgb[3] = {0, 0, 0};
  bool won = false;
  if (!BG) { best_t = lay[0]; best_id = __float_as_int(lay[lstride]); }
  // Listed geoms nearest first (tg.order): a hull whose bounding box starts behind the current hit is skipped without touching its
  // planes.  The winner is order independent: smallest t, then smallest geom id.
  const int n_items = BG ? M.n_static : tg.n_order;
  for (int it = 0; it < n_items; it++) {
    const int g = BG ? M.static_ids[it] : tg.order[it];
    const RGeom& G = M.geom[g];
    const bool dyn = !G.is_static;
    const float* pos = dyn ? tg.pos[g] : G.spos;
    const float* mat = dyn ? tg.mat[g] : G.smat;
    float t, n[3], rgb[3];
    bool hit;
    if (G.type == GEOM_BOX) { hit = ray_box(o, d, pos, mat, (dyn && G.prop >= 0) ? tg.size[g] : G.size, &t, n); }
    else if (G.type == GEOM_MESH) {
      const int lc = tg.pl_cnt[g];
      if (lc == 0) continue;                                                           // no ray of this tile meets the hull
      hit = ray_hull(o, d, pos, mat, G.cen, G.size, planes + G.padr, G.pnum, lc > 0 ? tg.plist + tg.pl_off[g] : nullptr, lc, best_t, &t, n);
    }
    else if (G.type == GEOM_PLANE) { hit = ray_ground(o, d, pos, mat, G.size, C.check, G.rgba, C.rgb2, &t, n, rgb); }
    else hit = false;
    if (hit && (t < best_t || (t == best_t && g < best_id))) {
      best_t = t; best_id = g; bn[0] = n[0]; bn[1] = n[1]; bn[2] = n[2]; won = true;
      if (G.type != GEOM_PLANE) { const float* c = (dyn && G.prop >= 0) ? tg.rgb[g] : G.rgba; rgb[0] = c[0]; rgb[1] = c[1]; rgb[2] = c[2]; }
      brgb[0] = rgb[0]; brgb[1] = rgb[1]; brgb[2] = rgb[2];
    }
  }
  float col[3] = {0.0f, 0.0f, 0.0f};
  if (BG || won) {
    if (best_id >= 0) {
      const float p[3] = {o[0] + best_t * d[0], o[1] + best_t * d[1], o[2] + best_t * d[2]};
      const float s = shade_factor(C, p, bn);
      for (int k = 0; k < 3; k++) col[k] = clamp01(clamp01(brgb[k]) * s);
    }
  } else { for (int k = 0; k < 3; k++) col[k] = lay[(2 + k) * lstride]; }                 // the static hit stands: its colour was shaded once
  if (BG) {
    lay[0] = best_t; lay[lstride] = __int_as_float(best_id);
    for (int k = 0; k < 3; k++) lay[(2 + k) * lstride] = col[k];
    const float ident[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    float zt = best_t; int zid = -1; float zn[3] = {0, 0, 0}, zterm[3] = {0, 0, 0};
    for (int z = 0; z < M.nzone; z++) {
      float t, n[3];
      if (ray_box(o, d, M.zone[z].pos, ident, M.zone[z].half, &t, n) && t < zt) { zt = t; zid = z; zn[0] = n[0]; zn[1] = n[1]; zn[2] = n[2]; }
    }
    if (zid >= 0) {
      const float p[3] = {o[0] + zt * d[0], o[1] + zt * d[1], o[2] + zt * d[2]};
      const float s = shade_factor(C, p, zn), al = M.zone[zid].rgba[3];
      for (int k = 0; k < 3; k++) { zterm[k] = al * clamp01(M.zone[zid].rgba[k] * s); col[k] = zterm[k] + (1.0f - al) * col[k]; }
      best_t = zt; best_id = M.ngeom + zid;
    }
    lay[5 * lstride] = zt; lay[6 * lstride] = __int_as_float(zid);
    for (int k = 0; k < 3; k++) lay[(7 + k) * lstride] = zterm[k];
  } else {
    // the nearest zone in front of the static hit is the nearest zone in front of any nearer hit too, if it is in front of it at all
    const float zt = lay[5 * lstride]; const int zid = __float_as_int(lay[6 * lstride]);
    if (zid >= 0 && zt < best_t) {
      const float al = M.zone[zid].rgba[3];
      for (int k = 0; k < 3; k++) col[k] = lay[(7 + k) * lstride] + (1.0f - al) * col[k];
      best_t = zt; best_id = M.ngeom + zid;
    }
  }
  for (int k = 0; k < 3; k++) rgb_out[k] = (unsigned char)(col[k] * 255.0f + 0.5f);
  *depth_out = best_t; *seg_out = best_id;
}

// ---------------------------------------------------------------------------------------- persistent frame kernel
// One WARP per work item (scene, 32x8 tile); warps never synchronise with each other after the model is staged.
struct WarpScratch { TileGeoms tg; __align__(16) unsigned char rgb[TH][TW * 3]; };

template <bool BG>
__global__ void __launch_bounds__(NT) render_frame_kernel(const RModel* __restrict__ gm, FrameArgs a) {
  __shared__ RModel M;
  __shared__ WarpScratch ws[NW];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < (int)(sizeof(RModel) / 4); i += NT) ((int*)&M)[i] = ((const int*)gm)[i];
  __syncthreads();
  WarpScratch& S = ws[warp];
  const int tiles_x = M.W / TW, ntile = tiles_x * (M.H / TH);
  const long long total = (long long)a.n_env * ntile;
  unsigned n_copy = 0, n_cast = 0;
  for (;;) {
    long long item = 0;
    if (lane == 0) item = atomicAdd(a.queue, 1ULL);
    item = __shfl_sync(0xffffffffu, item, 0);
    if (item >= total) break;
    const int env = (int)(item / ntile), tile = (int)(item % ntile), tx0 = (tile % tiles_x) * TW, ty0 = (tile / tiles_x) * TH;
    bool hit = false;
    if (!BG) { const short4 r = a.rects[(size_t)env * RMAXG + lane]; hit = r.x <= r.z && r.x <= tx0 + TW - 1 && r.z >= tx0 && r.y <= ty0 + TH - 1 && r.w >= ty0; }
    const unsigned mask = __ballot_sync(0xffffffffu, hit);
    const size_t img = (size_t)env * M.H * M.W;
    bool cast = BG || mask != 0;
    if (cast) {
      __syncwarp();
      if ((mask >> lane) & 1u) {
        const int g = lane; const size_t pb = (size_t)env * a.pose_stride;
        for (int k = 0; k < 3; k++) S.tg.pos[g][k] = ld_real(a.xpos, pb + 3 * g + k, a.dtype);
        for (int k = 0; k < 9; k++) S.tg.mat[g][k] = ld_real(a.xmat, pb + 9 * g + k, a.dtype);
        const int p = M.geom[g].prop;
        if (p >= 0) {
          const size_t q = (size_t)env * M.nprop + p;
          for (int k = 0; k < 3; k++) {
            S.tg.size[g][k] = a.prop_size ? ld_real(a.prop_size, q * 3 + k, a.dtype) : M.geom[g].size[k];
            S.tg.rgb[g][k] = a.prop_rgba ? a.prop_rgba[q * 4 + k] : M.geom[g].rgba[k];
          }
        }
      }
      {   // nearest-first order of the listed geoms (depth of the geom origin along the view axis; ties by id)
        const bool on = (mask >> lane) & 1u;
        float key = 3.0e38f;
        if (on) key = -(M.cam.mat[2] * (S.tg.pos[lane][0] - M.cam.pos[0]) + M.cam.mat[5] * (S.tg.pos[lane][1] - M.cam.pos[1]) + M.cam.mat[8] * (S.tg.pos[lane][2] - M.cam.pos[2]));
        int rank = 0;
        for (int j = 0; j < 32; j++) { const float kj = __shfl_sync(0xffffffffu, key, j); if (((mask >> j) & 1u) && (kj < key || (kj == key && j < lane))) rank++; }
        if (on) S.tg.order[rank] = (unsigned char)lane;
        if (lane == 0) S.tg.n_order = __popc(mask);
      }
      __syncwarp();
      {   // plane lists of the hulls this tile's rays are tested against (b2p_render_cull.h): lanes across planes
        int used = 0, alive = 0;
        const int n_items = BG ? M.n_static : S.tg.n_order;
        for (int it = 0; it < n_items; it++) {
          const int g = BG ? M.static_ids[it] : S.tg.order[it];
          const RGeom& G = M.geom[g];
          if (G.type != GEOM_MESH) { alive++; continue; }
          const int np = G.pnum;
          int total = -1;
          if (np <= 32 * PL_ROUNDS) {
            const bool dyn = !G.is_static;
            b2p_rcull::TileRays R;
            b2p_rcull::tile_rays(M.cam.pos, M.cam.mat, M.cam.f, M.cam.cx, M.cam.cy, tx0, ty0, TW, TH, dyn ? S.tg.pos[g] : G.spos, dyn ? S.tg.mat[g] : G.smat, R);
            const Plane4* pl = reinterpret_cast<const Plane4*>(a.planes) + G.padr;
            int cls[PL_ROUNDS]; float tmn[PL_ROUNDS], tmx[PL_ROUNDS];
            float e_lo = -3.0e38f, x_hi = 3.0e38f;
#pragma unroll
            for (int r = 0; r < PL_ROUNDS; r++) {
              const int i = r * 32 + lane; cls[r] = -1; tmn[r] = tmx[r] = 0.0f;
              if (i < np) {
                cls[r] = b2p_rcull::plane_class(b2p_rcull::ld_plane(pl + i), R, &tmn[r], &tmx[r]);
                if (cls[r] == b2p_rcull::PL_ENTRY) e_lo = fmaxf(e_lo, tmn[r]);
                if (cls[r] == b2p_rcull::PL_EXIT) x_hi = fminf(x_hi, tmx[r]);
              }
            }
            for (int o = 16; o; o >>= 1) { e_lo = fmaxf(e_lo, __shfl_xor_sync(0xffffffffu, e_lo, o)); x_hi = fminf(x_hi, __shfl_xor_sync(0xffffffffu, x_hi, o)); }
            if (b2p_rcull::tile_misses(e_lo, x_hi)) total = 0;       // the hull is not in this tile at all
            else {
              unsigned keep[PL_ROUNDS]; total = 0;
#pragma unroll
              for (int r = 0; r < PL_ROUNDS; r++) {
                keep[r] = __ballot_sync(0xffffffffu, cls[r] >= 0 && b2p_rcull::plane_keep(cls[r], tmn[r], tmx[r], e_lo, x_hi));
                total += __popc(keep[r]);
              }
              if (total == 0 || used + total > PL_CAP) total = -1;      // pool full (or nothing to leave out a list for): this hull keeps the full loop
              else {
                int at = used;
#pragma unroll
                for (int r = 0; r < PL_ROUNDS; r++) {
                  if ((keep[r] >> lane) & 1u) S.tg.plist[at + __popc(keep[r] & ((1u << lane) - 1u))] = (unsigned char)(r * 32 + lane);
                  at += __popc(keep[r]);
                }
              }
            }
          }
          if (lane == 0) { S.tg.pl_off[g] = (short)used; S.tg.pl_cnt[g] = (short)total; }
          if (total > 0) used += total;
          if (total != 0) alive++;
        }
        if (!BG && alive == 0) cast = false;       // every listed geom was a hull that no ray of the tile meets: a background tile after all
      }
      __syncwarp();
    }
    if (!cast) {
      // background tile: 16-byte vector copies L2 -> HBM
      if (a.rgb) for (int i = lane; i < TH * (TW * 3 / 16); i += 32) {
        const int row = i / (TW * 3 / 16), c = i % (TW * 3 / 16);
        const size_t off = ((size_t)(ty0 + row) * M.W + tx0) * 3 + (size_t)c * 16;
        *reinterpret_cast<uint4*>(a.rgb + img * 3 + off) = __ldg(reinterpret_cast<const uint4*>(a.bg_rgb + off));
      }
      for (int i = lane; i < TH * (TW / 4); i += 32) {
        const int row = i / (TW / 4), c = i % (TW / 4);
        const size_t off = (size_t)(ty0 + row) * M.W + tx0 + (size_t)c * 4;
        if (a.depth) *reinterpret_cast<uint4*>(a.depth + img + off) = __ldg(reinterpret_cast<const uint4*>(a.bg_depth + off));
        if (a.seg) *reinterpret_cast<uint4*>(a.seg + img + off) = __ldg(reinterpret_cast<const uint4*>(a.bg_seg + off));
      }
      n_copy++;
    } else {
      const size_t lstride = (size_t)M.W * M.H;
      for (int ly = 0; ly < TH; ly++) {
        float dep; int sg;
        const size_t off = (size_t)(ty0 + ly) * M.W + tx0 + lane;
        cast_pixel<BG>(M, S.tg, reinterpret_cast<const Plane4*>(a.planes), tx0 + lane, ty0 + ly, a.layers + off, lstride, &S.rgb[ly][lane * 3], &dep, &sg);
        if (a.depth) a.depth[img + off] = dep;
        if (a.seg) a.seg[img + off] = sg;
      }
      __syncwarp();
      if (a.rgb) for (int i = lane; i < TH * (TW * 3 / 16); i += 32) {
        const int row = i / (TW * 3 / 16), c = i % (TW * 3 / 16);
        const size_t off = ((size_t)(ty0 + row) * M.W + tx0) * 3 + (size_t)c * 16;
        *reinterpret_cast<uint4*>(a.rgb + img * 3 + off) = *reinterpret_cast<const uint4*>(&S.rgb[row][c * 16]);
      }
      n_cast++;
    }
  }
  if (lane == 0 && a.tile_stats) { if (n_copy) atomicAdd(a.tile_stats, (unsigned long long)n_copy); if (n_cast) atomicAdd(a.tile_stats + 1, (unsigned long long)n_cast); }
}

}  // namespace

struct b2p_render_s {
  int device, sms; long long launches;
  RModel hm; RModel* d_model; float4* d_planes; short4* d_rects; int rect_cap;
  uint8_t* bg_rgb; float* bg_depth; int* bg_seg; float* d_layers; unsigned long long* d_queue; unsigned long long* d_stats;
};

static int launch_frame(b2p_render_s* h, FrameArgs& a, cudaStream_t st) {
  a.planes = h->d_planes; a.queue = h->d_queue; a.tile_stats = h->d_stats;
  a.bg_rgb = h->bg_rgb; a.bg_depth = h->bg_depth; a.bg_seg = h->bg_seg; a.layers = h->d_layers;
  RCK(cudaMemsetAsync(h->d_queue, 0, sizeof(unsigned long long), st));
  const long long total = (long long)a.n_env * (h->hm.W / TW) * (h->hm.H / TH);
  long long grid = (long long)h->sms * 3;
  if (grid > (total + NW - 1) / NW) grid = (total + NW - 1) / NW;
  if (a.background) render_frame_kernel<true><<<(int)grid, NT, 0, st>>>(h->d_model, a);
  else render_frame_kernel<false><<<(int)grid, NT, 0, st>>>(h->d_model, a);
  RCK(cudaGetLastError());
  h->launches++;
  return B2P_OK;
}

extern "C" {

int b2p_render_create(const void* blob, size_t nbytes, int device, b2p_render_handle* out) {
  if (!blob || !out || nbytes < 64 + 128) return b2p::api_fail(B2P_ERR_ARG, "b2p_render_create: bad arguments");
  const int* hd = (const int*)blob;
  if ((unsigned)hd[0] != MAGIC || hd[1] != 1) return b2p::api_fail(B2P_ERR_MODEL, "b2p_render_create: not a B2PR v1 blob");
  const int ng = hd[2], np = hd[3], nz = hd[4], W = hd[5], H = hd[6];
  if (ng < 1 || ng > RMAXG || nz < 0 || nz > RMAXZ || np < 0) return b2p::api_fail(B2P_ERR_MODEL, "b2p_render_create: geom / zone count outside the render model class");
  if (W % TW || H % TH || W <= 0 || H <= 0 || W > 32000 || H > 32000) return b2p::api_fail(B2P_ERR_MODEL, "b2p_render_create: image size must be a multiple of the 32x8 tile");
  const size_t need = 64 + 128 + (size_t)ng * 128 + (size_t)nz * 48 + (size_t)np * 16;
  if (nbytes < need) return b2p::api_fail(B2P_ERR_MODEL, "b2p_render_create: truncated blob");
  b2p_render_s* h = new b2p_render_s();
  memset(h, 0, sizeof *h);
  h->device = device;
  RModel& M = h->hm;
  M.ngeom = ng; M.nplane = np; M.nzone = nz; M.W = W; M.H = H; M.nprop = hd[7];
  const char* p = (const char*)blob + 64;
  memcpy(&M.cam, p, 128); p += 128;
  memcpy(M.geom, p, (size_t)ng * 128); p += (size_t)ng * 128;
  memcpy(M.zone, p, (size_t)nz * 48); p += (size_t)nz * 48;
  M.n_static = 0;
  for (int g = 0; g < ng; g++) if (M.geom[g].is_static) M.static_ids[M.n_static++] = g;
  for (int g = 0; g < ng; g++) if (M.geom[g].type == GEOM_MESH && (M.geom[g].padr < 0 || M.geom[g].padr + M.geom[g].pnum > np)) { delete h; return b2p::api_fail(B2P_ERR_MODEL, "b2p_render_create: plane range"); }
  RCK(cudaSetDevice(device));
  RCK(cudaDeviceGetAttribute(&h->sms, cudaDevAttrMultiProcessorCount, device));
  RCK(cudaMalloc(&h->d_model, sizeof(RModel))); RCK(cudaMemcpy(h->d_model, &M, sizeof(RModel), cudaMemcpyHostToDevice));
  RCK(cudaMalloc(&h->d_planes, (size_t)(np + 1) * 16)); RCK(cudaMemcpy(h->d_planes, p, (size_t)np * 16, cudaMemcpyHostToDevice));
  const size_t px = (size_t)W * H;
  RCK(cudaMalloc(&h->bg_rgb, px * 3)); RCK(cudaMalloc(&h->bg_depth, px * 4)); RCK(cudaMalloc(&h->bg_seg, px * 4)); RCK(cudaMalloc(&h->d_layers, px * 4 * NLAYER));
  RCK(cudaMalloc(&h->d_queue, sizeof(unsigned long long))); RCK(cudaMalloc(&h->d_stats, 16)); RCK(cudaMemset(h->d_stats, 0, 16));
  // static layer: ground, table, arm base, zones -- ray-cast once, then L2-resident
  FrameArgs a; memset(&a, 0, sizeof a);
  a.n_env = 1; a.background = 1; a.rgb = h->bg_rgb; a.depth = h->bg_depth; a.seg = h->bg_seg;
  int rc = launch_frame(h, a, 0);
  if (rc != B2P_OK) return rc;
  RCK(cudaDeviceSynchronize());
  RCK(cudaMemset(h->d_stats, 0, 16));
  *out = h;
  return B2P_OK;
}

int b2p_render_destroy(b2p_render_handle h) {
  if (!h) return B2P_OK;
  cudaSetDevice(h->device);
  cudaFree(h->d_model); cudaFree(h->d_planes); cudaFree(h->d_rects); cudaFree(h->bg_rgb); cudaFree(h->bg_depth); cudaFree(h->bg_seg); cudaFree(h->d_layers); cudaFree(h->d_queue); cudaFree(h->d_stats);
  delete h;
  return B2P_OK;
}

int b2p_render_query(b2p_render_handle h, int what) {
  if (!h) return B2P_ERR_ARG;
  switch (what) {
    case B2P_RQ_WIDTH: return h->hm.W;
    case B2P_RQ_HEIGHT: return h->hm.H;
    case B2P_RQ_NGEOM: return h->hm.ngeom;
    case B2P_RQ_LAUNCHES: return (int)h->launches;
    case B2P_RQ_TILES_COPIED: case B2P_RQ_TILES_CAST: {
      unsigned long long s[2] = {0, 0};
      cudaSetDevice(h->device); cudaDeviceSynchronize(); cudaMemcpy(s, h->d_stats, 16, cudaMemcpyDeviceToHost);
      return (int)(s[what == B2P_RQ_TILES_CAST] & 0x7fffffff);
    }
  }
  return B2P_ERR_ARG;
}

int b2p_render(b2p_render_handle h, int n_env, const void* geom_xpos, const void* geom_xmat, int pose_stride, int dtype, const float* prop_rgba, const void* prop_size,
               const uint8_t* prop_active, uint8_t* rgb, float* depth, int32_t* seg, void* stream) {
  if (!h || n_env < 1 || !geom_xpos || !geom_xmat || (dtype != B2P_F64 && dtype != B2P_F32)) return b2p::api_fail(B2P_ERR_ARG, "b2p_render: bad arguments");
  if (((uintptr_t)rgb | (uintptr_t)depth | (uintptr_t)seg) & 15) return b2p::api_fail(B2P_ERR_ARG, "b2p_render: output buffers must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  RCK(cudaSetDevice(h->device));
  if (h->rect_cap < n_env) {
    if (h->d_rects) { RCK(cudaStreamSynchronize(st)); RCK(cudaFree(h->d_rects)); }
    RCK(cudaMalloc(&h->d_rects, (size_t)n_env * RMAXG * sizeof(short4))); h->rect_cap = n_env;
  }
  FrameArgs a; memset(&a, 0, sizeof a);
  a.n_env = n_env; a.pose_stride = pose_stride; a.dtype = dtype; a.xpos = geom_xpos; a.xmat = geom_xmat;
  a.prop_rgba = prop_rgba; a.prop_size = prop_size; a.prop_active = prop_active; a.rects = h->d_rects;
  a.rgb = rgb; a.depth = depth; a.seg = (int*)seg;
  const int n = n_env * RMAXG;
  render_bounds_kernel<<<(n + 255) / 256, 256, 0, st>>>(h->d_model, a, h->d_rects);
  RCK(cudaGetLastError());
  h->launches++;
  return launch_frame(h, a, st);
}

}  // extern "C"
