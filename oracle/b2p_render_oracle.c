/* CPU restatement of the overhead-camera ray-cast (TEST INFRASTRUCTURE ONLY: tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline leg may load this; the product path is csrc/b2p_render.cu and never calls it).
 *
 * What it restates: the observation the reference gets from mujoco.Renderer on "overhead_camera/overhead_camera"
 * (mujoco_robot_environments/tasks/rearrangement.py:182-186, :460-478 rgb + depth, :273-275 segmentation) for the scene of
 * rearrangement.py:54-210, with the pinhole model of `_get_camera_intrinsics` / `_get_camera_extrinsics` (:480-505):
 * pixel (u, v) looks along  xmat * ((u - cx) / f, (cy - v) / f, -1).  The reference rasterises through OpenGL (unavailable on
 * this box), so parity with GL images is unpinned; what this file pins is the CUDA kernel: same single-precision IEEE
 * arithmetic in the same order (compiled with -ffp-contract=off), brute force over every geom for every pixel -- no tiles, no
 * bounding rectangles, no AABB pre-test, no background cache -- so any culling error in the kernel shows up as a pixel diff.
 *
 * Blob layout: include/b2p_render.h.
 */
#include <math.h>
#include <stddef.h>
#include <string.h>

enum { G_PLANE = 0, G_BOX = 6, G_MESH = 7 };
typedef struct { int type, padr, pnum, prop, is_static, pad_i[3]; float rgba[4], size[3], cen[3], rbound, spos[3], smat[9], pad_f; } geom_t;
typedef struct { float pos[3], half[3], rgba[4], pad[2]; } zone_t;
typedef struct { float pos[3], mat[9], f, cx, cy, zfar, hl_amb, hl_diff, lpos[3], ldiff, rgb2[3], check, pad[6]; } cam_t;

static void to_local(const float* o, const float* d, const float* pos, const float* mat, float* ol, float* dl) {
  float w0 = o[0] - pos[0], w1 = o[1] - pos[1], w2 = o[2] - pos[2];
  for (int k = 0; k < 3; k++) {
    ol[k] = mat[k] * w0 + mat[3 + k] * w1 + mat[6 + k] * w2;
    dl[k] = mat[k] * d[0] + mat[3 + k] * d[1] + mat[6 + k] * d[2];
  }
}

static int hit_box(const float* o, const float* d, const float* pos, const float* mat, const float* half, float* t_out, float* n) {
  float ol[3], dl[3], tn = -1e30f, tf = 1e30f, sg = 0.0f;
  int ax = 0;
  to_local(o, d, pos, mat, ol, dl);
  for (int k = 0; k < 3; k++) {
    if (dl[k] != 0.0f) {
      float inv = 1.0f / dl[k];
      float t1 = (-half[k] - ol[k]) * inv, t2 = (half[k] - ol[k]) * inv;
      float s = dl[k] > 0.0f ? -1.0f : 1.0f;
      if (t1 > t2) { float q = t1; t1 = t2; t2 = q; }
      if (t1 > tn) { tn = t1; ax = k; sg = s; }
      if (t2 < tf) tf = t2;
    } else if (fabsf(ol[k]) > half[k]) return 0;
  }
  if (tn > tf || tn <= 0.0f) return 0;
  *t_out = tn;
  n[0] = sg * mat[ax]; n[1] = sg * mat[3 + ax]; n[2] = sg * mat[6 + ax];
  return 1;
}

static int hit_hull(const float* o, const float* d, const float* pos, const float* mat, const float* planes, int np, float* t_out, float* n) {
  /* entry / exit parameters as fractions num / den (entry den < 0, exit den > 0), compared by cross multiplication */
  float ol[3], dl[3], tn_n = 1e30f, tn_d = -1.0f, tf_n = 1e30f, tf_d = 1.0f;
  int best = -1;
  to_local(o, d, pos, mat, ol, dl);
  for (int i = 0; i < np; i++) {
    const float* P = planes + 4 * i;
    /* the two plane dot products are fused multiply-adds (correctly rounded fmaf), as in the kernel */
    float den = fmaf(P[2], dl[2], fmaf(P[1], dl[1], P[0] * dl[0]));
    float dist = P[3] - fmaf(P[2], ol[2], fmaf(P[1], ol[1], P[0] * ol[0]));
    if (den == 0.0f) { if (dist < 0.0f) return 0; continue; }
    if (den < 0.0f) { if (dist * tn_d > tn_n * den) { tn_n = dist; tn_d = den; best = i; } }
    else if (dist * tf_d < tf_n * den) { tf_n = dist; tf_d = den; }
    if (tn_n * tf_d < tf_n * tn_d) return 0;
  }
  if (best < 0) return 0;
  float tn = tn_n / tn_d;
  if (tn <= 0.0f) return 0;
  *t_out = tn;
  for (int k = 0; k < 3; k++) n[k] = mat[3 * k] * planes[4 * best] + mat[3 * k + 1] * planes[4 * best + 1] + mat[3 * k + 2] * planes[4 * best + 2];
  return 1;
}

static int hit_ground(const float* o, const float* d, const float* pos, const float* mat, const float* size, float check, const float* rgb1, const float* rgb2,
                      float* t_out, float* n, float* rgb) {
  float ol[3], dl[3];
  to_local(o, d, pos, mat, ol, dl);
  if (!(dl[2] < 0.0f) || !(ol[2] > 0.0f)) return 0;
  float t = -ol[2] / dl[2];
  float x = ol[0] + t * dl[0], y = ol[1] + t * dl[1];
  if (fabsf(x) > size[0] || fabsf(y) > size[1]) return 0;
  int cell = (int)floorf(x / check) + (int)floorf(y / check);
  const float* c = (cell & 1) ? rgb2 : rgb1;
  *t_out = t; n[0] = mat[2]; n[1] = mat[5]; n[2] = mat[8];
  rgb[0] = c[0]; rgb[1] = c[1]; rgb[2] = c[2];
  return 1;
}

static float shade(const cam_t* C, const float* p, const float* n) {
  float hl = fmaxf(0.0f, n[0] * C->mat[2] + n[1] * C->mat[5] + n[2] * C->mat[8]);
  float L0 = C->lpos[0] - p[0], L1 = C->lpos[1] - p[1], L2 = C->lpos[2] - p[2];
  float inv = 1.0f / sqrtf(L0 * L0 + L1 * L1 + L2 * L2);
  float sl = fmaxf(0.0f, (n[0] * L0 + n[1] * L1 + n[2] * L2) * inv);
  return C->hl_amb + C->hl_diff * hl + C->ldiff * sl;
}
static float c01(float x) { return fminf(1.0f, fmaxf(0.0f, x)); }

/* One scene.  geom_xpos [ngeom][3], geom_xmat [ngeom][9] (float64, world), prop_rgba [nprop][4] float32, prop_size [nprop][3]
 * float64 half sizes, prop_active [nprop] (NULL = all).  Renders the window [x0, x1) x [y0, y1) into full-size images
 * rgb [H][W][3], depth [H][W], seg [H][W].  Returns 0, or -1 for a malformed blob. */
int b2p_render_oracle(const void* blob, size_t nbytes, const double* geom_xpos, const double* geom_xmat, const float* prop_rgba, const double* prop_size,
                      const unsigned char* prop_active, unsigned char* rgb, float* depth, int* seg, int x0, int y0, int x1, int y1) {
  const int* hd = (const int*)blob;
  if (nbytes < 192 || (unsigned)hd[0] != 0x52503242u || hd[1] != 1) return -1;
  const int ng = hd[2], np = hd[3], nz = hd[4], W = hd[5], H = hd[6], nprop = hd[7];
  if (nbytes < 192 + (size_t)ng * 128 + (size_t)nz * 48 + (size_t)np * 16) return -1;
  const cam_t* C = (const cam_t*)((const char*)blob + 64);
  const geom_t* G = (const geom_t*)((const char*)blob + 192);
  const zone_t* Z = (const zone_t*)((const char*)blob + 192 + (size_t)ng * 128);
  const float* planes = (const float*)((const char*)blob + 192 + (size_t)ng * 128 + (size_t)nz * 48);
  (void)nprop;
  if (x1 <= x0) { x0 = 0; x1 = W; }
  if (y1 <= y0) { y0 = 0; y1 = H; }
  const float ident[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  for (int v = y0; v < y1; v++) for (int u = x0; u < x1; u++) {
    float dc0 = ((float)u - C->cx) / C->f, dc1 = (C->cy - (float)v) / C->f;
    float d[3], o[3] = {C->pos[0], C->pos[1], C->pos[2]};
    for (int k = 0; k < 3; k++) d[k] = C->mat[3 * k] * dc0 + C->mat[3 * k + 1] * dc1 - C->mat[3 * k + 2];
    float best_t = C->zfar, bn[3] = {0, 0, 0}, brgb[3] = {0, 0, 0};
    int best_id = -1;
    for (int g = 0; g < ng; g++) {
      float pos[3], mat[9], size[3], col[3], t, n[3], rgbg[3];
      int hit = 0;
      if (G[g].is_static) { memcpy(pos, G[g].spos, 12); memcpy(mat, G[g].smat, 36); }
      else {
        if (G[g].prop >= 0 && prop_active && !prop_active[G[g].prop]) continue;
        for (int k = 0; k < 3; k++) pos[k] = (float)geom_xpos[3 * g + k];
        for (int k = 0; k < 9; k++) mat[k] = (float)geom_xmat[9 * g + k];
      }
      for (int k = 0; k < 3; k++) {
        size[k] = (!G[g].is_static && G[g].prop >= 0 && prop_size) ? (float)prop_size[3 * G[g].prop + k] : G[g].size[k];
        col[k] = (!G[g].is_static && G[g].prop >= 0 && prop_rgba) ? prop_rgba[4 * G[g].prop + k] : G[g].rgba[k];
      }
      if (G[g].type == G_BOX) hit = hit_box(o, d, pos, mat, size, &t, n);
      else if (G[g].type == G_MESH) hit = hit_hull(o, d, pos, mat, planes + 4 * G[g].padr, G[g].pnum, &t, n);
      else if (G[g].type == G_PLANE) hit = hit_ground(o, d, pos, mat, G[g].size, C->check, G[g].rgba, C->rgb2, &t, n, rgbg);
      if (hit && t < best_t) {
        best_t = t; best_id = g;
        for (int k = 0; k < 3; k++) { bn[k] = n[k]; brgb[k] = G[g].type == G_PLANE ? rgbg[k] : col[k]; }
      }
    }
    float out[3] = {0.0f, 0.0f, 0.0f};
    if (best_id >= 0) {
      float p[3] = {o[0] + best_t * d[0], o[1] + best_t * d[1], o[2] + best_t * d[2]};
      float s = shade(C, p, bn);
      for (int k = 0; k < 3; k++) out[k] = c01(c01(brgb[k]) * s);
    }
    float zt = best_t, zn[3] = {0, 0, 0};
    int zid = -1;
    for (int z = 0; z < nz; z++) {
      float t, n[3];
      if (hit_box(o, d, Z[z].pos, ident, Z[z].half, &t, n) && t < zt) { zt = t; zid = z; zn[0] = n[0]; zn[1] = n[1]; zn[2] = n[2]; }
    }
    if (zid >= 0) {
      float p[3] = {o[0] + zt * d[0], o[1] + zt * d[1], o[2] + zt * d[2]};
      float s = shade(C, p, zn), al = Z[zid].rgba[3];
      for (int k = 0; k < 3; k++) out[k] = al * c01(Z[zid].rgba[k] * s) + (1.0f - al) * out[k];
      best_t = zt; best_id = ng + zid;
    }
    size_t px = (size_t)v * W + u;
    if (rgb) for (int k = 0; k < 3; k++) rgb[3 * px + k] = (unsigned char)(out[k] * 255.0f + 0.5f);
    if (depth) depth[px] = best_t;
    if (seg) seg[px] = best_id;
  }
  return 0;
}
