// Host check of the per-tile plane lists of the ray-caster (csrc/b2p_render_cull.h, the code the kernel runs): for random poses of
// every hull of the render model and the tiles around its image, the clip loop over the tile's list must return the very bits of
// the loop over all planes -- hit / miss, entry parameter, entry facet -- on every pixel of the tile.
//   b2p_render_cull_test <render blob> <seed> <poses per hull>      -> one line: rays hits mismatches planes kept
// Built by tests/test_render_cull.py with g++ -O2 -ffp-contract=off (the kernel's translation unit is compiled with -fmad=false).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../mujoco_robot_environments_b200/csrc/b2p_render_cull.h"

using namespace b2p_rcull;

struct RGeom { int type, padr, pnum, prop, is_static, pad_i[3]; float rgba[4], size[3], cen[3], rbound, spos[3], smat[9], pad_f; };
struct RCam { float pos[3], mat[9], f, cx, cy, zfar, hl_amb, hl_diff, lpos[3], ldiff, rgb2[3], check, pad[6]; };
static_assert(sizeof(RGeom) == 128 && sizeof(RCam) == 128, "B2PR record sizes");

static unsigned long long rng_state;
static double urand() { rng_state = rng_state * 6364136223846793005ULL + 1442695040888963407ULL; return (double)(rng_state >> 11) / 9007199254740992.0; }

int main(int argc, char** argv) {
  if (argc < 4) return 2;
  FILE* f = fopen(argv[1], "rb"); if (!f) return 3;
  std::vector<char> blob; { char buf[65536]; size_t n; while ((n = fread(buf, 1, sizeof buf, f)) > 0) blob.insert(blob.end(), buf, buf + n); } fclose(f);
  rng_state = 88172645463325252ULL + (unsigned long long)atoll(argv[2]) * 7919ULL;
  const int poses = atoi(argv[3]);
  const int* hd = (const int*)blob.data();
  const int ng = hd[2], nz = hd[4], W = hd[5], H = hd[6];
  RCam C; memcpy(&C, blob.data() + 64, 128);
  std::vector<RGeom> G(ng); memcpy(G.data(), blob.data() + 64 + 128, (size_t)ng * 128);
  const Plane4* planes = (const Plane4*)(blob.data() + 64 + 128 + (size_t)ng * 128 + (size_t)nz * 48);
  const int TW = 32, TH = 8;
  long long rays = 0, hits = 0, bad = 0, planes_all = 0, planes_kept = 0, tiles = 0, hit_tiles = 0, hit_planes = 0, hit_kept = 0, dropped = 0, empty_tiles = 0;
  for (int g = 0; g < ng; g++) {
    if (G[g].type != 7 || G[g].pnum > 128) continue;
    const Plane4* pl = planes + G[g].padr; const int np = G[g].pnum;
    for (int it = 0; it < poses; it++) {
      // a pose whose hull centre projects to a random pixel at a random depth, random orientation
      const float uc = (float)(urand() * W), vc = (float)(urand() * H), depth = 0.25f + (float)urand() * 1.1f;
      double q[4], nq = 0; for (int k = 0; k < 4; k++) { q[k] = urand() * 2 - 1; nq += q[k] * q[k]; }
      nq = 1.0 / sqrt(nq); for (int k = 0; k < 4; k++) q[k] *= nq;
      float mat[9] = {(float)(1 - 2 * (q[2] * q[2] + q[3] * q[3])), (float)(2 * (q[1] * q[2] - q[0] * q[3])), (float)(2 * (q[1] * q[3] + q[0] * q[2])),
                      (float)(2 * (q[1] * q[2] + q[0] * q[3])), (float)(1 - 2 * (q[1] * q[1] + q[3] * q[3])), (float)(2 * (q[2] * q[3] - q[0] * q[1])),
                      (float)(2 * (q[1] * q[3] - q[0] * q[2])), (float)(2 * (q[2] * q[3] + q[0] * q[1])), (float)(1 - 2 * (q[1] * q[1] + q[2] * q[2]))};
      const float dc0 = (uc - C.cx) / C.f, dc1 = (C.cy - vc) / C.f;
      float ctr[3], pos[3];
      for (int k = 0; k < 3; k++) ctr[k] = C.pos[k] + depth * (C.mat[3 * k] * dc0 + C.mat[3 * k + 1] * dc1 - C.mat[3 * k + 2]);
      for (int k = 0; k < 3; k++) pos[k] = ctr[k] - (mat[3 * k] * G[g].cen[0] + mat[3 * k + 1] * G[g].cen[1] + mat[3 * k + 2] * G[g].cen[2]);
      const int tcx = (int)uc / TW, tcy = (int)vc / TH;
      for (int dty = -6; dty <= 6; dty++) for (int dtx = -2; dtx <= 2; dtx++) {
        const int tx = tcx + dtx, ty = tcy + dty;
        if (tx < 0 || ty < 0 || tx >= W / TW || ty >= H / TH) continue;
        const int tx0 = tx * TW, ty0 = ty * TH;
        // ---- the tile's list, composed exactly as the kernel composes it (lanes -> a serial loop)
        TileRays R; tile_rays(C.pos, C.mat, C.f, C.cx, C.cy, tx0, ty0, TW, TH, pos, mat, R);
        int cls[128]; float tmn[128], tmx[128]; float e_lo = -3.0e38f, x_hi = 3.0e38f;
        for (int i = 0; i < np; i++) {
          tmn[i] = tmx[i] = 0.0f;
          cls[i] = plane_class(pl[i], R, &tmn[i], &tmx[i]);
          if (cls[i] == PL_ENTRY) e_lo = fmaxf(e_lo, tmn[i]);
          if (cls[i] == PL_EXIT) x_hi = fminf(x_hi, tmx[i]);
        }
        const bool drop = tile_misses(e_lo, x_hi);      // the kernel then skips the hull for the whole tile
        unsigned char list[128]; int cnt = 0;
        for (int i = 0; i < np; i++) if (plane_keep(cls[i], tmn[i], tmx[i], e_lo, x_hi)) list[cnt++] = (unsigned char)i;
        planes_all += np; planes_kept += cnt; tiles++;
        // ---- every pixel: full loop against listed loop
        long long tile_hits = 0;
        for (int v = ty0; v < ty0 + TH; v++) for (int u = tx0; u < tx0 + TW; u++) {
          const float c0 = ((float)u - C.cx) / C.f, c1 = (C.cy - (float)v) / C.f;
          float d[3], ol[3], dl[3];
          for (int k = 0; k < 3; k++) d[k] = C.mat[3 * k] * c0 + C.mat[3 * k + 1] * c1 - C.mat[3 * k + 2];
          const float w0 = C.pos[0] - pos[0], w1 = C.pos[1] - pos[1], w2 = C.pos[2] - pos[2];
          for (int k = 0; k < 3; k++) { ol[k] = mat[k] * w0 + mat[3 + k] * w1 + mat[6 + k] * w2; dl[k] = mat[k] * d[0] + mat[3 + k] * d[1] + mat[6 + k] * d[2]; }
          float t1 = 0, t2 = 0; int b1 = -1, b2 = -1;
          const bool h1 = hull_clip(ol, dl, pl, nullptr, np, &t1, &b1), h2 = hull_clip(ol, dl, pl, list, cnt, &t2, &b2);
          rays++; hits += h1; tile_hits += h1;
          if (h1 != h2 || (h1 && (memcmp(&t1, &t2, 4) != 0 || b1 != b2))) {
            if (bad < 5) fprintf(stderr, "mismatch geom %d pose %d pixel (%d, %d): full %d t %.9g facet %d, listed %d t %.9g facet %d\n", g, it, u, v, (int)h1, t1, b1, (int)h2, t2, b2);
            bad++;
          }
        }
        if (tile_hits) { hit_tiles++; hit_planes += np; hit_kept += cnt; } else empty_tiles++;
        if (drop) { dropped++; if (tile_hits) { if (bad < 5) fprintf(stderr, "geom %d pose %d tile (%d, %d) dropped with %lld hits\n", g, it, tx0, ty0, tile_hits); bad += tile_hits; } }
      }
    }
  }
  printf("{\"rays\": %lld, \"hits\": %lld, \"mismatches\": %lld, \"tiles\": %lld, \"planes\": %lld, \"kept\": %lld, \"hit_tiles\": %lld, \"hit_planes\": %lld, \"hit_kept\": %lld, \"empty_tiles\": %lld, \"dropped_tiles\": %lld}\n", rays, hits, bad, tiles, planes_all, planes_kept, hit_tiles, hit_planes, hit_kept, empty_tiles, dropped);
  return 0;
}
