/*
 * oracle/mc_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU marching cubes over a sparse voxel set {key -> (tsdf, weight)}.
 * PARITY UNPINNED: the reference has no mesh extraction at all (SURVEY 8(a)
 * A11, 8(c)); this file is the in-repo statement of the semantics the CUDA
 * kernel (csrc/mc.cu) is checked against:
 *   - a cube is rooted at every voxel key k, corners k + (c&1, c>>1&1, c>>2&1);
 *   - it is emitted only if all 8 corners exist with weight >= min_weight
 *     (validity rule of global_volume.py:99,119);
 *   - inside = tsdf < 0; case tables from oracle/gen_mc_tables.py;
 *   - vertex on edge (a,b), a < b: p = pa + (pb - pa) * ta / (ta - tb), in
 *     float32, corner position = (float)key * (float)voxel_size (voxel-corner
 *     convention: keys come from floor(p / voxel_size)).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define MC_TABLE_QUAL static const
#include "../mast3r_slam_quality_dualtsdf_b200/csrc/mc_tables.h"

typedef struct {
  int64_t k[3];
  int64_t idx;
} mc_ent;

static uint64_t mc_hash(const int64_t k[3]) {
  uint64_t h = (uint64_t)k[0] * 0x9E3779B97F4A7C15ull;
  h ^= ((uint64_t)k[1] + 0x7F4A7C15ull) * 0xC2B2AE3D27D4EB4Full;
  h = (h << 31) | (h >> 33);
  h ^= ((uint64_t)k[2] + 0x165667B1ull) * 0xD6E8FEB86659FD93ull;
  h ^= h >> 29;
  h *= 0xBF58476D1CE4E5B9ull;
  h ^= h >> 32;
  return h;
}

static int64_t mc_find(const mc_ent *tab, uint64_t mask, const int64_t k[3]) {
  uint64_t i = mc_hash(k) & mask;
  for (;;) {
    if (tab[i].idx < 0) return -1;
    if (tab[i].k[0] == k[0] && tab[i].k[1] == k[1] && tab[i].k[2] == k[2]) return tab[i].idx;
    i = (i + 1) & mask;
  }
}

/* Returns the triangle count; writes up to tri_cap triangles (9 floats each)
 * when verts != NULL. */
int64_t orc_marching_cubes(const int64_t *keys, const float *tsdf, const float *weight,
                           int64_t n, double voxel_size, double min_weight, float *verts,
                           int64_t tri_cap) {
  uint64_t cap = 16;
  while (cap < (uint64_t)n * 2 + 2) cap <<= 1;
  mc_ent *tab = (mc_ent *)malloc(cap * sizeof(mc_ent));
  for (uint64_t i = 0; i < cap; ++i) tab[i].idx = -1;
  for (int64_t i = 0; i < n; ++i) {
    uint64_t j = mc_hash(keys + 3 * i) & (cap - 1);
    while (tab[j].idx >= 0) j = (j + 1) & (cap - 1);
    tab[j].k[0] = keys[3 * i], tab[j].k[1] = keys[3 * i + 1], tab[j].k[2] = keys[3 * i + 2];
    tab[j].idx = i;
  }
  const float vs = (float)voxel_size, mw = (float)min_weight;
  int64_t ntri = 0;
  for (int64_t i = 0; i < n; ++i) {
    float t[8];
    int64_t ck[8][3];
    int ok = 1, mask = 0;
    for (int c = 0; c < 8 && ok; ++c) {
      ck[c][0] = keys[3 * i] + (c & 1);
      ck[c][1] = keys[3 * i + 1] + ((c >> 1) & 1);
      ck[c][2] = keys[3 * i + 2] + ((c >> 2) & 1);
      int64_t j = mc_find(tab, cap - 1, ck[c]);
      if (j < 0 || !(weight[j] >= mw)) {
        ok = 0;
        break;
      }
      t[c] = tsdf[j];
      if (t[c] < 0.0f) mask |= 1 << c;
    }
    if (!ok) continue;
    int nt = MC_NUM_TRIS[mask];
    for (int q = 0; q < nt; ++q) {
      if (verts && ntri < tri_cap) {
        for (int vtx = 0; vtx < 3; ++vtx) {
          int e = MC_TRI_TABLE[mask][3 * q + vtx];
          int a = MC_EDGE_CORNERS[e][0], b = MC_EDGE_CORNERS[e][1];
          float mu = t[a] / (t[a] - t[b]);
          for (int d = 0; d < 3; ++d) {
            float pa = (float)ck[a][d] * vs, pb = (float)ck[b][d] * vs;
            verts[9 * ntri + 3 * vtx + d] = pa + (pb - pa) * mu;
          }
        }
      }
      ntri++;
    }
  }
  free(tab);
  return ntri;
}
