/*
 * oracle/march_key_check.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU checker of the product's one-FMA voxel keys (csrc/march_key.h): for
 * random and adversarial rays it enumerates every in-band sample with the
 * reference operation (global_volume.py:54-70, :133-134, float32 as in
 * oracle/tsdf_oracle.c) and verifies that whenever the fast path accepts a
 * sample (max |u| < lim) its three keys equal the reference keys.  Returns the
 * number of mismatches (must be 0) and the accepted / total sample counts.
 * Build: g++ -O2 -ffp-contract=off -fno-fast-math -mfma (tests/test_march_key.py).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

#include "../mast3r_slam_quality_dualtsdf_b200/csrc/march_key.h"

static float norm3f(float x, float y, float z) {
  float px = x * x, py = y * y, pz = z * z;
  double acc = (double)px;
  acc += (double)py;
  acc += (double)pz;
  return sqrtf((float)acc);
}

/* xorshift64* */
static uint64_t rng_next(uint64_t *s) {
  uint64_t x = *s;
  x ^= x >> 12, x ^= x << 25, x ^= x >> 27;
  *s = x;
  return x * 0x2545F4914F6CDD1Dull;
}
static double rng_u(uint64_t *s) { return (double)(rng_next(s) >> 11) * (1.0 / 9007199254740992.0); }

extern "C" int64_t march_key_check(int64_t n_rays, uint64_t seed, double voxel, double trunc_d,
                                   double step_scale, double room_half, int mode,
                                   int64_t *out_total, int64_t *out_accepted) {
  const float vs = (float)voxel, trunc = (float)trunc_d;
  double step_d = voxel * step_scale;
  if (step_d < 1.0e-4) step_d = 1.0e-4;
  const float step = (float)step_d;
  volatile float ivs_v = 1.0f / vs;
  const float ivs = ivs_v;
  uint64_t st = seed * 0x9E3779B97F4A7C15ull + 1;
  int64_t bad = 0, total = 0, acc = 0;
  for (int64_t r = 0; r < n_rays; ++r) {
    float o[3], p[3];
    for (int a = 0; a < 3; ++a) {
      double oo = (rng_u(&st) - 0.5) * room_half;
      double pp = (rng_u(&st) * 2.0 - 1.0) * room_half;
      if (mode == 1) { /* origin and surface points on voxel faces / centres */
        oo = floor(oo / voxel) * voxel + (rng_next(&st) & 1 ? 0.0 : 0.5 * voxel);
        pp = floor(pp / voxel) * voxel;
      }
      o[a] = (float)oo;
      p[a] = (float)pp;
    }
    if (mode == 2) { /* axis-aligned rays onto a wall that lies on a voxel face */
      int ax = (int)(rng_next(&st) % 3);
      for (int a = 0; a < 3; ++a)
        if (a != ax) p[a] = o[a];
      p[ax] = (float)((rng_next(&st) & 1 ? 1.0 : -1.0) * floor(room_half / voxel) * voxel);
    }
    if (mode == 3) { /* walls of the room: one coordinate exactly +-half */
      int ax = (int)(rng_next(&st) % 3);
      p[ax] = (float)((rng_next(&st) & 1 ? 1.0 : -1.0) * room_half);
    }
    float rx = p[0] - o[0], ry = p[1] - o[1], rz = p[2] - o[2];
    float len = norm3f(rx, ry, rz);
    if (!isfinite(len) || len < 1.0e-4f) continue;
    float d[3] = {rx / len, ry / len, rz / len};
    float maxd = len + trunc;
    float q = maxd / step;
    int64_t n = (int64_t)q;
    if (n < 1) n = 1;
    float lstep = 0.0f;
    if (n > 1) lstep = maxd / (float)(n - 1);
    MarchKeyLine kl;
    march_key_line(kl, o, d, lstep, maxd, ivs);
    for (int64_t i = 0; i < n; ++i) {
      float dist;
      if (n == 1)
        dist = 0.0f;
      else if (i == n - 1)
        dist = maxd;
      else
        dist = (float)i * lstep;
      float sdf = len - dist;
      if (fabsf(sdf) > trunc) continue;
      volatile float mx = dist * d[0], my = dist * d[1], mz = dist * d[2];
      volatile float sx = o[0] + mx, sy = o[1] + my, sz = o[2] + mz;
      volatile float qx = sx / vs, qy = sy / vs, qz = sz / vs;
      int64_t key[3] = {(int64_t)floorf(qx), (int64_t)floorf(qy), (int64_t)floorf(qz)};
      int raw[3];
      float m = march_key_eval(kl, (float)i, raw);
      total++;
      if (m < kl.lim) {
        acc++;
        for (int a = 0; a < 3; ++a)
          if ((int64_t)(raw[a] - MARCH_KEY_MAGIC_BITS) != key[a]) bad++;
      }
    }
  }
  *out_total = total;
  *out_accepted = acc;
  return bad;
}
