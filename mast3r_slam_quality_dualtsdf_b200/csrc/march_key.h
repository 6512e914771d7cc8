// march_key.h -- one-FMA voxel keys of the float32 ray march, shared by the
// CUDA kernels and the CPU checker (tests/test_march_key.py compiles this file
// with gcc and compares it with the reference operation on millions of rays).
//
// Reference operation (mast3r_slam/tsdf/global_volume.py:63-64, :133-134, in
// the float32 types numpy >= 2 uses):
//     dist_i = fl(fl(i) * lstep)           (np.linspace; last sample = maxd)
//     s_a    = fl(o_a + fl(dist_i * d_a))
//     key_a  = floor(fl(s_a / vs))
// Fast path: qm_a = fma(fl(i), G_a, H_a) with
//     G_a = fl(fl(lstep * d_a) * ivs),  H_a = fma(o_a, ivs, -0.5),  ivs = fl(1 / vs)
// approximates Q_a - 0.5 (Q_a = s_a / vs), kf = rint(qm_a) by the 1.5 * 2^23
// magic constant, u = qm_a - kf.  Error budget in voxels, u24 = 2^-24, with
// Pv = maxd |d_a| / vs >= |dist_i d_a| / vs,  Ov = |o_a| / vs,  Sv <= Ov + Pv:
//   reference chain vs the real number Q:  dist u24 Pv, product u24 Pv, sum
//     u24 Sv, quotient u24 Sv, last sample (maxd instead of fl((n-1) lstep))
//     u24 Pv                                              <= u24 (3 Pv + 2 Sv)
//   fast path vs Q - 0.5:  G (3 roundings) 3 u24 Pv, H (ivs + fma) u24 (2 Ov
//     + 1), final fma u24 (Sv + 1)                   <= u24 (3 Pv + 2 Ov + Sv + 2)
//   together                      <= u24 (6 Pv + 3 Sv + 2 Ov + 2) <= u24 (9 Pv + 5 Ov + 2)
// So if |u| < 0.5 - E with E >= that bound, fl(s_a / vs) lies strictly inside
// (kf, kf + 1) and floor() of it is kf: the key is exact.  Samples that fail the
// test (~1e-3) take the reference operation itself.  MARCH_KEY_SAFETY widens
// E by 25 % for the second-order terms ((1 + u24)^k, |d_a| <= 1 + u24).
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define MK_HD __host__ __device__ __forceinline__
#else
#define MK_HD static inline
#endif

#define MARCH_KEY_MAGIC 12582912.0f   /* 1.5 * 2^23 */
#define MARCH_KEY_MAGIC_BITS 0x4B400000
#define MARCH_KEY_U24 5.9604644775390625e-8f
#define MARCH_KEY_SAFETY 1.25f

#if defined(__CUDA_ARCH__)
#define MK_FMA(a, b, c) __fmaf_rn((a), (b), (c))
#define MK_MUL(a, b) __fmul_rn((a), (b))
#define MK_ADD(a, b) __fadd_rn((a), (b))
#define MK_SUB(a, b) __fsub_rn((a), (b))
#define MK_BITS(x) __float_as_int(x)
#else
MK_HD float mk_hmul(float a, float b) { volatile float r = a * b; return r; }
MK_HD float mk_hadd(float a, float b) { volatile float r = a + b; return r; }
MK_HD int mk_hbits(float x) { int i; memcpy(&i, &x, 4); return i; }
#define MK_FMA(a, b, c) fmaf((a), (b), (c))
#define MK_MUL(a, b) mk_hmul((a), (b))
#define MK_ADD(a, b) mk_hadd((a), (b))
#define MK_SUB(a, b) mk_hadd((a), -(b))
#define MK_BITS(x) mk_hbits(x)
#endif

struct MarchKeyLine {
  float g[3], h[3]; // qm_a(i) = fma(fl(i), g[a], h[a])
  float lim;        // key exact when max_a |u_a| < lim; <= 0: never (take the reference operation)
};

// Error bound E (voxels) of the fast keys of one ray.  omax_v = max_a |o_a| * ivs.
MK_HD float march_key_err(float maxd, float ivs, float omax_v) {
  return MARCH_KEY_SAFETY * MARCH_KEY_U24 * (9.0f * maxd * ivs + 5.0f * omax_v + 2.0f);
}

MK_HD void march_key_line(MarchKeyLine &k, const float o[3], const float d[3], float lstep,
                          float maxd, float ivs) {
  float om = 0.f;
  for (int a = 0; a < 3; ++a) {
    k.g[a] = MK_MUL(MK_MUL(lstep, d[a]), ivs);
    k.h[a] = MK_FMA(o[a], ivs, -0.5f);
    om = fmaxf(om, fabsf(o[a]));
  }
  const float e = march_key_err(maxd, ivs, om * ivs);
  // the magic-constant rint needs |qm| < 2^22; keep well inside
  const float qmax = (om + maxd) * ivs;
  k.lim = (qmax < 2.0e6f && e < 0.25f) ? 0.5f - e : -1.0f;
}

// Raw keys (float bit patterns: key = raw - MARCH_KEY_MAGIC_BITS) of sample fi.
// Returns max_a |u_a| (compare with k.lim).
MK_HD float march_key_eval(const MarchKeyLine &k, float fi, int raw[3]) {
  float m = 0.f;
  for (int a = 0; a < 3; ++a) {
    const float qm = MK_FMA(fi, k.g[a], k.h[a]);
    const float y = MK_ADD(qm, MARCH_KEY_MAGIC);
    raw[a] = MK_BITS(y);
    const float u = MK_SUB(qm, MK_SUB(y, MARCH_KEY_MAGIC));
    m = fmaxf(m, fabsf(u));
  }
  return m;
}
