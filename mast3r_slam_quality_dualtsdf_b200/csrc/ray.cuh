// ray.cuh -- exact restatement of the reference ray-march sample positions.
//
// Follows TSDFVolume.integrate (mast3r_slam/tsdf/global_volume.py:53-70) and
// _world_to_voxel (:133-134) operation by operation, in the dtype numpy >= 2
// uses for the inputs the SLAM loop feeds (float32 points/origin:
// global_manager.py:99-106), or all-float64 when the caller passes float64.
// Voxel keys must be bit-exact, so every operation on the key path is an
// explicitly rounded intrinsic (no FMA contraction, true IEEE division).
// SURVEY.md appendix B gives the np.linspace recipe mirrored here.
#pragma once
#include <math.h>

template <typename T>
struct Real;

template <>
struct Real<float> {
  static __device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
  static __device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
  static __device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
  static __device__ __forceinline__ float div(float a, float b) { return __fdiv_rn(a, b); }
  static __device__ __forceinline__ float flr(float a) { return floorf(a); }
  static __device__ __forceinline__ float abs(float a) { return fabsf(a); }
  static __device__ __forceinline__ float fromi(int i) { return __int2float_rn(i); }
  // np.linalg.norm(float32[3]) = sqrt(sdot): x86-64 OpenBLAS rounds each
  // product to float32 and accumulates in double (oracle/tsdf_oracle.c:135-146)
  static __device__ __forceinline__ float norm3(float x, float y, float z) {
    double acc = (double)__fmul_rn(x, x);
    acc = __dadd_rn(acc, (double)__fmul_rn(y, y));
    acc = __dadd_rn(acc, (double)__fmul_rn(z, z));
    return __fsqrt_rn(__double2float_rn(acc));
  }
  static __device__ __forceinline__ float tiny() { return 1.0e-4f; }
};

template <>
struct Real<double> {
  static __device__ __forceinline__ double mul(double a, double b) { return __dmul_rn(a, b); }
  static __device__ __forceinline__ double add(double a, double b) { return __dadd_rn(a, b); }
  static __device__ __forceinline__ double sub(double a, double b) { return __dsub_rn(a, b); }
  static __device__ __forceinline__ double div(double a, double b) { return __ddiv_rn(a, b); }
  static __device__ __forceinline__ double flr(double a) { return floor(a); }
  static __device__ __forceinline__ double abs(double a) { return fabs(a); }
  static __device__ __forceinline__ double fromi(int i) { return (double)i; }
  // np.linalg.norm(float64[3]) = sqrt(ddot), FMA-contracted tail loop
  // (oracle/tsdf_oracle.c:204-212)
  static __device__ __forceinline__ double norm3(double x, double y, double z) {
    double acc = __dmul_rn(x, x);
    acc = __fma_rn(y, y, acc);
    acc = __fma_rn(z, z, acc);
    return __dsqrt_rn(acc);
  }
  static __device__ __forceinline__ double tiny() { return 1.0e-4; }
};

template <typename T>
struct Ray {
  T ox, oy, oz, dx, dy, dz, len, maxd, lstep;
  int n;    // np.linspace sample count           (global_volume.py:60-61)
  int i_lo; // first sample index that can lie inside the +-trunc band
};

// global_volume.py:54-56: ray = p - o, len = ||ray||.  Returns false when the
// reference `continue`s (:56).
template <typename T>
__device__ __forceinline__ bool ray_length(Ray<T> &r, T px, T py, T pz, T ox, T oy, T oz, T &rx,
                                           T &ry, T &rz) {
  using R = Real<T>;
  rx = R::sub(px, ox), ry = R::sub(py, oy), rz = R::sub(pz, oz);
  T len = R::norm3(rx, ry, rz);
  r.len = len;
  return isfinite(len) && !(len < R::tiny());
}

// global_volume.py:57-61: direction, sample count, np.linspace step.
template <typename T>
__device__ __forceinline__ void ray_finish(Ray<T> &r, T rx, T ry, T rz, T ox, T oy, T oz, T step,
                                           T trunc) {
  using R = Real<T>;
  const T len = r.len;
  r.ox = ox, r.oy = oy, r.oz = oz;
  r.dx = R::div(rx, len), r.dy = R::div(ry, len), r.dz = R::div(rz, len);
  r.maxd = R::add(len, trunc);
  T q = R::div(r.maxd, step);
  int n = (q < (T)1073741824.0) ? (int)q : 1073741824; // int() truncates
  if (n < 1) n = 1;
  r.n = n;
  r.lstep = (n > 1) ? R::div(r.maxd, R::fromi(n - 1)) : (T)0;
  // band: |len - dist_i| <= trunc  <=>  dist_i >= len - trunc (upper side is
  // dist_i <= maxd, true for every i).  Conservative start, exact test later.
  int lo = 0;
  T below = len - trunc;
  if (n > 1 && below > (T)0) {
    T f = below / r.lstep;
    lo = (int)f - 2;
    if (lo < 0) lo = 0;
    if (lo > n - 1) lo = n - 1;
  }
  r.i_lo = lo;
}

template <typename T>
__device__ __forceinline__ bool ray_setup(Ray<T> &r, T px, T py, T pz, T ox, T oy, T oz, T step,
                                          T trunc) {
  T rx, ry, rz;
  if (!ray_length<T>(r, px, py, pz, ox, oy, oz, rx, ry, rz)) return false;
  ray_finish<T>(r, rx, ry, rz, ox, oy, oz, step, trunc);
  return true;
}

// One sample of the ray (global_volume.py:63-69 + :133-134).  Returns false
// when the reference skips it (abs(sdf) > trunc).  t = clipped tsdf,
// a = -abs(sdf)/trunc (argument of the weight exponential).
template <typename T>
__device__ __forceinline__ bool ray_sample(const Ray<T> &r, int i, T trunc, T vs, int &kx,
                                           int &ky, int &kz, T &t, T &a) {
  using R = Real<T>;
  T dist;
  if (r.n == 1)
    dist = (T)0;
  else if (i == r.n - 1)
    dist = r.maxd;
  else
    dist = R::mul(R::fromi(i), r.lstep);
  T sdf = R::sub(r.len, dist);
  T asdf = R::abs(sdf);
  if (asdf > trunc) return false;
  T sx = R::add(r.ox, R::mul(dist, r.dx));
  T sy = R::add(r.oy, R::mul(dist, r.dy));
  T sz = R::add(r.oz, R::mul(dist, r.dz));
  t = R::div(sdf, trunc);
  t = t < (T)-1 ? (T)-1 : (t > (T)1 ? (T)1 : t);
  a = -R::div(asdf, trunc);
  T fx = R::flr(R::div(sx, vs)), fy = R::flr(R::div(sy, vs)), fz = R::flr(R::div(sz, vs));
  const T lim = (T)2.0e9;
  fx = fx < -lim ? -lim : (fx > lim ? lim : fx);
  fy = fy < -lim ? -lim : (fy > lim ? lim : fy);
  fz = fz < -lim ? -lim : (fz > lim ? lim : fz);
  kx = (int)fx, ky = (int)fy, kz = (int)fz;
  return true;
}
