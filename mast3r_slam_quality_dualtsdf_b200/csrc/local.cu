// local.cu -- the refiner's local (per-ROI, dense <= 64^3) TSDF on the GPU.
//
// Replaces TSDFRefiner._build_tsdf_robust (mast3r_slam/tsdf_refine.py:837-940),
// _extract_surface_safe (:942-1021) and _sample_tsdf_trilinear (:1023-1064).
// Discrete decisions (validity, sample count, sample positions, voxel index,
// zero crossing) follow the arithmetic torch executes in the reference
// (float32 tensor ops, float64 Python scalars; oracle/local_oracle.c states
// it): explicitly rounded intrinsics, fmaf where torch fuses.  The running
// mean of :924-930 has no clamp, i.e. it is sum(sdf*w)/sum(w): contributions
// are accumulated with float atomics (grid <= 2 MB, L2 resident) and divided
// once in a finalize pass (1e-5 tolerance instead of bit-exact values).
#include <math.h>

#include "common.cuh"

struct LocalParams {
  float lo[3], hi[3], av[3], o_unused;
  int nx, ny, nz;
  float voxel_f, min_conf_f;
  double voxel, trunc;
};

__device__ __forceinline__ float norm3_torch(float x, float y, float z) {
  float s = __fmul_rn(x, x);
  s = __fmaf_rn(y, y, s);
  s = __fmaf_rn(z, z, s);
  return __fsqrt_rn(s);
}

__device__ __forceinline__ float linspace_torch(float start, float end, int n, int i) {
  if (n == 1) return start;
  float step = __fdiv_rn(__fsub_rn(end, start), (float)(n - 1));
  if (i < n / 2) return __fmaf_rn(step, (float)i, start);
  return __fmaf_rn(-step, (float)(n - 1 - i), end);
}

__device__ __forceinline__ void sim3_act_local(const float *__restrict__ P, const float *x,
                                               float *y) {
  const float q0 = P[3], q1 = P[4], q2 = P[5], q3 = P[6], s = P[7];
  float uv0 = __fmul_rn(2.0f, __fsub_rn(__fmul_rn(q1, x[2]), __fmul_rn(q2, x[1])));
  float uv1 = __fmul_rn(2.0f, __fsub_rn(__fmul_rn(q2, x[0]), __fmul_rn(q0, x[2])));
  float uv2 = __fmul_rn(2.0f, __fsub_rn(__fmul_rn(q0, x[1]), __fmul_rn(q1, x[0])));
  float a0 = __fadd_rn(__fadd_rn(x[0], __fmul_rn(q3, uv0)),
                       __fsub_rn(__fmul_rn(q1, uv2), __fmul_rn(q2, uv1)));
  float a1 = __fadd_rn(__fadd_rn(x[1], __fmul_rn(q3, uv1)),
                       __fsub_rn(__fmul_rn(q2, uv0), __fmul_rn(q0, uv2)));
  float a2 = __fadd_rn(__fadd_rn(x[2], __fmul_rn(q3, uv2)),
                       __fsub_rn(__fmul_rn(q0, uv1), __fmul_rn(q1, uv0)));
  y[0] = __fadd_rn(__fmul_rn(a0, s), P[0]);
  y[1] = __fadd_rn(__fmul_rn(a1, s), P[1]);
  y[2] = __fadd_rn(__fmul_rn(a2, s), P[2]);
}

// S (in tsdf buffer) and W (in weights buffer) accumulate; ctr[0] = valid
// points, ctr[1] = fused points.
__global__ void __launch_bounds__(256)
    local_accumulate_kernel(LocalParams lp, const float *__restrict__ X,
                            const float *__restrict__ C, const float *__restrict__ pose,
                            long long n_pts, float *S, float *Wt, unsigned long long *ctr) {
  long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  bool ok = p < n_pts;
  float w[3] = {0, 0, 0};
  float c = 0.f;
  if (ok) {
    float x[3] = {X[3 * p], X[3 * p + 1], X[3 * p + 2]};
    sim3_act_local(pose, x, w); // :863
    c = C[p];
    ok = isfinite(w[0]) && isfinite(w[1]) && isfinite(w[2]);
#pragma unroll
    for (int a = 0; a < 3; ++a) ok = ok && w[a] >= lp.lo[a] && w[a] <= lp.hi[a];
    ok = ok && c > lp.min_conf_f; // :866-873
  }
  unsigned vm = __ballot_sync(0xFFFFFFFFu, ok);
  if ((threadIdx.x & 31) == 0 && vm) atomicAdd(ctr, (unsigned long long)__popc(vm));
  bool fused = false;
  if (ok) {
    const float o[3] = {pose[0], pose[1], pose[2]};
    const double conf = (double)c;
    float rv[3] = {__fsub_rn(w[0], o[0]), __fsub_rn(w[1], o[1]), __fsub_rn(w[2], o[2])};
    const double len = (double)norm3_torch(rv[0], rv[1], rv[2]); // :890
    const float den = (float)(len + 1.0e-8);                     // :891
    const float dir[3] = {__fdiv_rn(rv[0], den), __fdiv_rn(rv[1], den), __fdiv_rn(rv[2], den)};
    if (!(len < 0.05)) { // :893-894
      double t_start = len - lp.trunc * 2.0;
      if (t_start < 0.05) t_start = 0.05;
      const double t_end = len + lp.trunc * 2.0;
      int ns = (int)((t_end - t_start) / lp.voxel) + 1; // :899
      if (ns > 32) ns = 32;
      const int nd[3] = {lp.nx, lp.ny, lp.nz};
      for (int i = 0; i < ns; ++i) {
        const float t = linspace_torch((float)t_start, (float)t_end, ns, i);
        float s[3];
        bool out = false;
#pragma unroll
        for (int a = 0; a < 3; ++a) {
          s[a] = __fadd_rn(o[a], __fmul_rn(dir[a], t)); // :903
          out = out || s[a] < lp.lo[a] || s[a] > lp.hi[a];
        }
        if (out) continue; // :906-907
        int g[3];
#pragma unroll
        for (int a = 0; a < 3; ++a) { // :910-913
          float gp = __fdiv_rn(__fsub_rn(s[a], lp.lo[a]), lp.av[a]);
          float hi = (float)(nd[a] - 1);
          gp = gp < 0.0f ? 0.0f : (gp > hi ? hi : gp);
          g[a] = (int)gp;
        }
        double sdf = (len - (double)t) / lp.trunc; // :916-917
        sdf = sdf < -1.0 ? -1.0 : (sdf > 1.0 ? 1.0 : sdf);
        double dw = 1.0 - fabs(sdf);
        if (dw < 0.0) dw = 0.0;
        const double weight = conf * dw; // :920-921
        if (weight > 0.0) {
          const long long idx = ((long long)g[2] * lp.ny + g[1]) * lp.nx + g[0];
          atomicAdd(S + idx, (float)(sdf * weight));
          atomicAdd(Wt + idx, (float)weight);
        }
      }
      fused = true; // :932
    }
  }
  unsigned fm = __ballot_sync(0xFFFFFFFFu, fused);
  if ((threadIdx.x & 31) == 0 && fm) atomicAdd(ctr + 1, (unsigned long long)__popc(fm));
}

__global__ void local_finalize_kernel(long long nvox, float *S, float *Wt,
                                      const unsigned long long *ctr) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nvox) return;
  const bool enough = ctr[0] >= 5; // :877-878
  float w = Wt[i], s = S[i];
  if (enough && w > 1.0e-6f) { // :927-930
    S[i] = s / w;
  } else {
    S[i] = 1.0f; // :850-851 initial values
    Wt[i] = 0.0f;
  }
}

static void fill_params(LocalParams &lp, const float *box, const b200tsdf_local_cfg *cfg,
                        const int32_t dims[3]) {
  for (int a = 0; a < 3; ++a) {
    lp.lo[a] = box[a];
    lp.hi[a] = box[3 + a];
    lp.av[a] = (box[3 + a] - box[a]) / (float)dims[a]; // :880 / :958 actual_voxel
  }
  lp.nx = dims[0], lp.ny = dims[1], lp.nz = dims[2];
  lp.voxel = cfg->voxel_size;
  lp.trunc = cfg->trunc_dist;
  lp.voxel_f = (float)cfg->voxel_size;
  lp.min_conf_f = (float)cfg->min_confidence;
}

extern "C" int b200tsdf_local_dims(const float *box6_host, const b200tsdf_local_cfg *cfg,
                                   int32_t dims3_host[3]) {
  B200_NVTX("b200tsdf_local_dims");
  if (!box6_host || !cfg || !dims3_host || !(cfg->voxel_size > 0.0)) {
    b200_set_error("b200tsdf_local_dims: invalid argument");
    return B200TSDF_E_INVALID;
  }
  for (int a = 0; a < 3; ++a) { // :844-846
    float roi = box6_host[3 + a] - box6_host[a];
    long long g = (long long)ceilf(roi / (float)cfg->voxel_size);
    if (g > cfg->max_grid_dim) g = cfg->max_grid_dim;
    dims3_host[a] = (int32_t)g;
  }
  return B200TSDF_OK;
}

static unsigned long long *g_local_ctr[64] = {nullptr};

static int local_counters(unsigned long long **out) {
  int dev = 0;
  B200_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return B200TSDF_E_INVALID;
  if (!g_local_ctr[dev]) B200_CUDA(cudaMalloc(&g_local_ctr[dev], 2 * sizeof(unsigned long long)));
  *out = g_local_ctr[dev];
  return B200TSDF_OK;
}

extern "C" int b200tsdf_local_build(const float *X_dev, const float *C_dev, const float *pose_dev,
                                    const float *box6_host, int64_t n_pts,
                                    const b200tsdf_local_cfg *cfg, float *tsdf_dev,
                                    float *weights_dev, int64_t *fused_host, void *stream) {
  B200_NVTX("b200tsdf_local_build");
  if (!X_dev || !C_dev || !pose_dev || !box6_host || !cfg || !tsdf_dev || !weights_dev ||
      n_pts < 0) {
    b200_set_error("b200tsdf_local_build: invalid argument");
    return B200TSDF_E_INVALID;
  }
  int32_t dims[3];
  int rc = b200tsdf_local_dims(box6_host, cfg, dims);
  if (rc) return rc;
  if (dims[0] <= 0 || dims[1] <= 0 || dims[2] <= 0) {
    b200_set_error("b200tsdf_local_build: empty ROI grid %dx%dx%d", dims[0], dims[1], dims[2]);
    return B200TSDF_E_INVALID;
  }
  cudaStream_t st = (cudaStream_t)stream;
  unsigned long long *ctr;
  rc = local_counters(&ctr);
  if (rc) return rc;
  LocalParams lp;
  fill_params(lp, box6_host, cfg, dims);
  const long long nvox = (long long)dims[0] * dims[1] * dims[2];
  B200_CUDA(cudaMemsetAsync(tsdf_dev, 0, nvox * sizeof(float), st));
  B200_CUDA(cudaMemsetAsync(weights_dev, 0, nvox * sizeof(float), st));
  B200_CUDA(cudaMemsetAsync(ctr, 0, 2 * sizeof(unsigned long long), st));
  if (n_pts > 0) {
    local_accumulate_kernel<<<(unsigned)((n_pts + 255) / 256), 256, 0, st>>>(
        lp, X_dev, C_dev, pose_dev, (long long)n_pts, tsdf_dev, weights_dev, ctr);
    B200_CUDA(cudaGetLastError());
  }
  local_finalize_kernel<<<(unsigned)((nvox + 255) / 256), 256, 0, st>>>(nvox, tsdf_dev,
                                                                         weights_dev, ctr);
  B200_CUDA(cudaGetLastError());
  if (fused_host) {
    unsigned long long c[2];
    B200_CUDA(cudaMemcpyAsync(c, ctr, sizeof(c), cudaMemcpyDeviceToHost, st));
    B200_CUDA(cudaStreamSynchronize(st));
    *fused_host = c[0] < 5 ? -1 : (int64_t)c[1];
  }
  return B200TSDF_OK;
}

// ------------------------------------------------------------------ ray cast
__device__ __forceinline__ double trilinear_ref(const float *__restrict__ vol, int nx, int ny,
                                                int nz, float x, float y, float z) {
  float hx = (float)(nx - 1), hy = (float)(ny - 1), hz = (float)(nz - 1);
  x = x < 0.0f ? 0.0f : (x > hx ? hx : x);
  y = y < 0.0f ? 0.0f : (y > hy ? hy : y);
  z = z < 0.0f ? 0.0f : (z > hz ? hz : z);
  int x0 = (int)floorf(x), y0 = (int)floorf(y), z0 = (int)floorf(z);
  int x1 = min(x0 + 1, nx - 1), y1 = min(y0 + 1, ny - 1), z1 = min(z0 + 1, nz - 1);
  double xd = (double)__fsub_rn(x, (float)x0), yd = (double)__fsub_rn(y, (float)y0),
         zd = (double)__fsub_rn(z, (float)z0);
#define VV(zz, yy, xx) ((double)vol[((long long)(zz) * ny + (yy)) * nx + (xx)])
  double c00 = __dadd_rn(__dmul_rn(VV(z0, y0, x0), 1 - xd), __dmul_rn(VV(z0, y0, x1), xd));
  double c01 = __dadd_rn(__dmul_rn(VV(z0, y1, x0), 1 - xd), __dmul_rn(VV(z0, y1, x1), xd));
  double c10 = __dadd_rn(__dmul_rn(VV(z1, y0, x0), 1 - xd), __dmul_rn(VV(z1, y0, x1), xd));
  double c11 = __dadd_rn(__dmul_rn(VV(z1, y1, x0), 1 - xd), __dmul_rn(VV(z1, y1, x1), xd));
#undef VV
  double c0 = __dadd_rn(__dmul_rn(c00, 1 - yd), __dmul_rn(c01, yd));
  double c1 = __dadd_rn(__dmul_rn(c10, 1 - yd), __dmul_rn(c11, yd));
  return __dadd_rn(__dmul_rn(c0, 1 - zd), __dmul_rn(c1, zd));
}

__global__ void local_raycast_kernel(LocalParams lp, const float *__restrict__ tsdf,
                                     const float *__restrict__ X, const long long *pix,
                                     long long m, int n_samples, double max_disp,
                                     float *X_refined, uint8_t *hits) {
  long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= m) return;
  uint8_t hit = 0;
  const float P[3] = {X[3 * pix[q]], X[3 * pix[q] + 1], X[3 * pix[q] + 2]};
  const double depth = (double)P[2]; // :970
  if (!(depth < 0.05)) {
    double t_start = depth - 0.1; // :976-978
    if (t_start < 0.05) t_start = 0.05;
    const double t_end = depth + 0.1;
    bool have_prev = false;
    double prev_sdf = 0.0;
    float prev_t = 0.0f;
    for (int j = 0; j < n_samples; ++j) {
      const float t = linspace_torch((float)t_start, (float)t_end, n_samples, j);
      const float ratio = __fdiv_rn(t, (float)depth); // :986
      float s[3];
      bool out = false;
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        s[a] = __fmul_rn(P[a], ratio);
        out = out || s[a] < lp.lo[a] || s[a] > lp.hi[a];
      }
      if (out) { // :989-991
        have_prev = false;
        continue;
      }
      float gp[3];
#pragma unroll
      for (int a = 0; a < 3; ++a) gp[a] = __fdiv_rn(__fsub_rn(s[a], lp.lo[a]), lp.av[a]);
      const double sdf = trilinear_ref(tsdf, lp.nx, lp.ny, lp.nz, gp[0], gp[1], gp[2]);
      if (have_prev && prev_sdf * sdf < 0) { // :998-1000
        const double alpha = fabs(prev_sdf) / (fabs(prev_sdf) + fabs(sdf) + 1.0e-8);
        const float ts = __fadd_rn(prev_t, __fmul_rn((float)alpha, __fsub_rn(t, prev_t)));
        const float rs = __fdiv_rn(ts, (float)depth);
        const float sp[3] = {__fmul_rn(P[0], rs), __fmul_rn(P[1], rs), __fmul_rn(P[2], rs)};
        const double disp = (double)norm3_torch(__fsub_rn(sp[0], P[0]), __fsub_rn(sp[1], P[1]),
                                                __fsub_rn(sp[2], P[2]));
        if (disp <= max_disp) { // :1008-1011
          X_refined[3 * pix[q]] = sp[0];
          X_refined[3 * pix[q] + 1] = sp[1];
          X_refined[3 * pix[q] + 2] = sp[2];
          hit = 1;
        }
        break; // :1013
      }
      prev_sdf = sdf;
      prev_t = t;
      have_prev = true;
    }
  }
  hits[q] = hit;
}

extern "C" int b200tsdf_local_raycast(const float *tsdf_dev, const int32_t dims3_host[3],
                                      const float *box6_host, const float *X_dev,
                                      const int64_t *pix_dev, int64_t m,
                                      const b200tsdf_local_cfg *cfg, float *X_refined_dev,
                                      uint8_t *hits_dev, void *stream) {
  B200_NVTX("b200tsdf_local_raycast");
  if (!tsdf_dev || !dims3_host || !box6_host || !X_dev || !cfg || m < 0 ||
      (m > 0 && (!pix_dev || !X_refined_dev || !hits_dev))) {
    b200_set_error("b200tsdf_local_raycast: invalid argument");
    return B200TSDF_E_INVALID;
  }
  if (m == 0) return B200TSDF_OK;
  LocalParams lp;
  fill_params(lp, box6_host, cfg, dims3_host);
  cudaStream_t st = (cudaStream_t)stream;
  local_raycast_kernel<<<(unsigned)((m + 63) / 64), 64, 0, st>>>(
      lp, tsdf_dev, X_dev, (const long long *)pix_dev, (long long)m, cfg->ray_samples,
      cfg->max_displacement, X_refined_dev, hits_dev);
  B200_CUDA(cudaGetLastError());
  return B200TSDF_OK;
}
