// quality.cu -- per-patch quality maps on the device.
//
// Restates mast3r_slam/quality_core.py: the per-patch reductions of
// compute_batch (:127-140) -- u_from_CQ (:43-49), r_from_scalar (:51-52),
// valid_grid (:54-56), all through reduce_grid (:15-27) -- and classify
// (:63-109) with robust_z (:58-61).  The reference builds these grids with
// torch ops and ships them through numpy and a queue; here the grids stay in
// device memory and 1 - priority feeds b200tsdf_integrate_keyframes directly.
//
// Medians are torch's: nanmedian(dim) = LOWER median of the non-NaN entries
// (NaN if none), median() = lower median, NaN if any entry is NaN.  A warp owns
// a patch; the element of rank k is found by a bitwise select on the
// order-preserving integer image of the floats (largest key K with
// #{keys < K} <= k, built from the top bit down: 32 counting passes over the
// patch in shared memory, one warp reduction each) -- O(32 ps^2 / 32) compares
// per warp instead of the O(ps^4 / 32) of rank counting.
#include <math.h>
#include <stdint.h>

#include "common.cuh"

#define Q_WARPS 4
#define Q_MAX_PATCH 1024 // ps <= 32

// order-preserving map float -> uint32 (-x < +x, -0 < +0); NaN -> 0xFFFFFFFF (above everything)
__device__ __forceinline__ uint32_t ord_key(float x) {
  if (x != x) return 0xFFFFFFFFu;
  const uint32_t b = __float_as_uint(x);
  return b ^ ((b >> 31) ? 0xFFFFFFFFu : 0x80000000u);
}
__device__ __forceinline__ float ord_unkey(uint32_t k) {
  return __uint_as_float(k ^ ((k >> 31) ? 0x80000000u : 0xFFFFFFFFu));
}

// element of rank `target` (0-based, target < number of non-NaN keys) of the keys k[0..n)
__device__ __forceinline__ float warp_select(const uint32_t *k, int n, int target, int lane) {
  uint32_t prefix = 0;
#pragma unroll 1
  for (int bit = 31; bit >= 0; --bit) {
    const uint32_t cand = prefix | (1u << bit);
    int c = 0;
    for (int i = lane; i < n; i += 32) c += k[i] < cand;
    if (__reduce_add_sync(0xFFFFFFFFu, c) <= target) prefix = cand;
  }
  return ord_unkey(prefix);
}

__global__ void __launch_bounds__(32 * Q_WARPS)
    quality_patch_kernel(const float *__restrict__ C, const float *__restrict__ Q,
                         const float *__restrict__ r_pix, const uint8_t *__restrict__ valid, int H,
                         int W, int ps, float c_den, float q_den, float *__restrict__ u_grid,
                         float *__restrict__ r_grid, float *__restrict__ v_grid) {
  __shared__ uint32_t s_u[Q_WARPS][Q_MAX_PATCH], s_r[Q_WARPS][Q_MAX_PATCH];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int gh = H / ps, gw = W / ps, n = ps * ps;
  const int patch = blockIdx.x * Q_WARPS + warp;
  if (patch >= gh * gw) return;
  { // blockIdx.y = keyframe of the batch: maps are B x H*W, grids B x gh*gw
    const long long o_pix = (long long)blockIdx.y * H * W, o_grid = (long long)blockIdx.y * gh * gw;
    C += o_pix, Q += o_pix, r_pix += o_pix, valid += o_pix;
    u_grid += o_grid, r_grid += o_grid, v_grid += o_grid;
  }
  const int gy = patch / gw, gx = patch - gy * gw;
  int k_u = 0, k_r = 0, n_valid = 0;
  for (int i = lane; i < n; i += 32) {
    const int py = gy * ps + i / ps, px = gx * ps + i % ps;
    const long long pix = (long long)py * W + px;
    // :45-47  U = 1 - sqrt(clamp(clamp(C / (C_thr + 1e-8)) * clamp(Q / (Q_thr + 1e-8))))
    const float cn = fminf(fmaxf(__fdiv_rn(C[pix], c_den), 0.f), 1.f);
    const float qn = fminf(fmaxf(__fdiv_rn(Q[pix], q_den), 0.f), 1.f);
    float u = __fsub_rn(1.0f, __fsqrt_rn(fminf(fmaxf(__fmul_rn(cn, qn), 0.f), 1.f)));
    if (C[pix] != C[pix] || Q[pix] != Q[pix]) u = __int_as_float(0x7FC00000); // clamp keeps NaN in torch
    const bool ok = valid[pix] != 0;
    const float r = ok ? r_pix[pix] : __int_as_float(0x7FC00000); // :24 masked_fill(M < 0.5, nan)
    s_u[warp][i] = ord_key(u);
    s_r[warp][i] = ord_key(r);
    k_u += u == u;
    k_r += r == r;
    n_valid += ok;
  }
  k_u = __reduce_add_sync(0xFFFFFFFFu, k_u);
  k_r = __reduce_add_sync(0xFFFFFFFFu, k_r);
  n_valid = __reduce_add_sync(0xFFFFFFFFu, n_valid);
  __syncwarp();
  const float mu = k_u ? warp_select(s_u[warp], n, (k_u - 1) >> 1, lane) : __int_as_float(0x7FC00000);
  const float mr = k_r ? warp_select(s_r[warp], n, (k_r - 1) >> 1, lane) : 0.0f; // :26 nan_to_num
  if (lane == 0) {
    u_grid[patch] = mu;
    r_grid[patch] = mr;
    v_grid[patch] = n_valid > 0 ? 1.0f : 0.0f; // :55-56 mean > 0
  }
}

// ---- classify: one CTA; n = gh * gw patches (<= Q_MAX_N)
#define Q_MAX_N 3072 // 1024 x 768 at patch_size 16; 36 KB of shared memory
#define Q_CLS_THREADS 512

// lower median of s[0..n) (no NaN inside); every thread returns it.  Same
// bitwise select, the count of every pass reduced over the CTA.
__device__ float cta_median(const float *s, int n, float *s_res) {
  __shared__ int s_cnt[2][Q_CLS_THREADS / 32];
  const int target = (n - 1) >> 1, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  (void)s_res;
  uint32_t prefix = 0;
#pragma unroll 1
  for (int bit = 31; bit >= 0; --bit) {
    const uint32_t cand = prefix | (1u << bit);
    int c = 0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) c += ord_key(s[i]) < cand;
    c = __reduce_add_sync(0xFFFFFFFFu, c);
    if (lane == 0) s_cnt[bit & 1][warp] = c; // double-buffered: one barrier per pass
    __syncthreads();
    int total = 0;
#pragma unroll
    for (int w = 0; w < Q_CLS_THREADS / 32; ++w) total += s_cnt[bit & 1][w];
    if (total <= target) prefix = cand;
  }
  __syncthreads();
  return ord_unkey(prefix);
}

// :58-61 robust_z into z[] (shared); NaN input anywhere -> all NaN (torch.median propagates)
__device__ void cta_robust_z(const float *x, int n, float *z, float *tmp, float *s_res,
                             int *s_flag) {
  if (threadIdx.x == 0) *s_flag = 0;
  __syncthreads();
  for (int i = threadIdx.x; i < n; i += blockDim.x)
    if (x[i] != x[i]) *s_flag = 1;
  __syncthreads();
  const bool has_nan = *s_flag != 0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) tmp[i] = x[i];
  __syncthreads();
  const float qnan = __int_as_float(0x7FC00000);
  const float m = has_nan ? qnan : cta_median(tmp, n, s_res);
  for (int i = threadIdx.x; i < n; i += blockDim.x) z[i] = fabsf(__fsub_rn(x[i], m));
  __syncthreads();
  const float mad = has_nan ? qnan : __fadd_rn(cta_median(z, n, s_res), 1.0e-6f);
  __syncthreads();
  for (int i = threadIdx.x; i < n; i += blockDim.x) z[i] = __fdiv_rn(__fsub_rn(x[i], m), mad);
  __syncthreads();
}

// classify (:63-109) of one keyframe's n patches by one CTA
__device__ __forceinline__ void classify_cta(const float *dc, const float *__restrict__ r,
                                             const float *__restrict__ u, int n, float thr_zr,
                                             float thr_zu, float thr_dc,
                                             long long *__restrict__ cls_out,
                                             float *__restrict__ pri_out,
                                             float *__restrict__ q_out) {
  __shared__ float s_zr[Q_MAX_N], s_zu[Q_MAX_N], s_tmp[Q_MAX_N];
  __shared__ float s_res, s_max[Q_CLS_THREADS / 32];
  __shared__ int s_flag;
  cta_robust_z(r, n, s_zr, s_tmp, &s_res, &s_flag);
  cta_robust_z(u, n, s_zu, s_tmp, &s_res, &s_flag);
  float pmax = 0.f; // p >= 0 everywhere (class 0 -> 0)
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const float zr = s_zr[i], zu = s_zu[i], d = dc[i];
    int c = 0; // :78-86, later assignments win
    if (d < thr_dc && zu > thr_zu) c = 1;
    if (d >= thr_dc && zr > thr_zr && zu > thr_zu) c = 2;
    if (zr > thr_zr && zu <= thr_zu) c = 3;
    float p = 0.f; // :96-101
    if (c == 1) p = __fadd_rn(__fsub_rn(1.0f, fminf(fmaxf(d, 0.f), 1.f)), fmaxf(zu, 0.f));
    if (c == 2) p = __fadd_rn(fmaxf(zr, 0.f), fmaxf(zu, 0.f));
    if (c == 3) p = __fadd_rn(fmaxf(zr, 0.f), fmaxf(__fsub_rn(1.0f, u[i]), 0.f));
    cls_out[i] = c;
    s_tmp[i] = p;
    pmax = fmaxf(pmax, p);
  }
  pmax = fmaxf(pmax, __shfl_xor_sync(0xFFFFFFFFu, pmax, 16));
  pmax = fmaxf(pmax, __shfl_xor_sync(0xFFFFFFFFu, pmax, 8));
  pmax = fmaxf(pmax, __shfl_xor_sync(0xFFFFFFFFu, pmax, 4));
  pmax = fmaxf(pmax, __shfl_xor_sync(0xFFFFFFFFu, pmax, 2));
  pmax = fmaxf(pmax, __shfl_xor_sync(0xFFFFFFFFu, pmax, 1));
  if ((threadIdx.x & 31) == 0) s_max[threadIdx.x >> 5] = pmax;
  __syncthreads();
  pmax = 0.f;
  for (int w = 0; w < Q_CLS_THREADS / 32; ++w) pmax = fmaxf(pmax, s_max[w]);
  const float den = __fadd_rn(pmax, 1.0e-6f); // :104
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const float p = __fdiv_rn(s_tmp[i], den);
    pri_out[i] = p;
    if (q_out) q_out[i] = __fsub_rn(1.0f, fminf(fmaxf(p, 0.f), 1.f)); // A10: q = 1 - priority
  }
}

__global__ void __launch_bounds__(Q_CLS_THREADS)
    quality_classify_kernel(const float *__restrict__ dc, const float *__restrict__ r,
                            const float *__restrict__ u, int n, float thr_zr, float thr_zu,
                            float thr_dc, long long *__restrict__ cls_out,
                            float *__restrict__ pri_out, float *__restrict__ q_out) {
  classify_cta(dc, r, u, n, thr_zr, thr_zu, thr_dc, cls_out, pri_out, q_out);
}

// compute_batch (:127-140) after the per-patch medians, one CTA per keyframe:
// inc = valid_grid * view_weight (:131), cov_ewma / delta_cov = ema_delta
// (:39-41, torch's separately rounded float32 mul / add / sub), then classify.
__global__ void __launch_bounds__(Q_CLS_THREADS)
    quality_finish_batch_kernel(const float *__restrict__ v_grid, const float *__restrict__ view_w,
                                const float *__restrict__ prev, const float *__restrict__ r,
                                const float *__restrict__ u, int n, float alpha, float one_m_alpha,
                                float thr_zr, float thr_zu, float thr_dc,
                                float *__restrict__ ewma_out, float *dc_out,
                                long long *__restrict__ cls_out, float *__restrict__ pri_out,
                                float *__restrict__ q_out) {
  const long long o = (long long)blockIdx.x * n;
  const float vw = view_w[blockIdx.x];
  for (int i = threadIdx.x; i < n; i += blockDim.x) { // classify_cta reads dc[i] on the same thread
    const float pv = prev ? prev[o + i] : 0.0f;
    const float inc = __fmul_rn(v_grid[o + i], vw);
    const float nw = __fadd_rn(__fmul_rn(alpha, pv), __fmul_rn(one_m_alpha, inc));
    ewma_out[o + i] = nw;
    dc_out[o + i] = __fsub_rn(nw, pv);
  }
  __syncthreads();
  classify_cta(dc_out + o, r + o, u + o, n, thr_zr, thr_zu, thr_dc, cls_out + o, pri_out + o,
               q_out ? q_out + o : nullptr);
}

extern "C" int b200tsdf_quality_patch_grids(const float *C_dev, const float *Q_dev,
                                            const float *r_pix_dev, const uint8_t *valid_dev,
                                            int32_t H, int32_t W, int32_t patch_size, double C_thr,
                                            double Q_thr, float *u_grid_dev, float *r_grid_dev,
                                            float *valid_grid_dev, int32_t device, void *stream) {
  B200_NVTX("b200tsdf_quality_patch_grids");
  if (!C_dev || !Q_dev || !r_pix_dev || !valid_dev || !u_grid_dev || !r_grid_dev ||
      !valid_grid_dev || H <= 0 || W <= 0 || patch_size <= 0 ||
      patch_size * patch_size > Q_MAX_PATCH) {
    b200_set_error("b200tsdf_quality_patch_grids: invalid argument (patch_size <= 32)");
    return B200TSDF_E_INVALID;
  }
  const int n = (H / patch_size) * (W / patch_size);
  if (n == 0) return B200TSDF_OK;
  B200_CUDA(cudaSetDevice(device));
  quality_patch_kernel<<<(n + Q_WARPS - 1) / Q_WARPS, 32 * Q_WARPS, 0, (cudaStream_t)stream>>>(
      C_dev, Q_dev, r_pix_dev, valid_dev, H, W, patch_size, (float)(C_thr + 1e-8),
      (float)(Q_thr + 1e-8), u_grid_dev, r_grid_dev, valid_grid_dev);
  B200_CUDA(cudaGetLastError());
  return B200TSDF_OK;
}

extern "C" int b200tsdf_quality_classify(const float *delta_cov_dev, const float *r_dev,
                                         const float *u_dev, int32_t n, double thr_zr,
                                         double thr_zu, double thr_dc, int64_t *class_dev,
                                         float *priority_dev, float *q_grid_dev, int32_t device,
                                         void *stream) {
  B200_NVTX("b200tsdf_quality_classify");
  if (!delta_cov_dev || !r_dev || !u_dev || !class_dev || !priority_dev || n <= 0 ||
      n > Q_MAX_N) {
    b200_set_error("b200tsdf_quality_classify: invalid argument (1 <= n <= %d)", Q_MAX_N);
    return B200TSDF_E_INVALID;
  }
  B200_CUDA(cudaSetDevice(device));
  quality_classify_kernel<<<1, Q_CLS_THREADS, 0, (cudaStream_t)stream>>>(
      delta_cov_dev, r_dev, u_dev, n, (float)thr_zr, (float)thr_zu, (float)thr_dc,
      (long long *)class_dev, priority_dev, q_grid_dev);
  B200_CUDA(cudaGetLastError());
  return B200TSDF_OK;
}

extern "C" int b200tsdf_quality_batch(const float *C_dev, const float *Q_dev,
                                      const float *r_pix_dev, const uint8_t *valid_dev,
                                      const float *view_w_dev, const float *prev_ewma_dev,
                                      int32_t B, int32_t H, int32_t W, int32_t patch_size,
                                      double C_thr, double Q_thr, double alpha, double thr_zr,
                                      double thr_zu, double thr_dc, float *u_grid_dev,
                                      float *r_grid_dev, float *valid_grid_dev, float *ewma_dev,
                                      float *delta_cov_dev, int64_t *class_dev,
                                      float *priority_dev, float *q_grid_dev, int32_t device,
                                      void *stream) {
  B200_NVTX("b200tsdf_quality_batch");
  if (!C_dev || !Q_dev || !r_pix_dev || !valid_dev || !view_w_dev || !u_grid_dev || !r_grid_dev ||
      !valid_grid_dev || !ewma_dev || !delta_cov_dev || !class_dev || !priority_dev || B < 0 ||
      B > 65535 || H <= 0 || W <= 0 || patch_size <= 0 || patch_size * patch_size > Q_MAX_PATCH) {
    b200_set_error("b200tsdf_quality_batch: invalid argument (patch_size <= 32, B <= 65535)");
    return B200TSDF_E_INVALID;
  }
  const int n = (H / patch_size) * (W / patch_size);
  if (n > Q_MAX_N) {
    b200_set_error("b200tsdf_quality_batch: %d patches per keyframe (max %d)", n, Q_MAX_N);
    return B200TSDF_E_INVALID;
  }
  if (n == 0 || B == 0) return B200TSDF_OK;
  B200_CUDA(cudaSetDevice(device));
  cudaStream_t st = (cudaStream_t)stream;
  quality_patch_kernel<<<dim3((n + Q_WARPS - 1) / Q_WARPS, B), 32 * Q_WARPS, 0, st>>>(
      C_dev, Q_dev, r_pix_dev, valid_dev, H, W, patch_size, (float)(C_thr + 1e-8),
      (float)(Q_thr + 1e-8), u_grid_dev, r_grid_dev, valid_grid_dev);
  quality_finish_batch_kernel<<<B, Q_CLS_THREADS, 0, st>>>(
      valid_grid_dev, view_w_dev, prev_ewma_dev, r_grid_dev, u_grid_dev, n, (float)alpha,
      (float)(1.0 - alpha), (float)thr_zr, (float)thr_zu, (float)thr_dc, ewma_dev, delta_cov_dev,
      (long long *)class_dev, priority_dev, q_grid_dev);
  B200_CUDA(cudaGetLastError());
  return B200TSDF_OK;
}
