// sweep.cu -- streaming passes over the allocated voxel blocks: stats, voxel /
// block export, batched query.
// Replaces TSDFVolume.stats (mast3r_slam/tsdf/global_volume.py:136-140),
// TSDFVolume.query + _estimate_gradient (:93-128); export has no reference
// counterpart (it dumps what TSDFVolume._voxels holds, for parity tests).
#include <string.h>

#include "common.cuh"
#include "sim3.cuh"
#include "ray.cuh"

// One CTA per allocated block, 128 threads x 4 voxels, 128-bit loads.
__global__ void __launch_bounds__(128)
    stats_kernel(VolumeDev v, int which, unsigned long long *out /* [valid,total] */) {
  const uint32_t nb = (uint32_t)v.ctr[CTR_BLOCKS];
  uint32_t valid = 0, total = 0;
  uint32_t b = blockIdx.x;
  uint32_t slot = b < nb ? v.list[b] : 0u;
  while (b < nb) {
    // the next block's slot is fetched while this block streams (no dependent-load bubble)
    const uint32_t bn = b + gridDim.x;
    const uint32_t slot_n = bn < nb ? v.list[bn] : 0u;
    const size_t base = (size_t)slot * B200_BLOCK_VOX + threadIdx.x;
    float4 a[4];
    unsigned long long c[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) { // 8 independent 128/64-bit loads in flight per thread
      a[k] = v.vox[base + k * 128];
      c[k] = v.cnt[base + k * 128];
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float t, w;
      uint32_t hits;
      if (voxel_view(a[k], c[k], which, v.max_weight, t, w, hits)) {
        total++;
        if (w >= v.min_weight) valid++;
      }
    }
    b = bn, slot = slot_n;
  }
  valid = __reduce_add_sync(0xFFFFFFFFu, valid);
  total = __reduce_add_sync(0xFFFFFFFFu, total);
  if ((threadIdx.x & 31) == 0) {
    if (valid) atomicAdd(out, (unsigned long long)valid);
    if (total) atomicAdd(out + 1, (unsigned long long)total);
  }
}

extern "C" int b200tsdf_stats(b200tsdf_t *h, int which, int64_t *valid_host, int64_t *total_host,
                              void *stream) {
  B200_NVTX("b200tsdf_stats");
  if (!h || !valid_host || !total_host || (which != 0 && which != 1)) return B200TSDF_E_INVALID;
  cudaStream_t st = (cudaStream_t)stream;
  B200_CUDA(cudaSetDevice(h->cfg.device));
  B200_CUDA(cudaMemsetAsync(h->scratch, 0, 16, st));
  stats_kernel<<<(unsigned)(h->sm_count > 0 ? h->sm_count : 148) * 8, 128, 0, st>>>(h->d, which, h->scratch);
  B200_CUDA(cudaGetLastError());
  unsigned long long r[2];
  B200_CUDA(cudaMemcpyAsync(r, h->scratch, 16, cudaMemcpyDeviceToHost, st));
  B200_CUDA(cudaStreamSynchronize(st));
  *valid_host = (int64_t)r[0];
  *total_host = (int64_t)r[1];
  return B200TSDF_OK;
}

// Export: warp-scan compaction (ballot + popc), one atomicAdd per warp.
__global__ void __launch_bounds__(128)
    export_kernel(VolumeDev v, int which, long long *keys, float *tsdf, float *weight,
                  uint32_t *count, long long cap, unsigned long long *n_out) {
  const uint32_t nb = (uint32_t)v.ctr[CTR_BLOCKS];
  const int lane = threadIdx.x & 31;
  uint32_t b = blockIdx.x;
  uint32_t slot = b < nb ? v.list[b] : 0u;
  while (b < nb) {
    const uint32_t bn = b + gridDim.x;
    const uint32_t slot_n = bn < nb ? v.list[bn] : 0u; // fetched while this block streams
    const unsigned long long key = v.keys[slot];
    const size_t base = (size_t)slot * B200_BLOCK_VOX + threadIdx.x;
    float4 a[4];
    unsigned long long c[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) { // all loads of the block first
      a[k] = v.vox[base + k * 128];
      c[k] = v.cnt[base + k * 128];
    }
    int bx, by, bz;
    unpack_block_key(key, bx, by, bz);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int i = threadIdx.x + k * 128;
      float t, w;
      uint32_t hits;
      bool on = voxel_view(a[k], c[k], which, v.max_weight, t, w, hits);
      unsigned mask = __ballot_sync(0xFFFFFFFFu, on);
      if (mask == 0) continue;
      unsigned long long start = 0;
      if (lane == 0) start = atomicAdd(n_out, (unsigned long long)__popc(mask));
      start = __shfl_sync(0xFFFFFFFFu, start, 0);
      if (on && keys) {
        long long o = (long long)start + __popc(mask & ((1u << lane) - 1));
        if (o < cap) {
          keys[3 * o + 0] = (long long)bx * 8 + (i & 7);
          keys[3 * o + 1] = (long long)by * 8 + ((i >> 3) & 7);
          keys[3 * o + 2] = (long long)bz * 8 + (i >> 6);
          tsdf[o] = t;
          weight[o] = w;
          count[o] = hits;
        }
      }
    }
    b = bn, slot = slot_n;
  }
}

extern "C" int b200tsdf_export_voxels(b200tsdf_t *h, int which, int64_t *keys_dev, float *tsdf_dev,
                                      float *weight_dev, uint32_t *count_dev, int64_t cap,
                                      int64_t *n_host, void *stream) {
  B200_NVTX("b200tsdf_export_voxels");
  if (!h || !n_host || (which != 0 && which != 1)) return B200TSDF_E_INVALID;
  if (cap > 0 && (!keys_dev || !tsdf_dev || !weight_dev || !count_dev)) return B200TSDF_E_INVALID;
  cudaStream_t st = (cudaStream_t)stream;
  B200_CUDA(cudaSetDevice(h->cfg.device));
  B200_CUDA(cudaMemsetAsync(h->scratch, 0, 8, st));
  export_kernel<<<(unsigned)(h->sm_count > 0 ? h->sm_count : 148) * 8, 128, 0, st>>>(h->d, which, cap > 0 ? (long long *)keys_dev : nullptr,
                                         tsdf_dev, weight_dev, count_dev, (long long)cap,
                                         h->scratch);
  B200_CUDA(cudaGetLastError());
  unsigned long long n;
  B200_CUDA(cudaMemcpyAsync(&n, h->scratch, 8, cudaMemcpyDeviceToHost, st));
  B200_CUDA(cudaStreamSynchronize(st));
  *n_host = (int64_t)n;
  return B200TSDF_OK;
}

__global__ void export_blocks_kernel(VolumeDev v, long long *bkeys, long long cap) {
  uint32_t nb = (uint32_t)v.ctr[CTR_BLOCKS];
  for (long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x; b < nb && b < cap;
       b += (long long)gridDim.x * blockDim.x) {
    int bx, by, bz;
    unpack_block_key(v.keys[v.list[b]], bx, by, bz);
    bkeys[3 * b] = bx, bkeys[3 * b + 1] = by, bkeys[3 * b + 2] = bz;
  }
}

extern "C" int b200tsdf_export_blocks(b200tsdf_t *h, int64_t *bkeys_dev, int64_t cap,
                                      int64_t *n_host, void *stream) {
  B200_NVTX("b200tsdf_export_blocks");
  if (!h || !n_host || (cap > 0 && !bkeys_dev)) return B200TSDF_E_INVALID;
  cudaStream_t st = (cudaStream_t)stream;
  B200_CUDA(cudaSetDevice(h->cfg.device));
  unsigned long long n;
  B200_CUDA(cudaMemcpyAsync(&n, h->d.ctr + CTR_BLOCKS, 8, cudaMemcpyDeviceToHost, st));
  B200_CUDA(cudaStreamSynchronize(st));
  *n_host = (int64_t)n;
  if (cap > 0 && n > 0) {
    export_blocks_kernel<<<296, 256, 0, st>>>(h->d, (long long *)bkeys_dev, (long long)cap);
    B200_CUDA(cudaGetLastError());
    B200_CUDA(cudaStreamSynchronize(st));
  }
  return B200TSDF_OK;
}

// ------------------------------------------------------------------ query
__device__ __forceinline__ bool fetch(const VolumeDev &v, int which, int kx, int ky, int kz,
                                      float &t, float &w) {
  int bx = kx >> 3, by = ky >> 3, bz = kz >> 3;
  if (!block_in_range(bx, by, bz)) return false;
  uint32_t s = hash_find(v, pack_block_key(bx, by, bz));
  if (s == 0xFFFFFFFFu) return false;
  size_t off = (size_t)s * B200_BLOCK_VOX + (((kz & 7) << 6) | ((ky & 7) << 3) | (kx & 7));
  uint32_t hits;
  return voxel_view(v.vox[off], v.cnt[off], which, v.max_weight, t, w, hits);
}

// The 7 voxels TSDFVolume.query reads (global_volume.py:93-128): tap 0 the
// voxel of the point, taps 1..6 its +x, -x, +y, -y, +z, -z neighbours.
// OWNED_ONLY: report only voxels of this handle's z-slabs (multi-GPU: a halo
// copy of a foreign block is not reported, so every voxel has exactly one
// reporting rank).  Neighbour taps are fetched only when some rank may need
// them: always under OWNED_ONLY (the centre may live on another rank),
// otherwise only when the centre holds a value (the reference's early return).
struct QueryTaps {
  float t[7], w[7];
  uint32_t mask; // bit i: tap i present
};

template <typename T>
__device__ __forceinline__ void query_key(const T pt[3], T vs, int k[3]) {
  using R = Real<T>;
  for (int d = 0; d < 3; ++d) { // global_volume.py:133-134
    T f = R::flr(R::div(pt[d], vs));
    f = f < (T)-2.0e9 ? (T)-2.0e9 : (f > (T)2.0e9 ? (T)2.0e9 : f);
    k[d] = (int)f;
  }
}

template <bool OWNED_ONLY>
__device__ __forceinline__ void fetch_taps(const VolumeDev &v, int which, const int k[3],
                                           bool finite, QueryTaps &q) {
  q.mask = 0;
#pragma unroll
  for (int i = 0; i < 7; ++i) q.t[i] = 0.0f, q.w[i] = 0.0f;
  if (!finite) return;
  if ((!OWNED_ONLY || slab_owns_z(v, k[2])) && fetch(v, which, k[0], k[1], k[2], q.t[0], q.w[0]))
    q.mask = 1;
  if (!OWNED_ONLY && !(q.mask && q.w[0] >= v.min_weight)) return; // :98-100
#pragma unroll
  for (int i = 1; i < 7; ++i) {
    int kk[3] = {k[0], k[1], k[2]};
    kk[(i - 1) >> 1] += (i & 1) ? 1 : -1;
    if (OWNED_ONLY && !slab_owns_z(v, kk[2])) continue;
    if (fetch(v, which, kk[0], kk[1], kk[2], q.t[i], q.w[i])) q.mask |= 1u << i;
  }
}

// Value and unit gradient from the taps (global_volume.py:98-100, :107-128).
// Returns flags: 1 value, 3 value + gradient.
__device__ __forceinline__ uint8_t finish_query(const QueryTaps &q, float min_weight,
                                                double voxel_size, double &tv, double g[3]) {
  uint8_t fl = 0;
  tv = 0.0, g[0] = g[1] = g[2] = 0.0;
  if (!((q.mask & 1u) && q.w[0] >= min_weight)) return 0; // :98-100
  fl = 1;
  tv = (double)q.t[0];
  int denom = 0;
#pragma unroll
  for (int ax = 0; ax < 3; ++ax) { // :112-123
    const int ip = 1 + 2 * ax, in = 2 + 2 * ax;
    if (!((q.mask >> ip) & 1u) || !((q.mask >> in) & 1u)) continue;
    if (q.w[ip] < min_weight || q.w[in] < min_weight) continue;
    // voxels are stored as float32 sums: differences at the rounding-noise
    // level are the reference's exact zeros (equal tsdf on both sides)
    double diff = (double)q.t[ip] - (double)q.t[in];
    if (fabs(diff) <= 4.0e-7) diff = 0.0;
    g[ax] = diff / (2.0 * voxel_size);
    denom++;
  }
  if (denom) {
    double nrm = sqrt(fma(g[2], g[2], fma(g[1], g[1], g[0] * g[0])));
    if (nrm >= 1.0e-9) { // :126-128
      g[0] /= nrm, g[1] /= nrm, g[2] /= nrm;
      fl = 3;
    } else {
      g[0] = g[1] = g[2] = 0.0;
    }
  }
  return fl;
}

// 64-byte record of the taps: {t0,w0,t1,w1}, {t2,w2,t3,w3}, {t4,w4,t5,w5},
// {t6,w6,mask,0}; the mask travels as a float (<= 127, exact) so that one
// float SUM over ranks merges records with disjoint presence bits.
__device__ __forceinline__ void store_taps(float4 *dst, const QueryTaps &q) {
  dst[0] = make_float4(q.t[0], q.w[0], q.t[1], q.w[1]);
  dst[1] = make_float4(q.t[2], q.w[2], q.t[3], q.w[3]);
  dst[2] = make_float4(q.t[4], q.w[4], q.t[5], q.w[5]);
  dst[3] = make_float4(q.t[6], q.w[6], (float)q.mask, 0.0f);
}
__device__ __forceinline__ void load_taps(const float4 *src, QueryTaps &q) {
  const float4 a = src[0], b = src[1], c = src[2], d = src[3];
  q.t[0] = a.x, q.w[0] = a.y, q.t[1] = a.z, q.w[1] = a.w;
  q.t[2] = b.x, q.w[2] = b.y, q.t[3] = b.z, q.w[3] = b.w;
  q.t[4] = c.x, q.w[4] = c.y, q.t[5] = c.z, q.w[5] = c.w;
  q.t[6] = d.x, q.w[6] = d.y;
  q.mask = (uint32_t)d.z;
}

// Nearest-voxel tsdf and central-difference gradient of one point
// (TSDFVolume.query, global_volume.py:93-130).  Returns flags: 1 value, 3 value + gradient.
template <typename T>
__device__ __forceinline__ uint8_t query_point(const VolumeDev &v, int which, const T pt[3], T vs,
                                               double voxel_size, double &tv, double g[3]) {
  int k[3];
  query_key<T>(pt, vs, k);
  QueryTaps q;
  // a non-finite coordinate floors to an integer no voxel has (the reference returns
  // (None, None), global_volume.py:95-98): all three must be finite
  const bool fin = isfinite((double)pt[0]) && isfinite((double)pt[1]) && isfinite((double)pt[2]);
  fetch_taps<false>(v, which, k, fin, q);
  return finish_query(q, v.min_weight, voxel_size, tv, g);
}

template <typename T>
__global__ void query_kernel(VolumeDev v, int which, const T *__restrict__ pts, long long m,
                             T vs, double voxel_size, double *tsdf, double *grad,
                             uint8_t *flags) {
  long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= m) return;
  const T pt[3] = {pts[3 * p], pts[3 * p + 1], pts[3 * p + 2]};
  double tv, g[3];
  const uint8_t fl = query_point<T>(v, which, pt, vs, voxel_size, tv, g);
  tsdf[p] = tv;
  grad[3 * p] = (fl & 2) ? g[0] : 0.0;
  grad[3 * p + 1] = (fl & 2) ? g[1] : 0.0;
  grad[3 * p + 2] = (fl & 2) ? g[2] : 0.0;
  flags[p] = fl;
}

extern "C" int b200tsdf_query(b200tsdf_t *h, int which, const void *pts_dev, int dtype, int64_t m,
                              double *tsdf_dev, double *grad_dev, uint8_t *flags_dev,
                              void *stream) {
  B200_NVTX("b200tsdf_query");
  if (!h || m < 0 || (which != 0 && which != 1) ||
      (m > 0 && (!pts_dev || !tsdf_dev || !grad_dev || !flags_dev)) ||
      (dtype != B200TSDF_F32 && dtype != B200TSDF_F64)) {
    b200_set_error("b200tsdf_query: invalid argument");
    return B200TSDF_E_INVALID;
  }
  if (m == 0) return B200TSDF_OK;
  cudaStream_t st = (cudaStream_t)stream;
  B200_CUDA(cudaSetDevice(h->cfg.device));
  unsigned grid = (unsigned)((m + 127) / 128);
  if (dtype == B200TSDF_F32)
    query_kernel<float><<<grid, 128, 0, st>>>(h->d, which, (const float *)pts_dev, (long long)m,
                                              (float)h->cfg.voxel_size, h->cfg.voxel_size,
                                              tsdf_dev, grad_dev, flags_dev);
  else
    query_kernel<double><<<grid, 128, 0, st>>>(h->d, which, (const double *)pts_dev, (long long)m,
                                               h->cfg.voxel_size, h->cfg.voxel_size, tsdf_dev,
                                               grad_dev, flags_dev);
  B200_CUDA(cudaGetLastError());
  return B200TSDF_OK;
}

// ---------------------------------------------- pose optimiser (A9, 8(f) rank 4)
// One launch per Gauss-Newton iteration of TSDFPoseOptimizer._optimize_single
// (tsdf_optimizer.py:72-116): Sim3 act of the sampled keyframe points (float32,
// :76), query (:94-105), Jacobian J = [g, -p x g, p . g] and the weighted 7 x 7
// normal equations H = sum w J J^T, b = sum w r J with w = max(lambda C, 1e-6)
// (:107-116), reduced in float64: warp shuffles, then one atomicAdd per warp
// and entry.  out = [H upper triangle (28) | b (7) | used (1)].
template <bool TAPS>
__device__ __forceinline__ void pose_accumulate(const VolumeDev &v, int which,
                                                const float *__restrict__ pts_cam,
                                                const double *__restrict__ conf, long long n,
                                                const float *pose, float vs, double voxel_size,
                                                double lambda, double *out,
                                                const float4 *__restrict__ taps) {
  const long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  double acc[36];
#pragma unroll
  for (int i = 0; i < 36; ++i) acc[i] = 0.0;
  if (p < n) {
    float w0, w1, w2;
    sim3_act(pose, pts_cam[3 * p], pts_cam[3 * p + 1], pts_cam[3 * p + 2], w0, w1, w2);
    const float pt[3] = {w0, w1, w2};
    double t, g[3];
    uint8_t fl;
    if (TAPS) {
      QueryTaps q;
      load_taps(taps + 4 * p, q);
      fl = finish_query(q, v.min_weight, voxel_size, t, g);
    } else {
      fl = query_point<float>(v, which, pt, vs, voxel_size, t, g);
    }
    const double x = (double)w0, y = (double)w1, z = (double)w2;
    double J[7] = {g[0], g[1], g[2], -(y * g[2] - z * g[1]), -(z * g[0] - x * g[2]),
                   -(x * g[1] - y * g[0]), x * g[0] + y * g[1] + z * g[2]};
    bool ok = fl == 3 && isfinite(t);
#pragma unroll
    for (int i = 0; i < 7; ++i) ok = ok && isfinite(J[i]); // :111-112
    if (ok) {
      const double w = fmax(lambda * conf[p], 1.0e-6);
      int k = 0;
#pragma unroll
      for (int i = 0; i < 7; ++i)
#pragma unroll
        for (int j = i; j < 7; ++j) acc[k++] = w * J[i] * J[j];
#pragma unroll
      for (int i = 0; i < 7; ++i) acc[28 + i] = w * t * J[i];
      acc[35] = 1.0;
    }
  }
#pragma unroll
  for (int i = 0; i < 36; ++i) {
    double a = acc[i];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) a += __shfl_xor_sync(0xFFFFFFFFu, a, d);
    if ((threadIdx.x & 31) == 0 && a != 0.0) atomicAdd(out + i, a);
  }
}

template <bool TAPS>
__global__ void __launch_bounds__(128)
    pose_normal_eq_kernel(VolumeDev v, int which, const float *__restrict__ pts_cam,
                          const double *__restrict__ conf, long long n, const float *pose,
                          float vs, double voxel_size, double lambda, double *out,
                          const float4 *__restrict__ taps) {
  pose_accumulate<TAPS>(v, which, pts_cam, conf, n, pose, vs, voxel_size, lambda, out, taps);
}

static int pose_normal_equations_impl(b200tsdf_t *h, int which, const float *pts_cam_dev,
                                      const double *conf_dev, int64_t n, const float *pose_host,
                                      double lambda, double *H_host, double *b_host,
                                      int64_t *used_host, void *stream, const float *taps_dev) {
  if (!h || n < 0 || (which != 0 && which != 1) || !pose_host || !H_host || !b_host ||
      !used_host || (n > 0 && (!pts_cam_dev || !conf_dev))) {
    b200_set_error("b200tsdf_pose_normal_equations: invalid argument");
    return B200TSDF_E_INVALID;
  }
  cudaStream_t st = (cudaStream_t)stream;
  B200_CUDA(cudaSetDevice(h->cfg.device));
  // scratch (64 x 8 bytes): [24..60) accumulators, [60..64) the pose (8 floats)
  double *acc = reinterpret_cast<double *>(h->scratch + 24);
  float *pose_dev = reinterpret_cast<float *>(h->scratch + 60);
  B200_CUDA(cudaMemsetAsync(acc, 0, 36 * sizeof(double), st));
  B200_CUDA(cudaMemcpyAsync(pose_dev, pose_host, 8 * sizeof(float), cudaMemcpyHostToDevice, st));
  if (n > 0) {
    if (taps_dev)
      pose_normal_eq_kernel<true><<<(unsigned)((n + 127) / 128), 128, 0, st>>>(
          h->d, which, pts_cam_dev, conf_dev, (long long)n, pose_dev, (float)h->cfg.voxel_size,
          h->cfg.voxel_size, lambda, acc, (const float4 *)taps_dev);
    else
      pose_normal_eq_kernel<false><<<(unsigned)((n + 127) / 128), 128, 0, st>>>(
          h->d, which, pts_cam_dev, conf_dev, (long long)n, pose_dev, (float)h->cfg.voxel_size,
          h->cfg.voxel_size, lambda, acc, nullptr);
    B200_CUDA(cudaGetLastError());
  }
  double r[36];
  B200_CUDA(cudaMemcpyAsync(r, acc, sizeof(r), cudaMemcpyDeviceToHost, st));
  B200_CUDA(cudaStreamSynchronize(st));
  int k = 0;
  for (int i = 0; i < 7; ++i)
    for (int j = i; j < 7; ++j) H_host[i * 7 + j] = H_host[j * 7 + i] = r[k++];
  for (int i = 0; i < 7; ++i) b_host[i] = r[28 + i];
  *used_host = (int64_t)(r[35] + 0.5);
  return B200TSDF_OK;
}

extern "C" int b200tsdf_pose_normal_equations(b200tsdf_t *h, int which, const float *pts_cam_dev,
                                              const double *conf_dev, int64_t n,
                                              const float *pose_host, double lambda,
                                              double *H_host, double *b_host, int64_t *used_host,
                                              void *stream) {
  B200_NVTX("b200tsdf_pose_normal_equations");
  return pose_normal_equations_impl(h, which, pts_cam_dev, conf_dev, n, pose_host, lambda, H_host,
                                    b_host, used_host, stream, nullptr);
}

// ---- the whole Gauss-Newton iteration on the device -------
// TSDFPoseOptimizer._optimize_single (tsdf_optimizer.py:72-92): linearise, solve
// (H + damping I) delta = -b (:82, np.linalg.solve = LU with partial pivoting),
// cast delta to the point dtype float32 (:84), pose <- exp(delta) * pose (:85-86).
// One launch per iteration and no host round trip: the last block to finish the
// reduction (ticket counter behind a __threadfence) solves the 7 x 7 system in
// float64, retracts the pose and re-arms the accumulators for the next launch.
struct PoseOptState {
  double acc[36];  // upper triangle of H (28), b (7), number of points used
  double pose[8];  // float64 master copy [t, q_xyzw, s]
  float pose_f[8]; // what the next linearisation acts with (:76 is float32)
  unsigned int done_blocks;
  int stop;       // set when used < 6 (:78-79) or the system is singular (LinAlgError)
  int iterations; // retractions applied
  int used;       // points with value and gradient in the last linearisation
};

// Sim(3) exponential of a tangent [tau, phi, sigma] -> [t, q_xyzw, s]: the closed
// form of tsdf/sim3.py:exp (the lietorch operation of tsdf_optimizer.py:85).
__device__ void sim3_exp_d(const double d[7], double out[8]) {
  const double tau[3] = {d[0], d[1], d[2]}, phi[3] = {d[3], d[4], d[5]}, sigma = d[6];
  const double theta = sqrt(phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2]);
  const double s = exp(sigma), eps = 1.0e-9;
  double A, B, C;
  if (fabs(sigma) < eps) {
    C = 1.0;
    if (theta < eps) {
      A = 0.5, B = 1.0 / 6.0;
    } else {
      A = (1.0 - cos(theta)) / (theta * theta);
      B = (theta - sin(theta)) / (theta * theta * theta);
    }
  } else {
    C = (s - 1.0) / sigma;
    if (theta < eps) {
      A = ((sigma - 1.0) * s + 1.0) / (sigma * sigma);
      B = (s * 0.5 * sigma * sigma + s - 1.0 - sigma * s) / (sigma * sigma * sigma);
    } else {
      const double a = s * sin(theta), b = s * cos(theta), c = theta * theta + sigma * sigma;
      A = (a * sigma + (1.0 - b) * theta) / (theta * c);
      B = (C - ((b - 1.0) * sigma + a * theta) / c) / (theta * theta);
    }
  }
  // W = A [phi]x + B [phi]x^2 + C I ;  t = W tau
  const double O[3][3] = {{0.0, -phi[2], phi[1]}, {phi[2], 0.0, -phi[0]}, {-phi[1], phi[0], 0.0}};
  for (int i = 0; i < 3; ++i) {
    double t = 0.0;
    for (int j = 0; j < 3; ++j) {
      double o2 = 0.0;
      for (int k = 0; k < 3; ++k) o2 += O[i][k] * O[k][j];
      t += (A * O[i][j] + B * o2 + (i == j ? C : 0.0)) * tau[j];
    }
    out[i] = t;
  }
  double q[4];
  if (theta < eps) {
    q[0] = 0.5 * phi[0], q[1] = 0.5 * phi[1], q[2] = 0.5 * phi[2], q[3] = 1.0;
  } else {
    const double k = sin(0.5 * theta) / theta;
    q[0] = k * phi[0], q[1] = k * phi[1], q[2] = k * phi[2], q[3] = cos(0.5 * theta);
  }
  const double qn = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  for (int i = 0; i < 4; ++i) out[3 + i] = q[i] / qn;
  out[7] = s;
}

// Composition a * b (x -> a(b(x))) of [t, q_xyzw, s] poses: tsdf/sim3.py:mul.
__device__ void sim3_mul_d(const double a[8], const double b[8], double out[8]) {
  const double x = a[3], y = a[4], z = a[5], w = a[6];
  const double R[3][3] = {{1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)},
                          {2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)},
                          {2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)}};
  for (int i = 0; i < 3; ++i)
    out[i] = a[7] * (R[i][0] * b[0] + R[i][1] * b[1] + R[i][2] * b[2]) + a[i];
  const double bx = b[3], by = b[4], bz = b[5], bw = b[6];
  double q[4] = {w * bx + x * bw + y * bz - z * by, w * by - x * bz + y * bw + z * bx,
                 w * bz + x * by - y * bx + z * bw, w * bw - x * bx - y * by - z * bz};
  const double qn = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  for (int i = 0; i < 4; ++i) out[3 + i] = q[i] / qn;
  out[7] = a[7] * b[7];
}

// (H + damping I) x = -b by Gaussian elimination with partial pivoting (what
// LAPACK dgesv behind np.linalg.solve does).  false: singular.
__device__ bool solve7_d(const double *acc, double damping, double x[7]) {
  double M[7][8];
  int k = 0;
  for (int i = 0; i < 7; ++i)
    for (int j = i; j < 7; ++j) M[i][j] = M[j][i] = acc[k++];
  for (int i = 0; i < 7; ++i) {
    M[i][i] += damping;
    M[i][7] = -acc[28 + i];
  }
  for (int c = 0; c < 7; ++c) {
    int piv = c;
    for (int r = c + 1; r < 7; ++r)
      if (fabs(M[r][c]) > fabs(M[piv][c])) piv = r;
    if (M[piv][c] == 0.0 || !isfinite(M[piv][c])) return false;
    if (piv != c)
      for (int j = c; j < 8; ++j) {
        const double t = M[c][j];
        M[c][j] = M[piv][j], M[piv][j] = t;
      }
    for (int r = c + 1; r < 7; ++r) {
      const double f = M[r][c] / M[c][c];
      for (int j = c; j < 8; ++j) M[r][j] -= f * M[c][j];
    }
  }
  for (int i = 6; i >= 0; --i) {
    double t = M[i][7];
    for (int j = i + 1; j < 7; ++j) t -= M[i][j] * x[j];
    x[i] = t / M[i][i];
  }
  for (int i = 0; i < 7; ++i)
    if (!isfinite(x[i])) return false;
  return true;
}

__global__ void __launch_bounds__(128)
    pose_gn_iter_kernel(VolumeDev v, int which, const float *__restrict__ pts_cam,
                        const double *__restrict__ conf, long long n, PoseOptState *s, float vs,
                        double voxel_size, double lambda, double damping) {
  __shared__ int s_last;
  const int stop = *(volatile int *)&s->stop; // a finished optimisation: later launches idle
  if (!stop)
    pose_accumulate<false>(v, which, pts_cam, conf, n, s->pose_f, vs, voxel_size, lambda, s->acc,
                           nullptr);
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = atomicAdd(&s->done_blocks, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!s_last || threadIdx.x != 0) return;
  __threadfence();
  s->done_blocks = 0;
  if (stop) return;
  double acc[36];
  for (int i = 0; i < 36; ++i) {
    acc[i] = __ldcg(s->acc + i);
    s->acc[i] = 0.0; // re-armed for the next launch
  }
  const int used = (int)(acc[35] + 0.5);
  s->used = used;
  double delta[7];
  if (used < 6 || !solve7_d(acc, damping, delta)) { // :78-79, LinAlgError
    s->stop = 1;
    return;
  }
  for (int i = 0; i < 7; ++i) delta[i] = (double)(float)delta[i]; // :84 delta in the point dtype
  double e[8], np[8];
  sim3_exp_d(delta, e);
  sim3_mul_d(e, s->pose, np);
  for (int i = 0; i < 8; ++i) {
    s->pose[i] = np[i];
    s->pose_f[i] = (float)np[i];
  }
  s->iterations += 1;
}

extern "C" int b200tsdf_pose_optimize(b200tsdf_t *h, int which, const float *pts_cam_dev,
                                      const double *conf_dev, int64_t n, double *pose_inout_host,
                                      double lambda, double damping, int32_t iterations,
                                      int64_t *used_host, int32_t *iterations_done_host,
                                      void *stream) {
  B200_NVTX("b200tsdf_pose_optimize");
  if (!h || n < 0 || (which != 0 && which != 1) || !pose_inout_host || iterations < 0 ||
      (n > 0 && (!pts_cam_dev || !conf_dev))) {
    b200_set_error("b200tsdf_pose_optimize: invalid argument");
    return B200TSDF_E_INVALID;
  }
  cudaStream_t st = (cudaStream_t)stream;
  B200_CUDA(cudaSetDevice(h->cfg.device));
  if (!h->opt_state) B200_CUDA(cudaMalloc(&h->opt_state, sizeof(PoseOptState)));
  PoseOptState init;
  memset(&init, 0, sizeof(init));
  for (int i = 0; i < 8; ++i) {
    init.pose[i] = pose_inout_host[i];
    init.pose_f[i] = (float)pose_inout_host[i];
  }
  // pageable host source: the copy is staged before the call returns to the stream
  B200_CUDA(cudaMemcpyAsync(h->opt_state, &init, sizeof(init), cudaMemcpyHostToDevice, st));
  const unsigned grid = (unsigned)((n + 127) / 128);
  for (int it = 0; it < iterations && grid > 0; ++it)
    pose_gn_iter_kernel<<<grid, 128, 0, st>>>(h->d, which, pts_cam_dev, conf_dev, (long long)n,
                                              (PoseOptState *)h->opt_state,
                                              (float)h->cfg.voxel_size, h->cfg.voxel_size, lambda,
                                              damping);
  B200_CUDA(cudaGetLastError());
  PoseOptState fin;
  B200_CUDA(cudaMemcpyAsync(&fin, h->opt_state, sizeof(fin), cudaMemcpyDeviceToHost, st));
  B200_CUDA(cudaStreamSynchronize(st)); // the ONE synchronisation of the optimisation
  for (int i = 0; i < 8; ++i) pose_inout_host[i] = fin.pose[i];
  if (used_host) *used_host = fin.used;
  if (iterations_done_host) *iterations_done_host = fin.iterations;
  return B200TSDF_OK;
}

// ------------------------------------- sharded query (multi-GPU)
// A query point's 7 voxels may live on up to two ranks (z neighbours across a
// slab seam; the centre itself on any rank).  Step 1, every rank:
// query_taps_kernel writes the taps IT OWNS (64 bytes per point, absent = 0);
// the caller sums the records over the ranks (ownership is disjoint, so the
// sum is a bit-exact merge); step 2: query_finish / the TAPS variant of the
// pose kernel evaluate the reference arithmetic on the merged taps -- the
// same finish_query the single-volume query runs.
template <typename T, bool ACT>
__global__ void query_taps_kernel(VolumeDev v, int which, const T *__restrict__ pts, long long m,
                                  T vs, const float *pose, float4 *__restrict__ taps) {
  long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= m) return;
  T pt[3] = {pts[3 * p], pts[3 * p + 1], pts[3 * p + 2]};
  if (ACT) { // camera-frame float32 points: Sim3 act first (tsdf_optimizer.py:76)
    float w0, w1, w2;
    sim3_act(pose, (float)pt[0], (float)pt[1], (float)pt[2], w0, w1, w2);
    pt[0] = (T)w0, pt[1] = (T)w1, pt[2] = (T)w2;
  }
  int k[3];
  query_key<T>(pt, vs, k);
  QueryTaps q;
  const bool fin = isfinite((double)pt[0]) && isfinite((double)pt[1]) && isfinite((double)pt[2]);
  fetch_taps<true>(v, which, k, fin, q);
  store_taps(taps + 4 * p, q);
}

__global__ void query_finish_kernel(float min_weight, double voxel_size,
                                    const float4 *__restrict__ taps, long long m, double *tsdf,
                                    double *grad, uint8_t *flags) {
  long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= m) return;
  QueryTaps q;
  load_taps(taps + 4 * p, q);
  double tv, g[3];
  const uint8_t fl = finish_query(q, min_weight, voxel_size, tv, g);
  tsdf[p] = tv;
  grad[3 * p] = g[0], grad[3 * p + 1] = g[1], grad[3 * p + 2] = g[2];
  flags[p] = fl;
}

extern "C" int b200tsdf_query_taps(b200tsdf_t *h, int which, const void *pts_dev, int dtype,
                                   int64_t m, const float *pose_host, float *taps_dev,
                                   void *stream) {
  B200_NVTX("b200tsdf_query_taps");
  if (!h || m < 0 || (which != 0 && which != 1) || (m > 0 && (!pts_dev || !taps_dev)) ||
      (dtype != B200TSDF_F32 && dtype != B200TSDF_F64) || (pose_host && dtype != B200TSDF_F32) ||
      ((uintptr_t)taps_dev & 15)) {
    b200_set_error("b200tsdf_query_taps: invalid argument");
    return B200TSDF_E_INVALID;
  }
  if (m == 0) return B200TSDF_OK;
  cudaStream_t st = (cudaStream_t)stream;
  B200_CUDA(cudaSetDevice(h->cfg.device));
  const unsigned grid = (unsigned)((m + 127) / 128);
  float4 *taps = (float4 *)taps_dev;
  if (pose_host) {
    float *pose_dev = reinterpret_cast<float *>(h->scratch + 60);
    B200_CUDA(cudaMemcpyAsync(pose_dev, pose_host, 8 * sizeof(float), cudaMemcpyHostToDevice, st));
    query_taps_kernel<float, true><<<grid, 128, 0, st>>>(h->d, which, (const float *)pts_dev,
                                                        (long long)m, (float)h->cfg.voxel_size,
                                                        pose_dev, taps);
  } else if (dtype == B200TSDF_F32) {
    query_taps_kernel<float, false><<<grid, 128, 0, st>>>(h->d, which, (const float *)pts_dev,
                                                         (long long)m, (float)h->cfg.voxel_size,
                                                         nullptr, taps);
  } else {
    query_taps_kernel<double, false><<<grid, 128, 0, st>>>(h->d, which, (const double *)pts_dev,
                                                          (long long)m, h->cfg.voxel_size, nullptr,
                                                          taps);
  }
  B200_CUDA(cudaGetLastError());
  return B200TSDF_OK;
}

extern "C" int b200tsdf_query_finish(b200tsdf_t *h, const float *taps_dev, int64_t m,
                                     double *tsdf_dev, double *grad_dev, uint8_t *flags_dev,
                                     void *stream) {
  B200_NVTX("b200tsdf_query_finish");
  if (!h || m < 0 || (m > 0 && (!taps_dev || !tsdf_dev || !grad_dev || !flags_dev)) ||
      ((uintptr_t)taps_dev & 15)) {
    b200_set_error("b200tsdf_query_finish: invalid argument");
    return B200TSDF_E_INVALID;
  }
  if (m == 0) return B200TSDF_OK;
  cudaStream_t st = (cudaStream_t)stream;
  B200_CUDA(cudaSetDevice(h->cfg.device));
  query_finish_kernel<<<(unsigned)((m + 127) / 128), 128, 0, st>>>(
      h->d.min_weight, h->cfg.voxel_size, (const float4 *)taps_dev, (long long)m, tsdf_dev,
      grad_dev, flags_dev);
  B200_CUDA(cudaGetLastError());
  return B200TSDF_OK;
}

extern "C" int b200tsdf_pose_normal_equations_taps(b200tsdf_t *h, const float *taps_dev,
                                                   const float *pts_cam_dev,
                                                   const double *conf_dev, int64_t n,
                                                   const float *pose_host, double lambda,
                                                   double *H_host, double *b_host,
                                                   int64_t *used_host, void *stream) {
  B200_NVTX("b200tsdf_pose_normal_equations_taps");
  if ((n > 0 && !taps_dev) || ((uintptr_t)taps_dev & 15)) {
    b200_set_error("b200tsdf_pose_normal_equations_taps: invalid argument");
    return B200TSDF_E_INVALID;
  }
  static const float4 none = {0, 0, 0, 0}; // n == 0: the kernel is not launched
  return pose_normal_equations_impl(h, 0, pts_cam_dev, conf_dev, n, pose_host, lambda, H_host,
                                    b_host, used_host, stream,
                                    taps_dev ? taps_dev : (const float *)&none);
}

// ------------------------------------------------------- multi-GPU seams
// Marching cubes needs, for the top voxel layer of every z-slab, the first
// voxel layer of the slab above -- owned by another rank.  The
// owner exports the voxel blocks of the FIRST block layer of each of its slabs
// (export_seam_blocks); the rank below imports them as read-only halo blocks:
// present in the hash (so mc_kernel finds them as +z neighbours) but not in
// the block list (so no sweep meshes, counts or exports them).
__device__ __forceinline__ bool is_seam_block(const VolumeDev &v, int bz) {
  return v.slab_count > 1 && (((bz * 8 - v.slab_z0) & ((1 << v.slab_shift) - 1)) == 0);
}

__global__ void __launch_bounds__(128)
    export_seam_kernel(VolumeDev v, long long *bkeys, float4 *vox, unsigned long long *cnt,
                       long long cap, unsigned long long *n_out, int all_blocks) {
  __shared__ long long s_idx;
  const uint32_t nb = (uint32_t)v.ctr[CTR_BLOCKS];
  for (uint32_t b = blockIdx.x; b < nb; b += gridDim.x) {
    const uint32_t slot = v.list[b];
    int bx, by, bz;
    unpack_block_key(v.keys[slot], bx, by, bz);
    if (!all_blocks && !is_seam_block(v, bz)) continue; // uniform per CTA
    __syncthreads();
    if (threadIdx.x == 0) s_idx = (long long)atomicAdd(n_out, 1ull);
    __syncthreads();
    const long long o = s_idx;
    if (o >= cap) continue;
    if (threadIdx.x == 0) bkeys[3 * o] = bx, bkeys[3 * o + 1] = by, bkeys[3 * o + 2] = bz;
    const size_t base = (size_t)slot * B200_BLOCK_VOX;
    for (int i = threadIdx.x; i < B200_BLOCK_VOX; i += blockDim.x) {
      vox[o * B200_BLOCK_VOX + i] = v.vox[base + i];
      cnt[o * B200_BLOCK_VOX + i] = v.cnt[base + i];
    }
  }
}

static int export_block_payload(b200tsdf_t *h, int64_t *bkeys_dev, float *vox_dev,
                                uint64_t *cnt_dev, int64_t cap, int64_t *n_host, void *stream,
                                int all_blocks) {
  if (!h || !n_host || cap < 0 || (cap > 0 && (!bkeys_dev || !vox_dev || !cnt_dev))) {
    b200_set_error("b200tsdf_export_%s: invalid argument", all_blocks ? "block_data" : "seam_blocks");
    return B200TSDF_E_INVALID;
  }
  cudaStream_t st = (cudaStream_t)stream;
  B200_CUDA(cudaSetDevice(h->cfg.device));
  B200_CUDA(cudaMemsetAsync(h->scratch, 0, 8, st));
  export_seam_kernel<<<(unsigned)(h->sm_count > 0 ? h->sm_count : 148) * 8, 128, 0, st>>>(h->d, (long long *)bkeys_dev, (float4 *)vox_dev,
                                              (unsigned long long *)cnt_dev, (long long)cap,
                                              h->scratch, all_blocks);
  B200_CUDA(cudaGetLastError());
  unsigned long long n;
  B200_CUDA(cudaMemcpyAsync(&n, h->scratch, 8, cudaMemcpyDeviceToHost, st));
  B200_CUDA(cudaStreamSynchronize(st));
  *n_host = (int64_t)n;
  return B200TSDF_OK;
}

extern "C" int b200tsdf_export_seam_blocks(b200tsdf_t *h, int64_t *bkeys_dev, float *vox_dev,
                                           uint64_t *cnt_dev, int64_t cap, int64_t *n_host,
                                           void *stream) {
  B200_NVTX("b200tsdf_export_seam_blocks");
  return export_block_payload(h, bkeys_dev, vox_dev, cnt_dev, cap, n_host, stream, 0);
}

extern "C" int b200tsdf_export_block_data(b200tsdf_t *h, int64_t *bkeys_dev, float *vox_dev,
                                          uint64_t *cnt_dev, int64_t cap, int64_t *n_host,
                                          void *stream) {
  B200_NVTX("b200tsdf_export_block_data");
  return export_block_payload(h, bkeys_dev, vox_dev, cnt_dev, cap, n_host, stream, 1);
}

__global__ void __launch_bounds__(128)
    import_halo_kernel(VolumeDev v, const long long *__restrict__ bkeys,
                       const float4 *__restrict__ vox, const unsigned long long *__restrict__ cnt,
                       long long n, unsigned long long *overflow, int owned) {
  __shared__ uint32_t s_slot;
  for (long long b = blockIdx.x; b < n; b += gridDim.x) {
    __syncthreads();
    if (threadIdx.x == 0) {
      int bx = (int)bkeys[3 * b], by = (int)bkeys[3 * b + 1], bz = (int)bkeys[3 * b + 2];
      bool fresh;
      fresh = false;
      s_slot = block_in_range(bx, by, bz) ? hash_insert(v, pack_block_key(bx, by, bz), &fresh)
                                          : B200_NO_SLOT;
      if (s_slot == B200_NO_SLOT) atomicAdd(overflow, 1ull);
      if (owned && s_slot != B200_NO_SLOT) {
        if (fresh) v.list[atomicAdd(v.ctr + CTR_BLOCKS, 1ull)] = s_slot; // an owned block
        if (v.dir_on) {
          const int dx = bx - v.dir_b0[0], dy = by - v.dir_b0[1], dz = bz - v.dir_b0[2];
          if ((unsigned)dx < (unsigned)v.dir_n[0] && (unsigned)dy < (unsigned)v.dir_n[1] &&
              (unsigned)dz < (unsigned)v.dir_n[2])
            v.dir[((size_t)dz * v.dir_n[1] + dy) * v.dir_n[0] + dx] = s_slot;
        }
      }
    }
    __syncthreads();
    if (s_slot == B200_NO_SLOT) continue;
    const size_t base = (size_t)s_slot * B200_BLOCK_VOX;
    for (int i = threadIdx.x; i < B200_BLOCK_VOX; i += blockDim.x) {
      const float4 a = vox[b * B200_BLOCK_VOX + i];
      const unsigned long long c = cnt[b * B200_BLOCK_VOX + i];
      if (owned) { // the fold state (sums, hit counters) is additive: import = merge
        red_add_v4(v.vox + base + i, a.x, a.y, a.z, a.w);
        red_add_u64(v.cnt + base + i, c);
      } else { // overwrite: a halo block is a copy
        v.vox[base + i] = a;
        v.cnt[base + i] = c;
      }
    }
  }
}

static int import_block_payload(b200tsdf_t *h, const int64_t *bkeys_dev, const float *vox_dev,
                                const uint64_t *cnt_dev, int64_t n, void *stream, int owned) {
  if (!h || n < 0 || (n > 0 && (!bkeys_dev || !vox_dev || !cnt_dev))) {
    b200_set_error("b200tsdf_import_blocks: invalid argument");
    return B200TSDF_E_INVALID;
  }
  if (n == 0) return B200TSDF_OK;
  cudaStream_t st = (cudaStream_t)stream;
  B200_CUDA(cudaSetDevice(h->cfg.device));
  B200_CUDA(cudaMemsetAsync(h->scratch + 1, 0, 8, st));
  import_halo_kernel<<<(unsigned)(n < 1184 ? n : 1184), 128, 0, st>>>(
      h->d, (const long long *)bkeys_dev, (const float4 *)vox_dev,
      (const unsigned long long *)cnt_dev, (long long)n, h->scratch + 1, owned);
  B200_CUDA(cudaGetLastError());
  unsigned long long ov;
  B200_CUDA(cudaMemcpyAsync(&ov, h->scratch + 1, 8, cudaMemcpyDeviceToHost, st));
  B200_CUDA(cudaStreamSynchronize(st));
  if (ov) {
    b200_set_error("b200tsdf_import_blocks: voxel-block hash full (%llu blocks lost)", ov);
    return B200TSDF_E_CAPACITY;
  }
  return B200TSDF_OK;
}

extern "C" int b200tsdf_import_halo_blocks(b200tsdf_t *h, const int64_t *bkeys_dev,
                                           const float *vox_dev, const uint64_t *cnt_dev,
                                           int64_t n, void *stream) {
  B200_NVTX("b200tsdf_import_halo_blocks");
  return import_block_payload(h, bkeys_dev, vox_dev, cnt_dev, n, stream, 0);
}

extern "C" int b200tsdf_import_blocks(b200tsdf_t *h, const int64_t *bkeys_dev,
                                      const float *vox_dev, const uint64_t *cnt_dev, int64_t n,
                                      void *stream) {
  B200_NVTX("b200tsdf_import_blocks");
  return import_block_payload(h, bkeys_dev, vox_dev, cnt_dev, n, stream, 1);
}
