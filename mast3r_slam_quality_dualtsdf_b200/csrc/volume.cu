// volume.cu -- handle lifetime, counters, error strings of libb200tsdf.
// Replaces TSDFVolume.__init__ (mast3r_slam/tsdf/global_volume.py:18-30).
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <new>

#include "common.cuh"

static thread_local char g_err[512] = "";

void b200_set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

extern "C" const char *b200tsdf_last_error(void) { return g_err; }
extern "C" int b200tsdf_abi_version(void) { return B200TSDF_ABI_VERSION; }

static int clamp_i32(int64_t v) {
  if (v < -2000000000ll) return -2000000000;
  if (v > 2000000000ll) return 2000000000;
  return (int)v;
}

extern "C" int b200tsdf_create(const b200tsdf_config *cfg, b200tsdf_t **out) {
  B200_NVTX("b200tsdf_create");
  if (!cfg || !out) {
    b200_set_error("b200tsdf_create: null argument");
    return B200TSDF_E_INVALID;
  }
  if (!(cfg->voxel_size > 0.0) || !(cfg->truncation > 0.0) || cfg->capacity_log2 < 4 ||
      cfg->capacity_log2 > 30) {
    b200_set_error("b200tsdf_create: bad config (voxel_size %g truncation %g capacity_log2 %d)",
                   cfg->voxel_size, cfg->truncation, cfg->capacity_log2);
    return B200TSDF_E_INVALID;
  }
  int ndev = 0;
  B200_CUDA(cudaGetDeviceCount(&ndev));
  if (ndev <= 0) {
    b200_set_error("b200tsdf_create: no CUDA device (there is no CPU fallback)");
    return B200TSDF_E_CUDA;
  }
  B200_CUDA(cudaSetDevice(cfg->device));
  b200tsdf *h = new (std::nothrow) b200tsdf();
  if (!h) return B200TSDF_E_NOMEM;
  memset(h, 0, sizeof(*h));
  h->cfg = *cfg;
  h->cap = 1ull << cfg->capacity_log2;
  VolumeDev &d = h->d;
  cudaError_t e;
  e = cudaMalloc(&d.keys, h->cap * sizeof(unsigned long long));
  if (e == cudaSuccess) e = cudaMalloc(&d.vox, h->cap * B200_BLOCK_VOX * sizeof(float4));
  if (e == cudaSuccess) e = cudaMalloc(&d.cnt, h->cap * B200_BLOCK_VOX * sizeof(unsigned long long));
  if (e == cudaSuccess) e = cudaMalloc(&d.list, h->cap * sizeof(uint32_t));
  if (e == cudaSuccess) e = cudaMalloc(&d.ctr, CTR_COPIES * CTR_COUNT * sizeof(unsigned long long));
  if (e == cudaSuccess) e = cudaMalloc(&h->scratch, 64 * sizeof(unsigned long long));
  if (e == cudaSuccess) e = cudaMalloc(&h->sched, 64);
  if (e == cudaSuccess)
    e = cudaDeviceGetAttribute(&h->sm_count, cudaDevAttrMultiProcessorCount, cfg->device);
  if (e != cudaSuccess) {
    b200_set_error("b200tsdf_create: cudaMalloc failed for 2^%d blocks (%s)", cfg->capacity_log2,
                   cudaGetErrorString(e));
    b200tsdf_destroy(h);
    return B200TSDF_E_NOMEM;
  }
  d.cap_mask = (uint32_t)(h->cap - 1);
  d.bounded = cfg->bounded;
  for (int i = 0; i < 3; ++i) {
    d.kmin[i] = clamp_i32(cfg->key_min[i]);
    d.kmax[i] = clamp_i32(cfg->key_max[i]);
  }
  if (!cfg->bounded) // slabs on an unbounded volume run the bounded code path
    for (int i = 0; i < 3; ++i) d.kmin[i] = -2000000000, d.kmax[i] = 2000000000;
  d.max_weight = (float)cfg->max_weight;
  d.min_weight = (float)cfg->min_weight;
  // z-slab ownership
  d.slab_count = cfg->slab_count > 1 ? cfg->slab_count : 1;
  d.slab_index = cfg->slab_count > 1 ? cfg->slab_index : 0;
  d.slab_mask = (d.slab_count & (d.slab_count - 1)) == 0 ? d.slab_count - 1 : -1;
  d.slab_z0 = cfg->bounded ? d.kmin[2] : 0;
  d.slab_shift = 3;
  if (d.slab_count > 1) {
    int64_t w = cfg->slab_width;
    if (w < 8 || (w & (w - 1)) != 0 || w > (1ll << 30) || cfg->slab_index < 0 ||
        cfg->slab_index >= cfg->slab_count || (d.slab_z0 & 7) != 0) {
      b200_set_error("b200tsdf_create: bad slab config (count %d index %d width %lld; width must "
                     "be a power of two >= 8 and key_min[2] a multiple of 8)",
                     cfg->slab_count, cfg->slab_index, (long long)cfg->slab_width);
      b200tsdf_destroy(h);
      return B200TSDF_E_INVALID;
    }
    d.slab_shift = 0;
    while ((1ll << d.slab_shift) < w) d.slab_shift++;
  }
  // block directory for bounded volumes (<= 256 MB of u32 entries)
  d.dir = nullptr;
  d.dir_on = 0;
  if (cfg->bounded) {
    unsigned long long n = 1;
    bool ok = true;
    for (int i = 0; i < 3; ++i) {
      if (d.kmax[i] <= d.kmin[i]) ok = false;
      d.dir_b0[i] = d.kmin[i] >> 3;
      d.dir_n[i] = ok ? ((d.kmax[i] - 1) >> 3) - d.dir_b0[i] + 1 : 0;
      if (ok && (d.dir_n[i] <= 0 || d.dir_n[i] > 4096)) ok = false;
      if (ok) n *= (unsigned long long)d.dir_n[i];
    }
    if (ok && n <= (1ull << 26)) {
      h->dir_entries = n;
      if (cudaMalloc(&d.dir, n * sizeof(uint32_t)) == cudaSuccess)
        d.dir_on = 1;
      else {
        (void)cudaGetLastError();
        d.dir = nullptr;
      }
    }
  }
  int rc = b200tsdf_reset(h, nullptr);
  if (rc != 0) {
    b200tsdf_destroy(h);
    return rc;
  }
  B200_CUDA(cudaStreamSynchronize(nullptr));
  *out = h;
  return B200TSDF_OK;
}

extern "C" int b200tsdf_destroy(b200tsdf_t *h) {
  B200_NVTX("b200tsdf_destroy");
  if (!h) return B200TSDF_OK;
  cudaSetDevice(h->cfg.device);
  cudaFree(h->d.keys);
  cudaFree(h->d.vox);
  cudaFree(h->d.cnt);
  cudaFree(h->d.list);
  cudaFree(h->d.ctr);
  cudaFree(h->d.dir);
  cudaFree(h->scratch);
  cudaFree(h->sched);
  cudaFree(h->tile_list);
  cudaFree(h->sub_thr);
  cudaFree(h->opt_state);
  cudaFree(h->nbr);
  cudaFree(h->stage_dev);
  if (h->copy_stream) {
    for (int i = 0; i <= B200_H2D_CHUNKS; ++i) cudaEventDestroy(h->copy_ev[i]);
    cudaStreamDestroy(h->copy_stream);
  }
  delete h;
  return B200TSDF_OK;
}

extern "C" int b200tsdf_reset(b200tsdf_t *h, void *stream) {
  B200_NVTX("b200tsdf_reset");
  if (!h) return B200TSDF_E_INVALID;
  cudaStream_t st = (cudaStream_t)stream;
  B200_CUDA(cudaSetDevice(h->cfg.device));
  B200_CUDA(cudaMemsetAsync(h->d.keys, 0xFF, h->cap * sizeof(unsigned long long), st));
  B200_CUDA(cudaMemsetAsync(h->d.vox, 0, h->cap * B200_BLOCK_VOX * sizeof(float4), st));
  B200_CUDA(cudaMemsetAsync(h->d.cnt, 0, h->cap * B200_BLOCK_VOX * sizeof(unsigned long long), st));
  B200_CUDA(cudaMemsetAsync(h->d.ctr, 0, CTR_COPIES * CTR_COUNT * sizeof(unsigned long long), st));
  if (h->d.dir_on)
    B200_CUDA(cudaMemsetAsync(h->d.dir, 0xFF, h->dir_entries * sizeof(uint32_t), st));
  return B200TSDF_OK;
}

extern "C" int b200tsdf_counters(b200tsdf_t *h, int64_t out_host[8], void *stream) {
  B200_NVTX("b200tsdf_counters");
  if (!h || !out_host) return B200TSDF_E_INVALID;
  cudaStream_t st = (cudaStream_t)stream;
  unsigned long long tmp[CTR_COUNT];
  int rc = b200_read_counters(h, st, tmp);
  if (rc) return rc;
  for (int i = 0; i < 8; ++i) out_host[i] = (int64_t)tmp[i];
  if (tmp[CTR_OVERFLOW] != 0) { // the values above are valid; the volume is not
    b200_set_error("voxel-block hash full: %llu updates were lost (capacity_log2 = %d); raise "
                   "tsdf_global.capacity_log2", tmp[CTR_OVERFLOW], h->cfg.capacity_log2);
    return B200TSDF_E_CAPACITY;
  }
  return B200TSDF_OK;
}

// Sum of the counter replicas: one CTA, thread c holds replica c.
__global__ void __launch_bounds__(CTR_COPIES)
    sum_counters_kernel(const unsigned long long *__restrict__ ctr, unsigned long long *out) {
  __shared__ unsigned long long part[CTR_COPIES / 32][CTR_COUNT];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int i = 0; i < CTR_COUNT; ++i) {
    unsigned long long x = ctr[(size_t)threadIdx.x * CTR_COUNT + i];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) x += __shfl_down_sync(0xFFFFFFFFu, x, d);
    if (lane == 0) part[warp][i] = x;
  }
  __syncthreads();
  if (threadIdx.x < CTR_COUNT) {
    unsigned long long t = 0;
    for (int w = 0; w < CTR_COPIES / 32; ++w) t += part[w][threadIdx.x];
    out[threadIdx.x] = t;
  }
}

int b200_read_counters(b200tsdf *h, cudaStream_t st, unsigned long long out[CTR_COUNT]) {
  B200_CUDA(cudaSetDevice(h->cfg.device));
  sum_counters_kernel<<<1, CTR_COPIES, 0, st>>>(h->d.ctr, h->scratch + 16);
  B200_CUDA(cudaGetLastError());
  B200_CUDA(cudaMemcpyAsync(out, h->scratch + 16, CTR_COUNT * sizeof(unsigned long long),
                            cudaMemcpyDeviceToHost, st));
  B200_CUDA(cudaStreamSynchronize(st));
  return B200TSDF_OK;
}

// Grow the device staging area used by the *_host entry points.
int b200_ensure_stage(b200tsdf *h, size_t bytes) {
  if (h->stage_bytes >= bytes) return B200TSDF_OK;
  if (h->stage_dev) B200_CUDA(cudaFree(h->stage_dev));
  h->stage_dev = nullptr;
  h->stage_bytes = 0;
  B200_CUDA(cudaMalloc(&h->stage_dev, bytes));
  h->stage_bytes = bytes;
  return B200TSDF_OK;
}
