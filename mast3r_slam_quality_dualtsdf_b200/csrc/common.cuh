// common.cuh -- shared device/host definitions of libb200tsdf (sm_100a only).
//
// Data layout in HBM (DESIGN.md "Data layout"):
//   keys  [cap]        u64   packed block key, EMPTY = ~0     (open addressing)
//   vox   [cap * 512]  float4 {S_p, W_p, S_q, W_q}             (16 B / voxel)
//   cnt   [cap * 512]  u64   lo32 = hits primary, hi32 = hits quality volume
//   list  [cap]        u32   slots in allocation order (n_blocks entries)
// A block is 8x8x8 voxels; the payload of hash slot s is vox[s*512 ...].
// S = sum(tsdf_i * w_i), W = sum(w_i): the reference running average
// (global_volume.py:74-88) is exactly tsdf = S / W_eff with
// W_eff = (hits == 1 ? W : min(W, max_weight)) -- see DESIGN.md "the fold".
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b200tsdf.h"

#define B200_BLOCK_LOG2 3
#define B200_BLOCK_EDGE 8
#define B200_BLOCK_VOX 512
#define B200_EMPTY_KEY 0xFFFFFFFFFFFFFFFFull
#define B200_KEY_BIAS (1 << 20) /* block coords in [-2^20, 2^20) */
#define B200_MAX_PROBE 4096
#define B200_H2D_CHUNKS 8

enum {
  CTR_RAYS = 0,
  CTR_EV_P = 1,
  CTR_EV_Q = 2,
  CTR_BLOCKS = 3,
  CTR_DROPPED = 4,
  CTR_OVERFLOW = 5,
  CTR_CULLED = 6, // rays skipped because their band misses this rank's slabs
  CTR_COUNT = 8
};
// Statistics counters are replicated CTR_COPIES times (one 64-byte line per
// copy) and summed on read: a warp adds to copy (its global warp id mod
// CTR_COPIES), so no L2 slice sees a hot address.  CTR_BLOCKS -- the block
// allocation cursor -- lives in copy 0 only.
#define CTR_COPIES 512

struct VolumeDev {
  unsigned long long *keys;
  float4 *vox;
  unsigned long long *cnt;
  uint32_t *list;
  unsigned long long *ctr; // CTR_COPIES x CTR_COUNT counters (see CTR_COPIES)
  uint32_t cap_mask;
  int bounded;
  int kmin[3], kmax[3]; // voxel-key bounds (clamped to int32 range)
  float max_weight, min_weight;
  // bounded volumes: dense block -> slot directory, a CACHE of the hash
  // (entries go EMPTY -> slot once; the hash stays the source of truth)
  uint32_t *dir;
  int dir_on, dir_b0[3], dir_n[3]; // first block coordinate, blocks per axis
  // z-slab ownership (multi-GPU): slab s = (kz - slab_z0) >> slab_shift is
  // owned iff s mod slab_count == slab_index
  int slab_count, slab_index, slab_shift, slab_z0;
  int slab_mask; // slab_count - 1 when slab_count is a power of two, else -1
};

#define B200_NO_SLOT 0xFFFFFFFFu

__host__ __device__ __forceinline__ int slab_owner_of(const VolumeDev &v, int s) {
  if (v.slab_mask >= 0) return s & v.slab_mask; // two's complement: also right for s < 0
  int m = s % v.slab_count;
  return m < 0 ? m + v.slab_count : m;
}
__host__ __device__ __forceinline__ bool slab_owns_z(const VolumeDev &v, int kz) {
  if (v.slab_count <= 1) return true;
  return slab_owner_of(v, (kz - v.slab_z0) >> v.slab_shift) == v.slab_index;
}

struct b200tsdf {
  b200tsdf_config cfg;
  VolumeDev d;
  uint64_t cap;
  uint64_t dir_entries;
  // host staging for the *_host entry points
  void *stage_dev;
  size_t stage_bytes;
  cudaStream_t copy_stream;                 // H2D of chunk k+1 overlaps the fuse of chunk k
  cudaEvent_t copy_ev[B200_H2D_CHUNKS + 1];
  unsigned long long *scratch; // small device scratch (counters for sweeps)
  uint32_t *nbr;               // marching cubes: 8 neighbour slots per block (lazy)
  // multi-GPU fuse: tile list of the classify pass, [0] list length / [1] ticket cursor
  int *tile_list;
  size_t tile_cap;
  unsigned int *sched;
  int sm_count;
  // max_points_per_kf: per-keyframe {hash threshold, salt} of the current batch
  void *sub_thr;
  size_t sub_cap;
  void *opt_state; // PoseOptState of b200tsdf_pose_optimize (sweep.cu), lazy
};

// ---------------------------------------------------------------- hashing
__host__ __device__ __forceinline__ unsigned long long pack_block_key(int bx, int by,
                                                                      int bz) {
  return (unsigned long long)(uint32_t)(bx + B200_KEY_BIAS) |
         ((unsigned long long)(uint32_t)(by + B200_KEY_BIAS) << 21) |
         ((unsigned long long)(uint32_t)(bz + B200_KEY_BIAS) << 42);
}
__host__ __device__ __forceinline__ void unpack_block_key(unsigned long long k, int &bx,
                                                          int &by, int &bz) {
  bx = (int)(k & 0x1FFFFF) - B200_KEY_BIAS;
  by = (int)((k >> 21) & 0x1FFFFF) - B200_KEY_BIAS;
  bz = (int)((k >> 42) & 0x1FFFFF) - B200_KEY_BIAS;
}
__host__ __device__ __forceinline__ bool block_in_range(int bx, int by, int bz) {
  return bx >= -B200_KEY_BIAS && bx < B200_KEY_BIAS && by >= -B200_KEY_BIAS &&
         by < B200_KEY_BIAS && bz >= -B200_KEY_BIAS && bz < B200_KEY_BIAS;
}
__host__ __device__ __forceinline__ uint32_t hash_block_key(unsigned long long k) {
  k ^= k >> 33;
  k *= 0xff51afd7ed558ccdull;
  k ^= k >> 33;
  k *= 0xc4ceb9fe1a85ec53ull;
  k ^= k >> 33;
  return (uint32_t)k;
}

#ifdef __CUDACC__
// Lookup only.  Returns slot or 0xFFFFFFFF.
__device__ __forceinline__ uint32_t hash_find(const VolumeDev &v, unsigned long long key) {
  uint32_t s = hash_block_key(key) & v.cap_mask;
  for (int p = 0; p < B200_MAX_PROBE; ++p) {
    unsigned long long k = __ldcg(v.keys + s);
    if (k == key) return s;
    if (k == B200_EMPTY_KEY) return 0xFFFFFFFFu;
    s = (s + 1) & v.cap_mask;
  }
  return 0xFFFFFFFFu;
}

// Insert-if-absent.  Payload is indexed by slot, so a successful CAS is the
// whole allocation; the winner appends the slot to the block list.
// *fresh is set when this call allocated the block.
__device__ __forceinline__ uint32_t hash_insert(const VolumeDev &v, unsigned long long key,
                                                bool *fresh) {
  uint32_t s = hash_block_key(key) & v.cap_mask;
  *fresh = false;
  for (int p = 0; p < B200_MAX_PROBE; ++p) {
    unsigned long long k = __ldcg(v.keys + s);
    if (k == key) return s;
    if (k == B200_EMPTY_KEY) {
      unsigned long long old = atomicCAS(v.keys + s, B200_EMPTY_KEY, key);
      if (old == B200_EMPTY_KEY) {
        *fresh = true;
        return s;
      }
      if (old == key) return s;
    }
    s = (s + 1) & v.cap_mask;
  }
  return 0xFFFFFFFFu;
}

// 128-bit vector reduction into the packed voxel (sm_90+: red.global.v4.f32).
__device__ __forceinline__ void red_add_v4(float4 *p, float a, float b, float c, float d) {
  asm volatile("red.relaxed.gpu.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p),
               "f"(a), "f"(b), "f"(c), "f"(d)
               : "memory");
}
__device__ __forceinline__ void red_add_u64(unsigned long long *p, unsigned long long x) {
  asm volatile("red.relaxed.gpu.global.add.u64 [%0], %1;" ::"l"(p), "l"(x) : "memory");
}

// Reference view of one packed voxel: VoxelData{tsdf, weight} of volume
// `which` (global_volume.py:9-12, :84-88).  hits == 0 -> absent.
__device__ __forceinline__ bool voxel_view(const float4 &a, unsigned long long c, int which,
                                           float max_weight, float &tsdf, float &w,
                                           uint32_t &hits) {
  float S = which ? a.z : a.x, W = which ? a.w : a.y;
  hits = which ? (uint32_t)(c >> 32) : (uint32_t)c;
  if (hits == 0) return false;
  w = (hits == 1) ? W : fminf(W, max_weight);
  tsdf = S / fmaxf(1.0e-9f, w);
  return true;
}
#endif

#ifdef __CUDACC__
// This warp's replica of the statistics counters.
__device__ __forceinline__ unsigned long long *ctr_replica(const VolumeDev &v) {
  unsigned b = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
  unsigned w = b * (blockDim.x >> 5) + (threadIdx.x >> 5);
  return v.ctr + (size_t)(w & (CTR_COPIES - 1)) * CTR_COUNT;
}
#endif

// ---------------------------------------------------------------- tracing
// NVTX range around every C-ABI entry (the reference's analogue is
// the wall-clock Timer of mast3r_slam/tictoc.py:4-30).  Header-only NVTX3: a no-op
// unless a profiler (nsys / ncu --nvtx) is attached.
#ifndef __CUDA_ARCH__
#include <nvtx3/nvToolsExt.h>
struct B200NvtxRange {
  explicit B200NvtxRange(const char *name) { nvtxRangePushA(name); }
  ~B200NvtxRange() { nvtxRangePop(); }
};
#define B200_NVTX(name) B200NvtxRange b200_nvtx_range__(name)
#else
#define B200_NVTX(name)
#endif

// ---------------------------------------------------------------- errors
void b200_set_error(const char *fmt, ...);
// Sum of the counter replicas (synchronises the stream).
int b200_read_counters(struct b200tsdf *h, cudaStream_t st, unsigned long long out[CTR_COUNT]);
#define B200_CUDA(call)                                                          \
  do {                                                                           \
    cudaError_t e__ = (call);                                                    \
    if (e__ != cudaSuccess) {                                                    \
      b200_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call,                \
                     cudaGetErrorString(e__));                                   \
      return B200TSDF_E_CUDA;                                                    \
    }                                                                            \
  } while (0)
