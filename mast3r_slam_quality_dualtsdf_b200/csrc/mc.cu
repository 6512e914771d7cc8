// mc.cu -- marching-cubes extraction over the allocated voxel blocks.
//
// No reference counterpart (the reference has no mesh path):
// semantics are stated in oracle/mc_oracle.c and checked against it.
// A pre-pass resolves every block's 8 neighbour slots through the block hash
// (one thread per lookup).  Then one CTA per voxel block: the 9x9x9 corner
// lattice (block + one-voxel halo from the +x/+y/+z neighbour blocks) is
// staged in shared memory with 128-bit loads; every thread classifies two
// cubes; triangle slots are assigned by a warp scan (shfl_up) and ONE
// atomicAdd per warp on the global triangle counter (warp-scan compaction);
// triangles are then emitted one per lane (balanced within the warp).
#include <math.h>
#include <stdlib.h>

#include <mutex>

#include "common.cuh"

#define MC_TABLE_QUAL __device__ __align__(16) const
#include "mc_tables.h"

#define MC_THREADS 256

// Pass 0: the 8 block slots every block's 9^3 corner lattice draws from (itself
// and its +x/+y/+z neighbours), one thread per (block, neighbour): the
// dependent hash probes run massively parallel instead of serially per CTA.
__global__ void mc_neighbors_kernel(VolumeDev v, uint32_t *__restrict__ nbr,
                                    unsigned long long *__restrict__ n_tris,
                                    uint32_t *__restrict__ sched) {
  // the triangle counter and the block ticket of the pass after
  if (blockIdx.x == 0 && threadIdx.x == 0) *n_tris = 0ull, *sched = 0u;
  const uint32_t nb = (uint32_t)v.ctr[CTR_BLOCKS];
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < (long long)nb * 8;
       i += (long long)gridDim.x * blockDim.x) {
    const uint32_t b = (uint32_t)(i >> 3), k = (uint32_t)(i & 7);
    const uint32_t slot = v.list[b];
    uint32_t s = slot;
    if (k) {
      int bx, by, bz;
      unpack_block_key(v.keys[slot], bx, by, bz);
      int nx = bx + (k & 1), ny = by + ((k >> 1) & 1), nz = bz + (k >> 2);
      s = block_in_range(nx, ny, nz) ? hash_find(v, pack_block_key(nx, ny, nz)) : B200_NO_SLOT;
    }
    nbr[i] = s;
  }
}

// cp.async (LDGSTS) helpers: the raw voxel words of the NEXT block travel to
// shared memory while the current block is classified and emitted.
__device__ __forceinline__ void cp_async8(void *smem, const void *gmem) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(
                   (uint32_t)__cvta_generic_to_shared(smem)),
               "l"(gmem)
               : "memory");
}
__device__ __forceinline__ void cp_async4(void *smem, const void *gmem) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(
                   (uint32_t)__cvta_generic_to_shared(smem)),
               "l"(gmem)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

#define MC_CTAS_PER_SM 5
#define MC_PTS 3 // lattice points per thread: ceil(729 / 256)

// Persistent CTAs, one voxel block per iteration, double-buffered: while block
// b is converted / classified / emitted, the (S, W) pair and the hit counter of
// every corner of block b+1's 9^3 lattice (block + halo from the +x/+y/+z
// neighbours) are in flight to the other raw buffer (8-byte + 4-byte cp.async
// per corner: only the words of the selected volume), the neighbour slots of
// block b+2 and the key of block b+1 are in flight to registers.
template <bool WRITE>
__global__ void __launch_bounds__(MC_THREADS, MC_CTAS_PER_SM)
    mc_kernel(VolumeDev v, const uint32_t *__restrict__ nbr, int which, float vs,
              float *__restrict__ verts, long long tri_cap, unsigned long long *n_tris) {
  __shared__ float tv[9 * 9 * 9];
  __shared__ __align__(16) float2 r_sw[2][9 * 9 * 9 + 7];
  __shared__ uint32_t r_h[2][9 * 9 * 9 + 7];
  __shared__ uint32_t s_nb[8];
  __shared__ unsigned long long s_key;
  __shared__ __align__(16) signed char s_tri[256 * 16];
  __shared__ unsigned char s_num[256];
  const int tid = threadIdx.x, lane = tid & 31;
  // case tables -> shared memory, 128 bits per thread (one row of the table)
  if (tid < 256) {
    reinterpret_cast<int4 *>(s_tri)[tid] = reinterpret_cast<const int4 *>(MC_TRI_TABLE)[tid];
    s_num[tid] = MC_NUM_TRIS[tid];
  }
  // this thread's lattice points: local voxel index | neighbour id << 9 (block-invariant)
  int pk[MC_PTS];
#pragma unroll
  for (int j = 0; j < MC_PTS; ++j) {
    const int i = tid + j * MC_THREADS;
    const int x = i % 9, y = (i / 9) % 9, z = i / 81;
    const int nbk = (x == 8) | ((y == 8) << 1) | ((z == 8) << 2);
    pk[j] = i < 729 ? ((((z & 7) << 6) | ((y & 7) << 3) | (x & 7)) | (nbk << 9)) : -1;
  }
  const char *vox_b = reinterpret_cast<const char *>(v.vox) + which * 8;
  const char *cnt_b = reinterpret_cast<const char *>(v.cnt) + which * 4;
  auto prefetch = [&](int buf) { // raw words of the block whose neighbour slots are in s_nb
#pragma unroll
    for (int j = 0; j < MC_PTS; ++j) {
      if (pk[j] < 0) continue;
      const int i = tid + j * MC_THREADS;
      const uint32_t s = s_nb[pk[j] >> 9];
      if (s != B200_NO_SLOT) {
        const size_t off = (size_t)s * B200_BLOCK_VOX + (pk[j] & 511);
        cp_async8(&r_sw[buf][i], vox_b + off * sizeof(float4));
        cp_async4(&r_h[buf][i], cnt_b + off * sizeof(unsigned long long));
      } else {
        r_h[buf][i] = 0u; // no block there: never observed
      }
    }
    cp_async_commit();
  };
  const uint32_t nb = (uint32_t)v.ctr[CTR_BLOCKS];
  const uint32_t g = gridDim.x;
  uint32_t b = blockIdx.x;
  if (b >= nb) return;
  uint32_t nbA = B200_NO_SLOT;      // threads 0..7: neighbour slot of the block after the current one
  unsigned long long keyA = 0;      // thread 0: key of the current block
  if (tid < 8) s_nb[tid] = nbr[(size_t)b * 8 + tid];
  __syncthreads();
  prefetch(0);
  if (tid == 0) keyA = v.keys[s_nb[0]];
  if (tid < 8 && b + g < nb) nbA = nbr[(size_t)(b + g) * 8 + tid];
  __syncthreads(); // s_nb is rewritten next
  for (uint32_t it = 0; b < nb; b += g, ++it) {
    const int buf = it & 1;
    if (tid < 8) s_nb[tid] = nbA;
    if (tid == 0) s_key = keyA;
    __syncthreads(); // previous block fully emitted; s_nb / s_key of this iteration visible
    if (b + g < nb) {
      prefetch(buf ^ 1);
      if (tid == 0) keyA = v.keys[s_nb[0]];
      if (tid < 8) nbA = b + 2 * g < nb ? nbr[(size_t)(b + 2 * g) * 8 + tid] : B200_NO_SLOT;
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads(); // raw words of block b have landed for every thread
#pragma unroll
    for (int j = 0; j < MC_PTS; ++j) {
      if (pk[j] < 0) continue;
      const int i = tid + j * MC_THREADS;
      const uint32_t hits = r_h[buf][i];
      float val = NAN;
      if (hits != 0) { // voxel_view (common.cuh) on the staged words
        const float2 sw = r_sw[buf][i];
        const float w = (hits == 1) ? sw.y : fminf(sw.y, v.max_weight);
        if (w >= v.min_weight) val = sw.x / fmaxf(1.0e-9f, w);
      }
      tv[i] = val;
    }
    __syncthreads();
    int bx, by, bz;
    unpack_block_key(s_key, bx, by, bz);
    int mask[2], nt[2];
#pragma unroll
    for (int c2 = 0; c2 < 2; ++c2) {
      int c = tid + c2 * MC_THREADS;
      int x = c & 7, y = (c >> 3) & 7, z = c >> 6;
      int m = 0;
      bool ok = true;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        float t = tv[(z + (k >> 2)) * 81 + (y + ((k >> 1) & 1)) * 9 + x + (k & 1)];
        ok = ok && !isnan(t);
        m |= (t < 0.0f) ? (1 << k) : 0;
      }
      mask[c2] = m;
      nt[c2] = ok ? s_num[m] : 0;
    }
    // warp-scan compaction of the per-thread triangle counts: an inclusive
    // shfl_up scan, then ONE atomicAdd per warp reserves the warp's output run
    uint32_t mine = nt[0] + nt[1], incl = mine;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      uint32_t o = __shfl_up_sync(0xFFFFFFFFu, incl, d);
      if (lane >= d) incl += o;
    }
    const uint32_t wtot = __shfl_sync(0xFFFFFFFFu, incl, 31);
    if (wtot == 0) continue;
    unsigned long long wbase = 0;
    if (lane == 31) wbase = atomicAdd(n_tris, (unsigned long long)wtot);
    wbase = __shfl_sync(0xFFFFFFFFu, wbase, 31);
    if (WRITE) {
      // balanced emission: triangle k of the warp's run goes to lane k % 32; its
      // source cube is found by a binary search over the scan (5 shuffles)
      const uint32_t excl = incl - mine;
      for (uint32_t k0 = 0; k0 < wtot; k0 += 32) {
        const uint32_t k = k0 + lane;
        const bool act = k < wtot;
        int src = 0;
#pragma unroll
        for (int step = 16; step > 0; step >>= 1) {
          uint32_t vv = __shfl_sync(0xFFFFFFFFu, incl, src + step - 1);
          if (act && vv <= k) src += step;
        }
        const uint32_t e_s = __shfl_sync(0xFFFFFFFFu, excl, src);
        const int n0_s = __shfl_sync(0xFFFFFFFFu, nt[0], src);
        const int m0_s = __shfl_sync(0xFFFFFFFFu, mask[0], src);
        const int m1_s = __shfl_sync(0xFFFFFFFFu, mask[1], src);
        if (!act) continue;
        int q = (int)(k - e_s);
        const int c2 = q >= n0_s ? 1 : 0;
        if (c2) q -= n0_s;
        const int m = c2 ? m1_s : m0_s;
        const int c = (tid - lane + src) + c2 * MC_THREADS;
        const int x = c & 7, y = (c >> 3) & 7, z = c >> 6;
        const long long o = (long long)wbase + k;
        if (o >= tri_cap) continue;
#pragma unroll
        for (int vtx = 0; vtx < 3; ++vtx) {
          int e = s_tri[m * 16 + 3 * q + vtx];
          int a = MC_EDGE_CORNERS[e][0], bb = MC_EDGE_CORNERS[e][1];
          int ax = x + (a & 1), ay = y + ((a >> 1) & 1), az = z + (a >> 2);
          int cx = x + (bb & 1), cy = y + ((bb >> 1) & 1), cz = z + (bb >> 2);
          float ta = tv[az * 81 + ay * 9 + ax], tb = tv[cz * 81 + cy * 9 + cx];
          float mu = __fdiv_rn(ta, __fsub_rn(ta, tb));
          float pa[3] = {__fmul_rn((float)(bx * 8 + ax), vs), __fmul_rn((float)(by * 8 + ay), vs),
                         __fmul_rn((float)(bz * 8 + az), vs)};
          float pb[3] = {__fmul_rn((float)(bx * 8 + cx), vs), __fmul_rn((float)(by * 8 + cy), vs),
                         __fmul_rn((float)(bz * 8 + cz), vs)};
#pragma unroll
          for (int d = 0; d < 3; ++d)
            verts[o * 9 + vtx * 3 + d] = __fadd_rn(pa[d], __fmul_rn(__fsub_rn(pb[d], pa[d]), mu));
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------
// mc_kernel2: the same extraction with the classification done on BIT PLANES.
// mc_kernel spends 8 shared loads + 16 ALU per cube on the case index and three
// CTA barriers per block on unevenly loaded warps (ncu: barrier stall 5.6 per
// issue, 34 % of HBM).  Here the conversion pass also ballots a sign and a
// validity bit per lattice point into two 729-bit planes; 64 threads then
// classify the block's 64 cube ROWS with 9-bit extracts (a row of 8 cubes in
// ~40 instructions: all-valid / all-inside / all-outside by AND of shifted
// rows), list the few active cubes with their case index, and a two-warp scan
// gives every active cube its triangle offset -- ONE atomicAdd per block.
// Emission is one triangle per thread (binary search over the cube scan), so
// stores are coalesced and no lane idles behind a 5-triangle cube.  Blocks
// without a sign change (most of a band's blocks) stop after the row pass.
#define MC2_CUBE_CAP 512
#ifndef MC2_STAGES
#define MC2_STAGES 2 // raw-word buffers: the block being processed + MC2_STAGES - 1 in flight
#endif
template <bool WRITE>
__global__ void __launch_bounds__(MC_THREADS, MC_CTAS_PER_SM)
    mc_kernel2(VolumeDev v, const uint32_t *__restrict__ nbr, int which, float vs,
               float *__restrict__ verts, long long tri_cap, unsigned long long *n_tris) {
  __shared__ float tv[9 * 9 * 9 + 7];
  __shared__ __align__(16) float2 r_sw[MC2_STAGES][9 * 9 * 9 + 7];
  __shared__ uint32_t r_h[MC2_STAGES][9 * 9 * 9 + 7];
  __shared__ uint32_t s_nb[8];
  __shared__ __align__(8) unsigned long long s_key[MC2_STAGES], s_base; // block key per stage
  __shared__ __align__(16) signed char s_tri[256 * 16];
  __shared__ unsigned char s_num[256];
  __shared__ uint32_t p_sgn[24], p_val[24]; // bit planes over the lattice index z*81 + y*9 + x
  __shared__ uint16_t s_cube[MC2_CUBE_CAP];       // active cubes: cube id (9 bits)
  __shared__ unsigned char s_case[MC2_CUBE_CAP];  // ... and their case index
  __shared__ uint16_t s_scan[MC2_CUBE_CAP + 2];   // exclusive scan of their triangle counts
  __shared__ uint32_t s_w0, s_tot;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid < 256) {
    reinterpret_cast<int4 *>(s_tri)[tid] = reinterpret_cast<const int4 *>(MC_TRI_TABLE)[tid];
    s_num[tid] = MC_NUM_TRIS[tid];
  }
  int pk[MC_PTS];
#pragma unroll
  for (int j = 0; j < MC_PTS; ++j) {
    const int i = tid + j * MC_THREADS;
    const int x = i % 9, y = (i / 9) % 9, z = i / 81;
    const int nbk = (x == 8) | ((y == 8) << 1) | ((z == 8) << 2);
    pk[j] = i < 729 ? ((((z & 7) << 6) | ((y & 7) << 3) | (x & 7)) | (nbk << 9)) : -1;
  }
  const char *vox_b = reinterpret_cast<const char *>(v.vox) + which * 8;
  const char *cnt_b = reinterpret_cast<const char *>(v.cnt) + which * 4;
  // raw words + key of the block whose neighbour slots are in s_nb -> stage `buf`
  // (always one commit group per call, empty when there is no such block)
  auto prefetch = [&](int buf, bool live) {
    if (live) {
#pragma unroll
      for (int j = 0; j < MC_PTS; ++j) {
        if (pk[j] < 0) continue;
        const int i = tid + j * MC_THREADS;
        const uint32_t s = s_nb[pk[j] >> 9];
        if (s != B200_NO_SLOT) {
          const size_t off = (size_t)s * B200_BLOCK_VOX + (pk[j] & 511);
          cp_async8(&r_sw[buf][i], vox_b + off * sizeof(float4));
          cp_async4(&r_h[buf][i], cnt_b + off * sizeof(unsigned long long));
        } else {
          r_h[buf][i] = 0u;
        }
      }
      if (tid == 0) cp_async8(&s_key[buf], v.keys + s_nb[0]);
    }
    cp_async_commit();
  };
  const uint32_t nb = (uint32_t)v.ctr[CTR_BLOCKS];
  const uint32_t g = gridDim.x;
  uint32_t b = blockIdx.x;
  if (b >= nb) return;
  // prologue: blocks b, b + g, ... of this CTA into stages 0 .. MC2_STAGES - 2
#pragma unroll 1
  for (int pre = 0; pre < MC2_STAGES - 1; ++pre) {
    const uint32_t bb = b + (uint32_t)pre * g;
    if (tid < 8 && bb < nb) s_nb[tid] = nbr[(size_t)bb * 8 + tid];
    __syncthreads();
    prefetch(pre, bb < nb);
    __syncthreads(); // s_nb is rewritten next
  }
  uint32_t nbA = B200_NO_SLOT; // threads 0..7: neighbour slots of the block MC2_STAGES - 1 ahead
  if (tid < 8 && b + (MC2_STAGES - 1) * g < nb) nbA = nbr[(size_t)(b + (MC2_STAGES - 1) * g) * 8 + tid];
  int buf = 0, pbuf = MC2_STAGES - 1;
  for (; b < nb; b += g) {
    if (tid < 8) s_nb[tid] = nbA;
    __syncthreads(); // previous block fully emitted (its stage is free); s_nb visible
    prefetch(pbuf, b + (MC2_STAGES - 1) * g < nb);
    if (tid < 8)
      nbA = b + MC2_STAGES * g < nb ? nbr[(size_t)(b + MC2_STAGES * g) * 8 + tid] : B200_NO_SLOT;
    cp_async_wait<MC2_STAGES - 1>();
    __syncthreads(); // raw words of block b have landed for every thread
    const int cur = buf;
    buf = buf + 1 == MC2_STAGES ? 0 : buf + 1;
    pbuf = pbuf + 1 == MC2_STAGES ? 0 : pbuf + 1;
    // ---- conversion: VoxelData view of every lattice point + the two bit planes
#pragma unroll
    for (int j = 0; j < MC_PTS; ++j) {
      const int i = tid + j * MC_THREADS;
      float val = NAN;
      if (pk[j] >= 0) {
        const uint32_t hits = r_h[cur][i];
        if (hits != 0) { // voxel_view (common.cuh) on the staged words
          const float2 sw = r_sw[cur][i];
          const float w = (hits == 1) ? sw.y : fminf(sw.y, v.max_weight);
          if (w >= v.min_weight) val = sw.x / fmaxf(1.0e-9f, w);
        }
        tv[i] = val;
      }
      const unsigned bs = __ballot_sync(0xFFFFFFFFu, val < 0.0f);
      const unsigned bv = __ballot_sync(0xFFFFFFFFu, !isnan(val));
      if (lane == 0) p_sgn[warp + 8 * j] = bs, p_val[warp + 8 * j] = bv;
    }
    __syncthreads();
    // ---- rows: thread r < 64 owns the 8 cubes (x = 0..7, y = r & 7, z = r >> 3)
    if (warp < 2) {
      const int y = tid & 7, z = tid >> 3, base = 81 * z + 9 * y;
      auto ext = [&](const uint32_t *a, int pos) {
        const int w = pos >> 5;
        return __funnelshift_r(a[w], w + 1 < 24 ? a[w + 1] : 0u, pos & 31) & 0x1FFu;
      };
      const uint32_t s0 = ext(p_sgn, base), s1 = ext(p_sgn, base + 9), s2 = ext(p_sgn, base + 81),
                     s3 = ext(p_sgn, base + 90);
      const uint32_t v0 = ext(p_val, base), v1 = ext(p_val, base + 9), v2 = ext(p_val, base + 81),
                     v3 = ext(p_val, base + 90);
      const uint32_t vc = (v0 & (v0 >> 1)) & (v1 & (v1 >> 1)) & (v2 & (v2 >> 1)) & (v3 & (v3 >> 1));
      const uint32_t in1 = (s0 & (s0 >> 1)) & (s1 & (s1 >> 1)) & (s2 & (s2 >> 1)) & (s3 & (s3 >> 1));
      const uint32_t o0 = ~s0, o1 = ~s1, o2 = ~s2, o3 = ~s3;
      const uint32_t in0 = (o0 & (o0 >> 1)) & (o1 & (o1 >> 1)) & (o2 & (o2 >> 1)) & (o3 & (o3 >> 1));
      uint32_t act = vc & ~(in1 | in0) & 0xFFu; // all 8 corners observed, signs mixed
      uint32_t ta = 0;
      for (uint32_t a = act; a; a &= a - 1) {
        const int x = __ffs(a) - 1;
        const uint32_t m = ((s0 >> x) & 3u) | (((s1 >> x) & 3u) << 2) | (((s2 >> x) & 3u) << 4) |
                           (((s3 >> x) & 3u) << 6);
        ta += s_num[m];
      }
      // two-warp scan of (cubes << 16 | triangles) over the 64 rows
      const uint32_t mine = ((uint32_t)__popc(act) << 16) | ta;
      uint32_t incl = mine;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, incl, d);
        if (lane >= d) incl += o;
      }
      if (tid == 31) s_w0 = incl;
      asm volatile("bar.sync 1, 64;" ::: "memory"); // the two row warps only
      if (warp == 1) incl += s_w0;
      uint32_t co = (incl - mine) >> 16, to = (incl - mine) & 0xFFFFu;
      for (uint32_t a = act; a; a &= a - 1) {
        const int x = __ffs(a) - 1;
        const uint32_t m = ((s0 >> x) & 3u) | (((s1 >> x) & 3u) << 2) | (((s2 >> x) & 3u) << 4) |
                           (((s3 >> x) & 3u) << 6);
        s_cube[co] = (uint16_t)(tid * 8 + x);
        s_case[co] = (unsigned char)m;
        s_scan[co] = (uint16_t)to;
        ++co;
        to += s_num[m];
      }
      if (tid == 63) { // totals: ONE atomicAdd reserves the block's output run
        const uint32_t A = incl >> 16, T = incl & 0xFFFFu;
        s_scan[A] = (uint16_t)T;
        s_tot = incl;
        if (T) s_base = atomicAdd(n_tris, (unsigned long long)T);
      }
    }
    __syncthreads();
    if (!WRITE) continue;
    const uint32_t A = s_tot >> 16, T = s_tot & 0xFFFFu;
    if (T == 0) continue;
    int bx, by, bz;
    unpack_block_key(s_key[cur], bx, by, bz);
    const unsigned long long base_o = s_base;
    // ---- emission: one triangle per thread
    for (uint32_t k = tid; k < T; k += MC_THREADS) {
      uint32_t lo = 0, hi = A; // largest c with s_scan[c] <= k
      while (hi - lo > 1) {
        const uint32_t mid = (lo + hi) >> 1;
        if (s_scan[mid] <= k) lo = mid; else hi = mid;
      }
      const int q = (int)(k - s_scan[lo]), c = (int)s_cube[lo], m = (int)s_case[lo];
      const int x = c & 7, y = (c >> 3) & 7, z = c >> 6;
      const long long o = (long long)base_o + k;
      if (o >= tri_cap) continue;
#pragma unroll
      for (int vtx = 0; vtx < 3; ++vtx) {
        const int e = s_tri[m * 16 + 3 * q + vtx];
        const int a = MC_EDGE_CORNERS[e][0], bb = MC_EDGE_CORNERS[e][1];
        const int ax = x + (a & 1), ay = y + ((a >> 1) & 1), az = z + (a >> 2);
        const int cx = x + (bb & 1), cy = y + ((bb >> 1) & 1), cz = z + (bb >> 2);
        const float ta = tv[az * 81 + ay * 9 + ax], tb = tv[cz * 81 + cy * 9 + cx];
        const float mu = __fdiv_rn(ta, __fsub_rn(ta, tb));
        const float pa[3] = {__fmul_rn((float)(bx * 8 + ax), vs), __fmul_rn((float)(by * 8 + ay), vs),
                             __fmul_rn((float)(bz * 8 + az), vs)};
        const float pb[3] = {__fmul_rn((float)(bx * 8 + cx), vs), __fmul_rn((float)(by * 8 + cy), vs),
                             __fmul_rn((float)(bz * 8 + cz), vs)};
#pragma unroll
        for (int d = 0; d < 3; ++d)
          verts[o * 9 + vtx * 3 + d] = __fadd_rn(pa[d], __fmul_rn(__fsub_rn(pb[d], pa[d]), mu));
      }
    }
  }
}

#include "mc3.cuh" // mc_kernel3: one warp per voxel block (the default)

static int mc_launch(b200tsdf_t *h, int which, float *verts_dev, int64_t tri_cap,
                     unsigned long long *n_tris_dev, cudaStream_t st) {
  B200_CUDA(cudaSetDevice(h->cfg.device));
  const float vs = (float)h->cfg.voxel_size;
  if (!h->nbr) B200_CUDA(cudaMalloc(&h->nbr, h->cap * 8 * sizeof(uint32_t)));
  const unsigned sms = (unsigned)(h->sm_count > 0 ? h->sm_count : 148);
  uint32_t *sched = reinterpret_cast<uint32_t *>(h->scratch + 8); // block ticket of mc_kernel3
  mc_neighbors_kernel<<<sms * 8, 256, 0, st>>>(h->d, h->nbr, n_tris_dev, sched);
  static const int ver = [] {
    // A/B knob: B200TSDF_MC = 1 per-cube kernel, 2 bit-plane kernel (CTA per block), 3 warp per block
    const char *e = getenv("B200TSDF_MC");
    const char *o = getenv("B200TSDF_MC_V1");
    return e ? atoi(e) : (o && atoi(o) ? 1 : 3);
  }();
  const bool wr = verts_dev && tri_cap > 0;
  const unsigned grid = sms * MC_CTAS_PER_SM;
  if (ver == 1) {
    if (wr)
      mc_kernel<true><<<grid, MC_THREADS, 0, st>>>(h->d, h->nbr, which, vs, verts_dev,
                                                   (long long)tri_cap, n_tris_dev);
    else
      mc_kernel<false><<<grid, MC_THREADS, 0, st>>>(h->d, h->nbr, which, vs, nullptr, 0, n_tris_dev);
  } else if (ver == 2) {
    if (wr)
      mc_kernel2<true><<<grid, MC_THREADS, 0, st>>>(h->d, h->nbr, which, vs, verts_dev,
                                                    (long long)tri_cap, n_tris_dev);
    else
      mc_kernel2<false><<<grid, MC_THREADS, 0, st>>>(h->d, h->nbr, which, vs, nullptr, 0, n_tris_dev);
  } else {
    static bool attr_set = false;
    if (!attr_set) {
      B200_CUDA(cudaFuncSetAttribute(mc_kernel3<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)sizeof(Mc3Smem)));
      B200_CUDA(cudaFuncSetAttribute(mc_kernel3<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)sizeof(Mc3Smem)));
      attr_set = true;
    }
    const unsigned g3 = sms * MC3_CTAS_PER_SM;
    if (wr)
      mc_kernel3<true><<<g3, MC3_WARPS * 32, sizeof(Mc3Smem), st>>>(
          h->d, h->nbr, which, vs, verts_dev, (long long)tri_cap, n_tris_dev, sched);
    else
      mc_kernel3<false><<<g3, MC3_WARPS * 32, sizeof(Mc3Smem), st>>>(h->d, h->nbr, which, vs, nullptr,
                                                                    0, n_tris_dev, sched);
  }
  B200_CUDA(cudaGetLastError());
  return B200TSDF_OK;
}

extern "C" int b200tsdf_extract_mesh_async(b200tsdf_t *h, int which, float *verts_dev,
                                           int64_t tri_cap, uint64_t *n_tris_dev, void *stream) {
  B200_NVTX("b200tsdf_extract_mesh_async");
  if (!h || !n_tris_dev || (which != 0 && which != 1) || tri_cap < 0) {
    b200_set_error("b200tsdf_extract_mesh_async: invalid argument");
    return B200TSDF_E_INVALID;
  }
  return mc_launch(h, which, verts_dev, tri_cap, (unsigned long long *)n_tris_dev, (cudaStream_t)stream);
}

extern "C" int b200tsdf_extract_mesh(b200tsdf_t *h, int which, float *verts_dev, int64_t tri_cap,
                                     int64_t *n_tris_host, void *stream) {
  B200_NVTX("b200tsdf_extract_mesh");
  if (!h || !n_tris_host || (which != 0 && which != 1) || tri_cap < 0) {
    b200_set_error("b200tsdf_extract_mesh: invalid argument");
    return B200TSDF_E_INVALID;
  }
  cudaStream_t st = (cudaStream_t)stream;
  int rc = mc_launch(h, which, verts_dev, tri_cap, h->scratch, st);
  if (rc) return rc;
  unsigned long long n;
  B200_CUDA(cudaMemcpyAsync(&n, h->scratch, 8, cudaMemcpyDeviceToHost, st));
  B200_CUDA(cudaStreamSynchronize(st));
  *n_tris_host = (int64_t)n;
  return B200TSDF_OK;
}

// ---------------------------------------------------------------------------
// b200tsdf_extract_mesh_static: the extraction pass for callers that may not
// allocate (CUDA graph capture, tight serving loops).  No counterpart in the
// reference.
// b200tsdf_extract_mesh[_async] allocates the neighbour table of the pass lazily
// (cudaMalloc of 32 bytes per hash slot on the first call).  Here the table and
// the block ticket of mc_kernel3 live in a file-scope device array of a fixed
// size, so the entry calls no cudaMalloc / cudaFree at all; volumes with more
// than B200TSDF_STATIC_MESH_MAX_SLOTS hash slots are rejected.  Every GPU that
// loads the library holds the array (2 MB + 256 B) until the process exits.
// All calls on a device share it: two calls that overlap on the GPU (different
// streams, or different volumes) race on the table.  Nothing in it has to
// survive a call: the neighbour pass rewrites the table rows it uses and zeroes
// the ticket and the caller's triangle counter.
#define MC_STATIC_HEAD 256 // bytes in front of the table: [0] block ticket
static __device__ __align__(256) unsigned char
    mc_static_scratch[MC_STATIC_HEAD + (size_t)B200TSDF_STATIC_MESH_MAX_SLOTS * 8 * sizeof(uint32_t)];

#define MC_STATIC_MAX_DEVICES 64
// Device address of mc_static_scratch per device.  In host code the array's name is only a
// stand-in for the symbol: it must never be handed to a kernel.
static void *mc_static_ptr[MC_STATIC_MAX_DEVICES];
static std::mutex mc_static_mu;

// Set-up, once per device, with that device current.
static int mc_static_setup(int device, void **out) {
  if (device < 0 || device >= MC_STATIC_MAX_DEVICES) {
    b200_set_error("b200tsdf_extract_mesh_static: device %d out of range", device);
    return B200TSDF_E_INVALID;
  }
  std::lock_guard<std::mutex> lk(mc_static_mu);
  if (!mc_static_ptr[device]) {
    void *p = nullptr;
    B200_CUDA(cudaGetSymbolAddress(&p, mc_static_scratch));
    B200_CUDA(cudaFuncSetAttribute(mc_kernel3<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)sizeof(Mc3Smem)));
    B200_CUDA(cudaFuncSetAttribute(mc_kernel3<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   (int)sizeof(Mc3Smem)));
    mc_static_ptr[device] = p;
  }
  *out = mc_static_ptr[device];
  return B200TSDF_OK;
}

extern "C" int b200tsdf_extract_mesh_static(b200tsdf_t *h, int which, float *verts_dev,
                                            int64_t tri_cap, uint64_t *n_tris_dev, void *stream) {
  B200_NVTX("b200tsdf_extract_mesh_static");
  if (!h || !n_tris_dev || (which != 0 && which != 1) || tri_cap < 0) {
    b200_set_error("b200tsdf_extract_mesh_static: invalid argument");
    return B200TSDF_E_INVALID;
  }
  if ((uint64_t)h->cap > (uint64_t)B200TSDF_STATIC_MESH_MAX_SLOTS) {
    b200_set_error("b200tsdf_extract_mesh_static: the volume has %llu hash slots, the static "
                   "neighbour table holds %d (capacity_log2 <= 16); use b200tsdf_extract_mesh",
                   (unsigned long long)h->cap, B200TSDF_STATIC_MESH_MAX_SLOTS);
    return B200TSDF_E_INVALID;
  }
  cudaStream_t st = (cudaStream_t)stream;
  B200_CUDA(cudaSetDevice(h->cfg.device));
  void *base = nullptr;
  int rc = mc_static_setup(h->cfg.device, &base);
  if (rc) return rc;
  uint32_t *sched = reinterpret_cast<uint32_t *>(base);
  uint32_t *nbr = reinterpret_cast<uint32_t *>(static_cast<unsigned char *>(base) + MC_STATIC_HEAD);
  unsigned long long *n_tris = reinterpret_cast<unsigned long long *>(n_tris_dev);
  const float vs = (float)h->cfg.voxel_size;
  const unsigned sms = (unsigned)(h->sm_count > 0 ? h->sm_count : 148);
  mc_neighbors_kernel<<<sms * 8, 256, 0, st>>>(h->d, nbr, n_tris, sched);
  const unsigned g3 = sms * MC3_CTAS_PER_SM;
  if (verts_dev && tri_cap > 0)
    mc_kernel3<true><<<g3, MC3_WARPS * 32, sizeof(Mc3Smem), st>>>(h->d, nbr, which, vs, verts_dev,
                                                                 (long long)tri_cap, n_tris, sched);
  else
    mc_kernel3<false><<<g3, MC3_WARPS * 32, sizeof(Mc3Smem), st>>>(h->d, nbr, which, vs, nullptr, 0,
                                                                  n_tris, sched);
  B200_CUDA(cudaGetLastError());
  return B200TSDF_OK;
}
