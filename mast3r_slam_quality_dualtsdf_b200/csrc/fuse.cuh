// fuse.cuh -- warp-converged dual-volume ray march (float32).
//
// Same arithmetic as march_fast.cuh (exact voxel keys, tolerance-level tsdf /
// weight), restructured so that the 32 rays of a warp run ONE converged
// instruction stream:
//   * warp-uniform trip count (max band length over the lanes), lanes that
//     ran out of samples are predicated off -- no `continue`, no per-lane
//     branches in the sample math;
//   * the fold of a finished voxel (address, 128-bit vector reduction +
//     64-bit counter reduction) is executed by the whole warp under a
//     predicate whenever any lane leaves a voxel, instead of once per
//     divergent lane group;
//   * voxel-block slots come from a per-warp direct-mapped cache in shared
//     memory; misses are de-duplicated with __match_any_sync so that one lane
//     per distinct block probes / inserts into the global hash
//     (warp-ballot allocation, one aggregated append to the block list).
// Reference semantics: TSDFVolume.integrate / _update_voxel
// (mast3r_slam/tsdf/global_volume.py:53-88).
#pragma once
#include "common.cuh"
#include "march_fast.cuh"
#include "ray.cuh"

#define FUSE_CACHE_WAYS 32

struct WarpCache { // one per warp, in shared memory
  unsigned long long key[FUSE_CACHE_WAYS];
  uint32_t slot[FUSE_CACHE_WAYS];
};

__device__ __forceinline__ void warp_cache_init(WarpCache &wc) {
  wc.key[threadIdx.x & 31] = B200_EMPTY_KEY;
  wc.slot[threadIdx.x & 31] = 0;
  __syncwarp();
}

__device__ __forceinline__ float ex2_ftz(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// Converged fold: every lane runs this; only lanes with `pred` reduce.
template <bool BOUNDED>
__device__ __forceinline__ void fold_warp(const VolumeDev &v, WarpCache &wc, bool pred, int kx,
                                          int ky, int kz, float su, float wu, uint32_t n,
                                          float conf, float conf_q, uint32_t cp, uint32_t cq,
                                          bool check_bounds, Tally &tl) {
  bool ok = pred;
  if (BOUNDED && check_bounds) {
    bool in = (uint32_t)(kx - v.kmin[0]) < (uint32_t)(v.kmax[0] - v.kmin[0]) &&
              (uint32_t)(ky - v.kmin[1]) < (uint32_t)(v.kmax[1] - v.kmin[1]) &&
              (uint32_t)(kz - v.kmin[2]) < (uint32_t)(v.kmax[2] - v.kmin[2]);
    if (ok && !in) tl.dropped += n * (cp + cq);
    ok = ok && in;
  }
  const int bx = kx >> 3, by = ky >> 3, bz = kz >> 3;
  const unsigned long long key = pack_block_key(bx, by, bz);
  const int h = (bx + 3 * by + 5 * bz) & (FUSE_CACHE_WAYS - 1);
  uint32_t slot = wc.slot[h];
  bool miss = ok && (wc.key[h] != key);
  unsigned mm = __ballot_sync(0xFFFFFFFFu, miss);
  if (mm) { // rare once the warp's blocks are cached
    if (miss) {
      unsigned peers = __match_any_sync(mm, key);
      int leader = __ffs(peers) - 1;
      bool fresh = false;
      if ((int)(threadIdx.x & 31) == leader) slot = hash_insert(v, key, &fresh);
      // warp-aggregated append of the freshly allocated blocks
      unsigned fm = __ballot_sync(mm, fresh);
      if (fm) {
        int first = __ffs(fm) - 1;
        unsigned long long base = 0;
        if ((int)(threadIdx.x & 31) == first)
          base = atomicAdd(v.ctr + CTR_BLOCKS, (unsigned long long)__popc(fm));
        base = __shfl_sync(mm, base, first);
        if (fresh) v.list[base + __popc(fm & ((1u << (threadIdx.x & 31)) - 1))] = slot;
      }
      slot = __shfl_sync(peers, slot, leader);
      unsigned same_way = __match_any_sync(mm, h); // one writer per cache way
      if (slot == 0xFFFFFFFFu) {
        tl.overflow += 1;
        ok = false;
      } else if ((int)(threadIdx.x & 31) == __ffs(same_way) - 1) {
        wc.key[h] = key;
        wc.slot[h] = slot;
      }
    }
    __syncwarp();
  }
  const uint32_t li = ((kz & 7) << 6) | ((ky & 7) << 3) | (kx & 7);
  const size_t off = (size_t)slot * B200_BLOCK_VOX + li;
  if (ok) {
    red_add_v4(v.vox + off, conf * su, conf * wu, conf_q * su, conf_q * wu);
    red_add_u64(v.cnt + off, (unsigned long long)(n * cp) | ((unsigned long long)(n * cq) << 32));
  }
}

// March the 32 rays of a warp in lockstep.  `have` = this lane holds a ray
// (already counted in tl.rays by the caller).
template <bool BOUNDED>
__device__ __forceinline__ void march_warp(const VolumeDev &v, WarpCache &wc, const MarchConst &mc,
                                           bool have, const Ray<float> &r, float conf,
                                           float conf_q, bool use_q, Tally &tl) {
  const uint32_t cp = (have && conf > 0.0f) ? 1u : 0u; // :75  weight <= 0 -> return
  const uint32_t cq = (have && use_q && conf_q > 0.0f) ? 1u : 0u;
  const bool live = (cp | cq) != 0;
  if (!cp) conf = 0.0f;
  if (!cq) conf_q = 0.0f;
  // |s / vs| bound along the ray -> integer-proximity threshold of the fast
  // floor; rays whose keys could leave the +-2^23 voxel range of the block hash
  // are dropped whole.  Bounded volumes: a ray whose band lies inside the
  // bounds takes the launch-wide threshold and skips the per-voxel bound test.
  float lim;
  bool far = false, check_bounds = false;
  bool inside = false;
  if (BOUNDED) {
    float d0 = fmaxf(r.len - mc.trunc, 0.0f);
    float ax0 = fmaf(d0, r.dx, r.ox) * mc.inv_vs, ax1 = fmaf(r.maxd, r.dx, r.ox) * mc.inv_vs;
    float ay0 = fmaf(d0, r.dy, r.oy) * mc.inv_vs, ay1 = fmaf(r.maxd, r.dy, r.oy) * mc.inv_vs;
    float az0 = fmaf(d0, r.dz, r.oz) * mc.inv_vs, az1 = fmaf(r.maxd, r.dz, r.oz) * mc.inv_vs;
    inside = fminf(ax0, ax1) >= mc.blo[0] && fmaxf(ax0, ax1) < mc.bhi[0] &&
             fminf(ay0, ay1) >= mc.blo[1] && fmaxf(ay0, ay1) < mc.bhi[1] &&
             fminf(az0, az1) >= mc.blo[2] && fmaxf(az0, az1) < mc.bhi[2];
    check_bounds = !__all_sync(0xFFFFFFFFu, inside || !live);
  }
  if (BOUNDED && inside) {
    lim = mc.lim_bounded;
  } else {
    float ex = fabsf(r.ox) + r.maxd * fabsf(r.dx), ey = fabsf(r.oy) + r.maxd * fabsf(r.dy),
          ez = fabsf(r.oz) + r.maxd * fabsf(r.dz);
    float qmax = fmaxf(ex, fmaxf(ey, ez)) * mc.inv_vs * 1.0001f + 2.0f;
    lim = (qmax < 4.0e6f) ? 0.5f - qmax * 2.3841858e-7f : -1.0f; // 2^-22
    far = !(qmax < 8.0e6f);
  }
  // first in-band sample: i_lo is conservative (ray.cuh), advance it exactly
  int i = r.i_lo;
  if (live && !far) {
#pragma unroll 1
    while (i < r.n - 1 && fabsf(__fsub_rn(r.len, __fmul_rn((float)i, r.lstep))) > mc.trunc) ++i;
  }
  const int cnt = (live && !far) ? (r.n - i) : 0;
  const int J = __reduce_max_sync(0xFFFFFFFFu, cnt);
  if (live && far) tl.dropped += 1;
  const float MAGIC = 12582912.0f; // 1.5 * 2^23
  const float last_dist = (r.n == 1) ? 0.0f : r.maxd; // np.linspace endpoint / n == 1
  int ax = 0, ay = 0, az = 0;
  float su = 0.f, wu = 0.f;
  uint32_t n = 0, total = 0;
  float fi = (float)i;
  for (int j = 0; j < J; ++j, ++i, fi += 1.0f) {
    float dist = (i == r.n - 1) ? last_dist : __fmul_rn(fi, r.lstep);
    float sdf = __fsub_rn(r.len, dist);
    const bool inb = (j < cnt) && !(fabsf(sdf) > mc.trunc); // :66-67
    float sx = __fadd_rn(r.ox, __fmul_rn(dist, r.dx));
    float sy = __fadd_rn(r.oy, __fmul_rn(dist, r.dy));
    float sz = __fadd_rn(r.oz, __fmul_rn(dist, r.dz));
    float qx = __fmaf_rn(sx, mc.inv_vs, -0.5f), qy = __fmaf_rn(sy, mc.inv_vs, -0.5f),
          qz = __fmaf_rn(sz, mc.inv_vs, -0.5f);
    float yx = __fadd_rn(qx, MAGIC), yy = __fadd_rn(qy, MAGIC), yz = __fadd_rn(qz, MAGIC);
    int kx = __float_as_int(yx) - 0x4B400000, ky = __float_as_int(yy) - 0x4B400000,
        kz = __float_as_int(yz) - 0x4B400000;
    float ux = __fsub_rn(qx, __fsub_rn(yx, MAGIC)), uy = __fsub_rn(qy, __fsub_rn(yy, MAGIC)),
          uz = __fsub_rn(qz, __fsub_rn(yz, MAGIC));
    float m = fmaxf(fabsf(ux), fmaxf(fabsf(uy), fabsf(uz)));
    const bool slow = inb && !(m < lim);
    if (__any_sync(0xFFFFFFFFu, slow)) {
      if (slow) { // the reference operation itself (:133-134)
        float fx = floorf(__fdiv_rn(sx, mc.vs)), fy = floorf(__fdiv_rn(sy, mc.vs)),
              fz = floorf(__fdiv_rn(sz, mc.vs));
        kx = (int)fx, ky = (int)fy, kz = (int)fz;
      }
    }
    const bool chg = inb && n != 0 && (kx != ax || ky != ay || kz != az);
    if (__any_sync(0xFFFFFFFFu, chg)) {
      fold_warp<BOUNDED>(v, wc, chg, ax, ay, az, su, wu, n, conf, conf_q, cp, cq, check_bounds,
                         tl);
      if (chg) {
        su = wu = 0.f;
        n = 0;
      }
    }
    if (inb) {
      ax = kx, ay = ky, az = kz;
      float e = ex2_ftz(fabsf(sdf) * mc.nl2e_inv_trunc); // :69 exp(-|sdf| / trunc)
      su = fmaf(sdf * mc.inv_trunc, e, su);               // :68 (|t| <= 1 inside the band)
      wu += e;
      n += 1;
      total += 1;
    }
  }
  const bool fin = n != 0;
  if (__any_sync(0xFFFFFFFFu, fin))
    fold_warp<BOUNDED>(v, wc, fin, ax, ay, az, su, wu, n, conf, conf_q, cp, cq, check_bounds, tl);
  tl.ev_p += total * cp;
  tl.ev_q += total * cq;
}
