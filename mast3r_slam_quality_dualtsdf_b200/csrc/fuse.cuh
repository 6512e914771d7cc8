// fuse.cuh -- warp-converged dual-volume ray march (float32), v3.
//
// Same arithmetic as march_fast.cuh (exact voxel keys, tolerance-level tsdf /
// weight), structured so that the 32 rays of a warp run ONE converged
// instruction stream:
//   * the in-band sample range [i0, i1] of every ray is found ONCE with the
//     reference's exact test (abs(sdf) > trunc, global_volume.py:66-67; the
//     sdf sequence is monotone, so the band is contiguous) -- the loop itself
//     carries no band test and no event counting;
//   * warp-uniform trip count (longest band of the warp), finished lanes are
//     predicated off;
//   * the fold of a finished voxel (address, ONE 128-bit vector reduction +
//     one 64-bit counter reduction) is executed by the whole warp under a
//     predicate whenever any lane leaves a voxel;
//   * voxel-block slots: bounded volumes read a dense block->slot directory
//     (an L1/L2-resident cache of the block hash); unbounded volumes use a
//     per-warp direct-mapped cache in shared memory.  Misses are
//     de-duplicated with __match_any_sync so that one lane per distinct block
//     probes / inserts into the global hash (warp-ballot allocation, one
//     warp-aggregated append to the block list);
//   * multi-GPU: rays whose band misses this rank's z-slabs are culled before
//     the march; voxels of foreign slabs are dropped at the fold.
// Reference semantics: TSDFVolume.integrate / _update_voxel
// (mast3r_slam/tsdf/global_volume.py:53-88).
#pragma once
#include "common.cuh"
#include "march_fast.cuh"
#include "ray.cuh"

#define FUSE_CACHE_WAYS 32
#ifndef FUSE_SLOT_REUSE
#define FUSE_SLOT_REUSE 1
#endif
#ifndef FUSE_EX2_EARLY
#define FUSE_EX2_EARLY 1
#endif
#ifndef FUSE_MINB
#define FUSE_MINB 4
#endif
#ifndef FUSE_SEAM_TRIM
#define FUSE_SEAM_TRIM 1
#endif

struct WarpCache { // one per warp, in shared memory (unbounded volumes)
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

// One lane per distinct missing block inserts into the hash; fresh slots are
// appended to the block list with one aggregated atomicAdd.  `mm` = ballot of
// the lanes that call this (all of them hold a valid `key`).
__device__ __forceinline__ uint32_t alloc_blocks_warp(const VolumeDev &v, unsigned mm,
                                                      unsigned long long key) {
  const int lane = threadIdx.x & 31;
  unsigned peers = __match_any_sync(mm, key);
  int leader = __ffs(peers) - 1;
  bool fresh = false;
  uint32_t slot = B200_NO_SLOT;
  if (lane == leader) slot = hash_insert(v, key, &fresh);
  unsigned fm = __ballot_sync(mm, fresh);
  if (fm) {
    int first = __ffs(fm) - 1;
    unsigned long long base = 0;
    if (lane == first) base = atomicAdd(v.ctr + CTR_BLOCKS, (unsigned long long)__popc(fm));
    base = __shfl_sync(mm, base, first);
    if (fresh) v.list[base + __popc(fm & ((1u << lane) - 1))] = slot;
  }
  return __shfl_sync(peers, slot, leader);
}

// Converged fold: every lane runs this; only lanes with `pred` reduce.
// DIR: bounded volume with a block directory.
template <bool BOUNDED, bool DIR>
__device__ __forceinline__ void fold_warp(const VolumeDev &v, WarpCache &wc, bool pred, int kx,
                                          int ky, int kz, float su, float wu, uint32_t n,
                                          float conf, float conf_q, float inv_trunc, uint32_t cp,
                                          uint32_t cq, uint32_t mlo, uint32_t mhi,
                                          bool check_bounds, Tally &tl, int &cdi,
                                          uint32_t &cslot) {
  bool ok = pred;
  if (BOUNDED && check_bounds) {
    bool in = (uint32_t)(kx - v.kmin[0]) < (uint32_t)(v.kmax[0] - v.kmin[0]) &&
              (uint32_t)(ky - v.kmin[1]) < (uint32_t)(v.kmax[1] - v.kmin[1]) &&
              (uint32_t)(kz - v.kmin[2]) < (uint32_t)(v.kmax[2] - v.kmin[2]);
    if (v.slab_count > 1) in = in && slab_owns_z(v, kz);
    if (ok && !in) {
      tl.drop_p += n * cp;
      tl.drop_q += n * cq;
    }
    ok = ok && in;
  }
  const int bx = kx >> 3, by = ky >> 3, bz = kz >> 3;
  uint32_t slot;
  if (DIR) {
    const int di = ((bz - v.dir_b0[2]) * v.dir_n[1] + (by - v.dir_b0[1])) * v.dir_n[0] +
                   (bx - v.dir_b0[0]);
#if FUSE_SLOT_REUSE
    // consecutive voxels of a ray mostly share a block: re-read the directory
    // only when the lane's block changed
    slot = (ok && di != cdi) ? v.dir[di] : cslot;
    if (ok) cdi = di;
#else
    slot = ok ? v.dir[di] : 0u;
#endif
    const bool miss = ok && slot == B200_NO_SLOT;
    const unsigned mm = __ballot_sync(0xFFFFFFFFu, miss);
    if (mm) { // only while the batch reaches fresh blocks
      if (miss) {
        slot = alloc_blocks_warp(v, mm, pack_block_key(bx, by, bz));
        if (slot == B200_NO_SLOT) {
          tl.overflow += 1;
          ok = false;
        } else {
          v.dir[di] = slot; // every writer stores the same value
        }
      }
    }
#if FUSE_SLOT_REUSE
    if (ok) cslot = slot;
    else if (di == cdi) cdi = -1;
#endif
  } else {
    const unsigned long long key = pack_block_key(bx, by, bz);
    const int h = (bx + 3 * by + 5 * bz) & (FUSE_CACHE_WAYS - 1);
    slot = wc.slot[h];
    const bool miss = ok && (wc.key[h] != key);
    const unsigned mm = __ballot_sync(0xFFFFFFFFu, miss);
    if (mm) { // rare once the warp's blocks are cached
      if (miss) {
        slot = alloc_blocks_warp(v, mm, key);
        unsigned same_way = __match_any_sync(mm, h); // one writer per cache way
        if (slot == B200_NO_SLOT) {
          tl.overflow += 1;
          ok = false;
        } else if ((int)(threadIdx.x & 31) == __ffs(same_way) - 1) {
          wc.key[h] = key;
          wc.slot[h] = slot;
        }
      }
      __syncwarp();
    }
  }
  if (ok) {
    const uint32_t li = ((kz & 7) << 6) | ((ky & 7) << 3) | (kx & 7);
    const size_t off = (size_t)slot * B200_BLOCK_VOX + li;
#ifndef FUSE_DBG_NO_VOXRED
    const float st = su * inv_trunc; // su = sum of sdf * e; t = sdf / trunc (:68)
    red_add_v4(v.vox + off, conf * st, conf * wu, conf_q * st, conf_q * wu);
#endif
#ifndef FUSE_DBG_NO_CNTRED
    // n hits for each volume that takes the ray: n * (cp + 2^32 cq) mod 2^64, with
    // the multiplier negated for de-integration (an exact 64-bit subtract)
    const unsigned long long hits =
        (unsigned long long)n * mlo + ((unsigned long long)(n * mhi) << 32);
    red_add_u64(v.cnt + off, hits);
#endif
  }
}

// March the 32 rays of a warp in lockstep.  `have` = this lane holds a ray
// (already counted in tl.rays by the caller).
template <bool BOUNDED, bool DIR>
__device__ __forceinline__ void march_warp(const VolumeDev &v, WarpCache &wc, const MarchConst &mc,
                                           bool have, const Ray<float> &r, float conf,
                                           float conf_q, bool use_q, Tally &tl) {
  uint32_t cp = (have && conf > 0.0f) ? 1u : 0u; // :75  weight <= 0 -> return
  uint32_t cq = (have && use_q && conf_q > 0.0f) ? 1u : 0u;
  bool live = (cp | cq) != 0;
  // |s / vs| bound along the ray -> integer-proximity threshold of the fast
  // floor; rays whose keys could leave the +-2^23 voxel range of the block hash
  // are skipped whole.  Bounded volumes: a ray whose band lies inside the
  // bounds (and inside one owned slab) takes the launch-wide threshold and
  // skips the per-voxel bound test.
  float lim;
  bool far = false, check_bounds = false, inside = false;
  // multi-GPU seam trim: z range [zt_lo, zt_hi) (voxel layers) that holds every
  // slab of this rank the band can reach; samples clearly outside it are not
  // marched (the per-voxel ownership test at the fold stays the exact one)
  bool trim = false;
  float zt_lo = 0.f, zt_hi = 0.f;
  if (BOUNDED) {
    float d0 = fmaxf(r.len - mc.trunc, 0.0f);
    float ax0 = fmaf(d0, r.dx, r.ox) * mc.inv_vs, ax1 = fmaf(r.maxd, r.dx, r.ox) * mc.inv_vs;
    float ay0 = fmaf(d0, r.dy, r.oy) * mc.inv_vs, ay1 = fmaf(r.maxd, r.dy, r.oy) * mc.inv_vs;
    float az0 = fmaf(d0, r.dz, r.oz) * mc.inv_vs, az1 = fmaf(r.maxd, r.dz, r.oz) * mc.inv_vs;
    inside = fminf(ax0, ax1) >= mc.blo[0] && fmaxf(ax0, ax1) < mc.bhi[0] &&
             fminf(ay0, ay1) >= mc.blo[1] && fmaxf(ay0, ay1) < mc.bhi[1] &&
             fminf(az0, az1) >= mc.blo[2] && fmaxf(az0, az1) < mc.bhi[2];
    if (v.slab_count > 1 && live) {
      float zl = fminf(az0, az1) - 2.0f, zh = fmaxf(az0, az1) + 2.0f;
      bool any = true, all = false;
      if (fabsf(zl) < 1.0e9f && fabsf(zh) < 1.0e9f) { // else: keep, per-voxel test decides
        int s_lo = ((int)floorf(zl) - v.slab_z0) >> v.slab_shift;
        int s_hi = ((int)floorf(zh) - v.slab_z0) >> v.slab_shift;
        int need = v.slab_index - slab_owner_of(v, s_lo); // slabs to my first one at/after s_lo
        if (need < 0) need += v.slab_count;
        any = need <= s_hi - s_lo;
        all = (s_hi == s_lo) && need == 0;
#if FUSE_SEAM_TRIM
        if (any && !all && fabsf(zl) < 4.0e6f && fabsf(zh) < 4.0e6f) {
          const int s_own = s_lo + need, span = s_hi - s_own; // first / last owned slab in reach
          const int s_last = s_own + (v.slab_mask >= 0 ? (span & ~v.slab_mask)
                                                       : span - span % v.slab_count);
          const float wz = (float)(1 << v.slab_shift);
          zt_lo = fmaf((float)s_own, wz, (float)v.slab_z0);
          zt_hi = fmaf((float)(s_last + 1), wz, (float)v.slab_z0);
          trim = true;
        }
#endif
      }
      if (!any) { // the band misses every slab of this rank
        live = false;
        tl.culled += 1;
      }
      inside = inside && all;
    }
    check_bounds = !__all_sync(0xFFFFFFFFu, inside || !live);
  }
  if (!live) cp = cq = 0u;
  if (!cp) conf = 0.0f;
  if (!cq) conf_q = 0.0f;
  uint32_t mlo = cp, mhi = cq; // hit-counter multiplier (see fold_warp)
  if (mc.negate != 0) { // de-integration: subtract what the same call added
    conf = -conf, conf_q = -conf_q;
    const unsigned long long m = 0ull - ((unsigned long long)cp | ((unsigned long long)cq << 32));
    mlo = (uint32_t)m, mhi = (uint32_t)(m >> 32);
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
  if (live && far) tl.culled += 1;
  // exact in-band range [i0, i1]: sdf_i = fl(len - dist_i) is non-increasing in
  // i, so {i : !(abs(sdf_i) > trunc)} (global_volume.py:66-67) is contiguous.
  const float last_dist = (r.n == 1) ? 0.0f : r.maxd; // np.linspace endpoint / n == 1
  int i0 = r.i_lo, cnt = 0;
  if (live && !far) {
    int i1 = r.n - 1;
#pragma unroll 1
    while (i0 < i1 && __fsub_rn(r.len, __fmul_rn((float)i0, r.lstep)) > mc.trunc) ++i0;
    const float s_last = __fsub_rn(r.len, last_dist);
    if (s_last < -mc.trunc) { // the endpoint itself fell out of the band by rounding
      --i1;
#pragma unroll 1
      while (i1 > i0 && __fsub_rn(r.len, __fmul_rn((float)i1, r.lstep)) < -mc.trunc) --i1;
    }
    if (i1 >= i0) {
      float sa = __fsub_rn(r.len, (i0 == r.n - 1) ? last_dist : __fmul_rn((float)i0, r.lstep));
      float sb = __fsub_rn(r.len, (i1 == r.n - 1) ? last_dist : __fmul_rn((float)i1, r.lstep));
      if (!(fabsf(sa) > mc.trunc) && !(fabsf(sb) > mc.trunc)) cnt = i1 - i0 + 1;
    }
#if FUSE_SEAM_TRIM
    if (BOUNDED && trim && cnt > 0) {
      // sample i sits at z ~ zq0 + i * g voxel layers (the exact recipe differs by
      // rounding only: < 1e-3 layers).  With |g| >= 4e-3 that is < 0.3 samples, and
      // the range keeps one more guard sample on either side; flatter rays are not
      // trimmed.  Everything cut here lies in a foreign slab (or off the grid), so
      // applied = ev - drop is unchanged.
      const float g = r.dz * r.lstep * mc.inv_vs, zq0 = r.oz * mc.inv_vs;
      if (fabsf(g) >= 4.0e-3f) {
        const float ig = 1.0f / g;
        const float ia = (zt_lo - zq0) * ig, ib = (zt_hi - zq0) * ig;
        const float flo = fmaxf(fminf(ia, ib), -1.0e9f), fhi = fminf(fmaxf(ia, ib), 1.0e9f);
        const int j0 = max(i0, (int)floorf(flo) - 1), j1 = min(i1, (int)ceilf(fhi) + 1);
        i0 = j0;
        cnt = j1 >= j0 ? j1 - j0 + 1 : 0;
      }
    }
#endif
  }
  const int J = __reduce_max_sync(0xFFFFFFFFu, cnt);
  const float MAGIC = 12582912.0f; // 1.5 * 2^23
  const int i_end = r.n - 1;
  int ax = 0, ay = 0, az = 0;
  float su = 0.f, wu = 0.f;
  uint32_t n = 0;
  int i = i0, cdi = -1;
  uint32_t cslot = 0;
  float fi = (float)i0;
  for (int j = 0; j < J; ++j, ++i, fi += 1.0f) {
    const bool act = j < cnt;
    float dist = (i == i_end) ? last_dist : __fmul_rn(fi, r.lstep);
    float sdf = __fsub_rn(r.len, dist);
#if FUSE_EX2_EARLY
    const float e = ex2_ftz(fabsf(sdf) * mc.nl2e_inv_trunc); // :69 exp(-|sdf| / trunc)
#endif
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
    const bool slow = act && !(m < lim);
    if (__any_sync(0xFFFFFFFFu, slow)) {
      if (slow) { // the reference operation itself (:133-134)
        float fx = floorf(__fdiv_rn(sx, mc.vs)), fy = floorf(__fdiv_rn(sy, mc.vs)),
              fz = floorf(__fdiv_rn(sz, mc.vs));
        kx = (int)fx, ky = (int)fy, kz = (int)fz;
      }
    }
    const bool chg = act && n != 0 && (((kx ^ ax) | (ky ^ ay) | (kz ^ az)) != 0);
    if (__any_sync(0xFFFFFFFFu, chg)) {
      fold_warp<BOUNDED, DIR>(v, wc, chg, ax, ay, az, su, wu, n, conf, conf_q, mc.inv_trunc, cp, cq,
                              mlo, mhi, check_bounds, tl, cdi, cslot);
      if (chg) {
        su = wu = 0.f;
        n = 0;
      }
    }
    if (act) {
      ax = kx, ay = ky, az = kz;
#if !FUSE_EX2_EARLY
      const float e = ex2_ftz(fabsf(sdf) * mc.nl2e_inv_trunc); // :69 exp(-|sdf| / trunc)
#endif
      su = fmaf(sdf, e, su); // :68 t * w with t = sdf / trunc (|t| <= 1 inside the band)
      wu += e;
      n += 1;
    }
  }
  const bool fin = n != 0;
  if (__any_sync(0xFFFFFFFFu, fin))
    fold_warp<BOUNDED, DIR>(v, wc, fin, ax, ay, az, su, wu, n, conf, conf_q, mc.inv_trunc, cp, cq, mlo,
                            mhi, check_bounds, tl, cdi, cslot);
  tl.ev_p += (uint32_t)cnt * cp;
  tl.ev_q += (uint32_t)cnt * cq;
}
