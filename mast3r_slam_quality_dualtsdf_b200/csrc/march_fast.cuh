// march_fast.cuh -- constants and tallies of the float32 ray march.
//
// How each sample's key floor(fl(s / vs)) is obtained without a division:
//   fast path  qm = fma(s, 1/vs, -0.5); kf = rint(qm) via the 1.5*2^23 magic
//              constant; u = qm - kf.  With Q = s/vs:  Q = kf + 0.5 + u + err,
//              |err| <= |Q| * 2^-23, and |fl(Q) - Q| <= |Q| * 2^-24, so if
//              |u| < 0.5 - thr, thr = 2^-22 * (|Q|max + 1), then fl(Q) lies
//              strictly inside (kf, kf + 1) and floor(fl(Q)) == kf.  No
//              division, no MUFU, no F2I/I2F.
//   slow path  (any axis within thr of an integer, ~1e-4 of the samples) the
//              reference operation itself: floorf(__fdiv_rn(s, vs)).
// The sample positions s = o + dist * d are computed exactly as in ray.cuh.
// tsdf and weight use a reciprocal multiply and ex2.approx: relative error
// < 4e-7, far inside the 1e-5 parity tolerance (global_volume.py:68-69).
#pragma once
#include "common.cuh"
#include "ray.cuh"

struct MarchConst {
  float trunc, vs, inv_vs, inv_trunc, nl2e_inv_trunc; // -log2(e) / trunc
  // bounded volumes: float key bounds shrunk by 2 voxels, launch-wide floor threshold
  float blo[3], bhi[3], lim_bounded;
  // quality grid cell size in pixels (H / gh, W / gw) and shifts when a power of two (-1: divide)
  int cell_h, cell_w, cell_h_shift, cell_w_shift;
  int use_dir; // bounded volume with a block directory
  int negate;  // 1: de-integration (sums and hit counters are subtracted)
};

struct Tally {
  uint32_t rays, ev_p, ev_q, drop_p, drop_q, overflow, culled;
};

// ev_* arrive as ENUMERATED in-band samples; applied = enumerated - dropped.
// Every vote / reduction is skipped when no lane has the quantity, so a warp
// whose rays were all culled pays for two reductions and two atomics only.
__device__ __forceinline__ void commit_tally(const VolumeDev &v, Tally tl) {
  const uint32_t lost = tl.drop_p | tl.drop_q | tl.overflow;
  const bool any_lost = __any_sync(0xFFFFFFFFu, lost != 0);
  const bool any_culled = __any_sync(0xFFFFFFFFu, tl.culled != 0);
  const bool any_ev = __any_sync(0xFFFFFFFFu, (tl.ev_p | tl.ev_q) != 0);
  tl.rays = __reduce_add_sync(0xFFFFFFFFu, tl.rays);
  // nothing to report (warp-uniform); a warp can hold updates of rays it did not
  // count itself once rays are re-dealt between warps
  if (tl.rays == 0 && !any_lost && !any_culled && !any_ev) return;
  // applied = enumerated - dropped; signed, because the fast path commits the
  // enumerated samples before the march and the drops after it
  const bool any_app = any_ev || any_lost;
  const int app_p = any_app ? __reduce_add_sync(0xFFFFFFFFu, (int)tl.ev_p - (int)tl.drop_p) : 0;
  const int app_q = any_app ? __reduce_add_sync(0xFFFFFFFFu, (int)tl.ev_q - (int)tl.drop_q) : 0;
  uint32_t dropped = 0, overflow = 0;
  if (any_lost) { // out-of-bounds / foreign-slab drops, hash overflow
    dropped = __reduce_add_sync(0xFFFFFFFFu, tl.drop_p + tl.drop_q);
    overflow = __reduce_add_sync(0xFFFFFFFFu, tl.overflow);
  }
  tl.culled = any_culled ? __reduce_add_sync(0xFFFFFFFFu, tl.culled) : 0u;
  const bool any_rare = any_lost || any_culled;
#ifdef FUSE_DBG_NO_TALLY
  if (tl.rays == 0xFFFFFFFFu)
#else
  if ((threadIdx.x & 31) == 0)
#endif
  {
    unsigned long long *ctr = ctr_replica(v);
    if (tl.rays) atomicAdd(ctr + CTR_RAYS, (unsigned long long)tl.rays);
    if (app_p) atomicAdd(ctr + CTR_EV_P, (unsigned long long)(long long)app_p);
    if (app_q) atomicAdd(ctr + CTR_EV_Q, (unsigned long long)(long long)app_q);
    if (any_rare) {
      if (dropped) atomicAdd(ctr + CTR_DROPPED, (unsigned long long)dropped);
      if (overflow) atomicAdd(ctr + CTR_OVERFLOW, (unsigned long long)overflow);
      if (tl.culled) atomicAdd(ctr + CTR_CULLED, (unsigned long long)tl.culled);
    }
  }
}
