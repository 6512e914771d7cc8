// march_fast.cuh -- float32 ray march with the instruction diet.
//
// Same results as march_ray<float> (integrate.cu): voxel keys bit-exact,
// tsdf / weight within the 1e-5 tolerance.  What changes is HOW each sample's
// key floor(fl(s / vs)) is obtained:
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
  int dbg; // experiments only (B200TSDF_DEBUG): bit0 skip count red, bit1 skip voxel red
};

struct Tally {
  uint32_t rays, ev_p, ev_q, dropped, overflow;
};

struct VoxCache {
  int kx, ky, kz; // any voxel key inside the cached block
  uint32_t slot;
  bool valid;
};

// Fold one voxel's partial sums into the packed voxel: one 128-bit vector
// reduction + one 64-bit reduction (hit counters).
template <bool BOUNDED>
__device__ __forceinline__ void flush_fast(const VolumeDev &v, int kx, int ky, int kz, float su,
                                           float wu, uint32_t n, float conf, float conf_q,
                                           uint32_t cp, uint32_t cq, VoxCache &bc, Tally &tl,
                                           int dbg = 0) {
  if (BOUNDED) {
    bool in = (uint32_t)(kx - v.kmin[0]) < (uint32_t)(v.kmax[0] - v.kmin[0]) &&
              (uint32_t)(ky - v.kmin[1]) < (uint32_t)(v.kmax[1] - v.kmin[1]) &&
              (uint32_t)(kz - v.kmin[2]) < (uint32_t)(v.kmax[2] - v.kmin[2]);
    if (!in) {
      tl.dropped += n * (cp + cq);
      return;
    }
  }
  if (!bc.valid || ((((kx ^ bc.kx) | (ky ^ bc.ky) | (kz ^ bc.kz)) & ~7) != 0)) {
    int bx = kx >> 3, by = ky >> 3, bz = kz >> 3;
    if (!block_in_range(bx, by, bz)) {
      tl.dropped += n * (cp + cq);
      return;
    }
    bool fresh;
    uint32_t s = hash_insert(v, pack_block_key(bx, by, bz), &fresh);
    if (s == 0xFFFFFFFFu) {
      tl.overflow += 1;
      return;
    }
    if (fresh) {
      unsigned long long idx = atomicAdd(v.ctr + CTR_BLOCKS, 1ull);
      v.list[idx] = s;
    }
    bc.kx = kx, bc.ky = ky, bc.kz = kz, bc.slot = s, bc.valid = true;
  }
  uint32_t li = ((kz & 7) << 6) | ((ky & 7) << 3) | (kx & 7);
  size_t off = (size_t)bc.slot * B200_BLOCK_VOX + li;
  if (!(dbg & 2)) red_add_v4(v.vox + off, conf * su, conf * wu, conf_q * su, conf_q * wu);
  if (!(dbg & 1)) red_add_u64(v.cnt + off, (unsigned long long)(n * cp) | ((unsigned long long)(n * cq) << 32));
}

// global_volume.py:63-70 for one ray; conf / conf_q are the reference `conf`
// of the primary / quality volume (conf_q ignored when !use_q).
template <bool BOUNDED>
__device__ __forceinline__ void march_ray_fast(const VolumeDev &v, const Ray<float> &r,
                                               const MarchConst &mc, float conf, float conf_q,
                                               bool use_q, Tally &tl) {
  const uint32_t cp = conf > 0.0f ? 1u : 0u;            // :75  weight <= 0 -> return
  const uint32_t cq = (use_q && conf_q > 0.0f) ? 1u : 0u;
  if ((cp | cq) == 0) return;
  if (!cp) conf = 0.0f;
  if (!cq) conf_q = 0.0f;
  // bound on |s / vs| along the ray -> integer-proximity threshold
  float ex = fabsf(r.ox) + r.maxd * fabsf(r.dx), ey = fabsf(r.oy) + r.maxd * fabsf(r.dy),
        ez = fabsf(r.oz) + r.maxd * fabsf(r.dz);
  float qmax = fmaxf(ex, fmaxf(ey, ez)) * mc.inv_vs * 1.0001f + 2.0f;
  float lim = (qmax < 4.0e6f) ? 0.5f - qmax * 2.3841858e-7f : -1.0f; // 2^-22
  const float MAGIC = 12582912.0f; // 1.5 * 2^23
  const float last_dist = (r.n == 1) ? 0.0f : r.maxd; // np.linspace endpoint / n == 1
  VoxCache bc;
  bc.valid = false;
  bc.kx = bc.ky = bc.kz = 0, bc.slot = 0;
  int ax = 0, ay = 0, az = 0;
  float su = 0.f, wu = 0.f;
  uint32_t n = 0, total = 0;
  float fi = (float)r.i_lo;
  for (int i = r.i_lo; i < r.n; ++i, fi += 1.0f) {
    float dist = (i == r.n - 1) ? last_dist : __fmul_rn(fi, r.lstep);
    float sdf = __fsub_rn(r.len, dist);
    if (fabsf(sdf) > mc.trunc) continue; // :66-67
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
    if (!(m < lim)) { // the reference operation itself (:133-134)
      float fx = floorf(__fdiv_rn(sx, mc.vs)), fy = floorf(__fdiv_rn(sy, mc.vs)),
            fz = floorf(__fdiv_rn(sz, mc.vs));
      fx = fminf(fmaxf(fx, -2.0e9f), 2.0e9f);
      fy = fminf(fmaxf(fy, -2.0e9f), 2.0e9f);
      fz = fminf(fmaxf(fz, -2.0e9f), 2.0e9f);
      kx = (int)fx, ky = (int)fy, kz = (int)fz;
    }
    if (n && (kx != ax || ky != ay || kz != az)) {
      flush_fast<BOUNDED>(v, ax, ay, az, su, wu, n, conf, conf_q, cp, cq, bc, tl, mc.dbg);
      su = wu = 0.f;
      n = 0;
    }
    ax = kx, ay = ky, az = kz;
    float t = sdf * mc.inv_trunc;                       // :68 (|t| <= 1 inside the band)
    t = fminf(fmaxf(t, -1.0f), 1.0f);
    float e = exp2f(fabsf(sdf) * mc.nl2e_inv_trunc);    // :69 exp(-|sdf| / trunc)
    su = fmaf(t, e, su);
    wu += e;
    n += 1;
    total += 1;
  }
  if (n) flush_fast<BOUNDED>(v, ax, ay, az, su, wu, n, conf, conf_q, cp, cq, bc, tl, mc.dbg);
  tl.ev_p += total * cp;
  tl.ev_q += total * cq;
}

__device__ __forceinline__ void commit_tally(const VolumeDev &v, Tally tl) {
  tl.rays = __reduce_add_sync(0xFFFFFFFFu, tl.rays);
  tl.ev_p = __reduce_add_sync(0xFFFFFFFFu, tl.ev_p);
  tl.ev_q = __reduce_add_sync(0xFFFFFFFFu, tl.ev_q);
  tl.dropped = __reduce_add_sync(0xFFFFFFFFu, tl.dropped);
  tl.overflow = __reduce_add_sync(0xFFFFFFFFu, tl.overflow);
  if ((threadIdx.x & 31) == 0) {
    if (tl.rays) atomicAdd(v.ctr + CTR_RAYS, (unsigned long long)tl.rays);
    if (tl.ev_p) atomicAdd(v.ctr + CTR_EV_P, (unsigned long long)tl.ev_p);
    if (tl.ev_q) atomicAdd(v.ctr + CTR_EV_Q, (unsigned long long)tl.ev_q);
    if (tl.dropped) atomicAdd(v.ctr + CTR_DROPPED, (unsigned long long)tl.dropped);
    if (tl.overflow) atomicAdd(v.ctr + CTR_OVERFLOW, (unsigned long long)tl.overflow);
  }
}
