// integrate.cu -- dual-volume ray-march fusion (direct-to-global variant).
//
// Replaces TSDFVolume.integrate + _update_voxel
// (mast3r_slam/tsdf/global_volume.py:35-88) and, for keyframes, the staging
// of TSDFGlobalIntegrator._integrate_single (tsdf/global_manager.py:81-106).
// One thread marches one ray over the samples inside the +-trunc band,
// pre-combines consecutive samples that land in the same voxel in registers,
// and folds each voxel's partial sums {dS_p, dW_p, dS_q, dW_q} into the packed
// voxel with ONE 128-bit vector reduction (red.global.add.v4.f32) plus one
// 64-bit reduction for the two hit counters.  The fold is commutative
// (DESIGN.md "the fold"), so no ordering between rays is needed.
#include <cooperative_groups.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "ray.cuh"
#include "march_fast.cuh"
#include "fuse.cuh"
#include "fuse_fast.cuh"
#include "sim3.cuh"

struct Acc {
  int kx, ky, kz;
  float sp, wp, sq, wq;
  uint32_t np, nq;
  bool live;
};

struct BlockCache {
  unsigned long long key;
  uint32_t slot;
};

__device__ __forceinline__ void flush_voxel(const VolumeDev &v, const Acc &a, BlockCache &bc,
                                            uint32_t &ev_p, uint32_t &ev_q, uint32_t &dropped,
                                            uint32_t &overflow) {
  if (!a.live) return;
  const bool out =
      (v.bounded && (a.kx < v.kmin[0] || a.kx >= v.kmax[0] || a.ky < v.kmin[1] ||
                     a.ky >= v.kmax[1] || a.kz < v.kmin[2] || a.kz >= v.kmax[2])) ||
      !slab_owns_z(v, a.kz);
  int bx = a.kx >> B200_BLOCK_LOG2, by = a.ky >> B200_BLOCK_LOG2, bz = a.kz >> B200_BLOCK_LOG2;
  if (out || !block_in_range(bx, by, bz)) { // counters hold APPLIED updates
    ev_p -= a.np, ev_q -= a.nq;
    dropped += a.np + a.nq;
    return;
  }
  unsigned long long key = pack_block_key(bx, by, bz);
  if (key != bc.key) {
    bool fresh;
    uint32_t s = hash_insert(v, key, &fresh);
    if (s == 0xFFFFFFFFu) {
      overflow += 1;
      return;
    }
    if (fresh) {
      unsigned long long idx = atomicAdd(v.ctr + CTR_BLOCKS, 1ull);
      v.list[idx] = s;
    }
    bc.key = key;
    bc.slot = s;
  }
  uint32_t li = ((a.kz & 7) << 6) | ((a.ky & 7) << 3) | (a.kx & 7);
  size_t off = (size_t)bc.slot * B200_BLOCK_VOX + li;
  red_add_v4(v.vox + off, a.sp, a.wp, a.sq, a.wq);
  red_add_u64(v.cnt + off, (unsigned long long)a.np | ((unsigned long long)a.nq << 32));
}

// March one ray.  conf / conf_q are the reference `conf` of the two volumes;
// use_q = 0 leaves the quality volume untouched.
template <typename T>
__device__ __forceinline__ void march_ray(const VolumeDev &v, const Ray<T> &r, T trunc, T vs,
                                          float conf, float conf_q, bool use_q, uint32_t &ev_p,
                                          uint32_t &ev_q, uint32_t &dropped,
                                          uint32_t &overflow) {
  const bool on_p = conf > 0.0f;            // global_volume.py:75 weight <= 0 -> return
  const bool on_q = use_q && conf_q > 0.0f; // same test on the C*q volume
  if (!on_p && !on_q) return;
  Acc acc;
  acc.live = false;
  acc.kx = acc.ky = acc.kz = 0;
  acc.sp = acc.wp = acc.sq = acc.wq = 0.f;
  acc.np = acc.nq = 0;
  BlockCache bc;
  bc.key = B200_EMPTY_KEY;
  bc.slot = 0;
  for (int i = r.i_lo; i < r.n; ++i) {
    int kx, ky, kz;
    T t, a;
    if (!ray_sample(r, i, trunc, vs, kx, ky, kz, t, a)) continue;
    float e = expf((float)a); // global_volume.py:69 (f64 there; 1e-5 tolerance)
    float tf = (float)t;
    if (!acc.live || kx != acc.kx || ky != acc.ky || kz != acc.kz) {
      flush_voxel(v, acc, bc, ev_p, ev_q, dropped, overflow);
      acc.live = true;
      acc.kx = kx, acc.ky = ky, acc.kz = kz;
      acc.sp = acc.wp = acc.sq = acc.wq = 0.f;
      acc.np = acc.nq = 0;
    }
    if (on_p) {
      float w = conf * e;
      acc.sp += tf * w;
      acc.wp += w;
      acc.np += 1;
      ev_p += 1;
    }
    if (on_q) {
      float w = conf_q * e;
      acc.sq += tf * w;
      acc.wq += w;
      acc.nq += 1;
      ev_q += 1;
    }
  }
  flush_voxel(v, acc, bc, ev_p, ev_q, dropped, overflow);
}

__device__ __forceinline__ void commit_counters(const VolumeDev &v, uint32_t rays, uint32_t ev_p,
                                                uint32_t ev_q, uint32_t dropped,
                                                uint32_t overflow) {
  rays = __reduce_add_sync(0xFFFFFFFFu, rays);
  ev_p = __reduce_add_sync(0xFFFFFFFFu, ev_p);
  ev_q = __reduce_add_sync(0xFFFFFFFFu, ev_q);
  dropped = __reduce_add_sync(0xFFFFFFFFu, dropped);
  overflow = __reduce_add_sync(0xFFFFFFFFu, overflow);
  if ((threadIdx.x & 31) == 0) {
    unsigned long long *ctr = ctr_replica(v);
    if (rays) atomicAdd(ctr + CTR_RAYS, (unsigned long long)rays);
    if (ev_p) atomicAdd(ctr + CTR_EV_P, (unsigned long long)ev_p);
    if (ev_q) atomicAdd(ctr + CTR_EV_Q, (unsigned long long)ev_q);
    if (dropped) atomicAdd(ctr + CTR_DROPPED, (unsigned long long)dropped);
    if (overflow) atomicAdd(ctr + CTR_OVERFLOW, (unsigned long long)overflow);
  }
}

// ------------------------------------------------------------ points entry
// float64 inputs: the generic march (every intermediate float64, as numpy
// does for float64 arrays); float32 inputs: the fast march.
__global__ void __launch_bounds__(256)
    integrate_points_f64_kernel(VolumeDev v, const double *__restrict__ pts,
                                const double *__restrict__ conf,
                                const double *__restrict__ conf_q, double ox, double oy,
                                double oz, long long m, double step, double trunc, double vs) {
  long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t rays = 0, ev_p = 0, ev_q = 0, dropped = 0, overflow = 0;
  if (p < m) {
    Ray<double> r;
    if (ray_setup<double>(r, pts[3 * p], pts[3 * p + 1], pts[3 * p + 2], ox, oy, oz, step,
                          trunc)) {
      rays = 1; // global_volume.py:71
      float c = (float)conf[p];
      float cq = conf_q ? (float)conf_q[p] : 0.f;
      march_ray<double>(v, r, trunc, vs, c, cq, conf_q != nullptr, ev_p, ev_q, dropped,
                        overflow);
    }
  }
  commit_counters(v, rays, ev_p, ev_q, dropped, overflow);
}

template <bool BOUNDED, bool DIR>
__global__ void __launch_bounds__(256)
    integrate_points_f32_kernel(VolumeDev v, const float *__restrict__ pts,
                                const double *__restrict__ conf,
                                const double *__restrict__ conf_q, float ox, float oy, float oz,
                                long long m, float step, MarchConst mc) {
  __shared__ WarpCache wcs[DIR ? 1 : 8];
  WarpCache &wc = wcs[DIR ? 0 : (threadIdx.x >> 5)];
  if (!DIR) warp_cache_init(wc);
  long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  Tally tl = {0, 0, 0, 0, 0, 0, 0};
  Ray<float> r = {};
  bool have = false;
  float c = 0.f, cq = 0.f;
  if (p < m) {
    have = ray_setup<float>(r, pts[3 * p], pts[3 * p + 1], pts[3 * p + 2], ox, oy, oz, step,
                            mc.trunc);
    if (have) {
      tl.rays = 1; // global_volume.py:71
      c = (float)conf[p];
      cq = conf_q ? (float)conf_q[p] : 0.f;
    }
  }
  march_warp<BOUNDED, DIR>(v, wc, mc, have, r, c, cq, conf_q != nullptr, tl);
  commit_tally(v, tl);
}

// ---------------------------------------------------------- keyframe entry
// ---- keyframe staging: TMA bulk copies (cp.async.bulk -> UBLKCP) of the CTA's
// 32 x 8 pixel tile into shared memory, completion on an mbarrier.
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes,
                                         uint32_t bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
          "r"(dst),
      "l"(src), "r"(bytes), "r"(bar)
      : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t phase) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra WAIT_DONE;\n"
      "bra WAIT_LOOP;\n"
      "WAIT_DONE:\n"
      "}\n" ::"r"(bar),
      "r"(phase)
      : "memory");
}

// multi-GPU: the band's z extent lies within trunc (+3 voxels) of the surface
// point; true when no slab of this rank is within that reach (conservative).
__device__ __forceinline__ bool slab_zmiss_vox(const VolumeDev &v, float zc, float ext) {
  if (!(fabsf(zc) < 1.0e9f)) return false; // the per-voxel test decides
  const int s_lo = ((int)floorf(zc - ext) - v.slab_z0) >> v.slab_shift;
  const int s_hi = ((int)floorf(zc + ext) - v.slab_z0) >> v.slab_shift;
  int need = v.slab_index - slab_owner_of(v, s_lo);
  if (need < 0) need += v.slab_count;
  return need > s_hi - s_lo;
}
// The in-band samples of a ray lie within trunc * |dz| of the surface point in
// z (dz = z component of the unit direction); + 2.5 voxels for the floor and
// rounding.  d2 = squared ray length, az = its z component.
__device__ __forceinline__ float band_reach_vox(const MarchConst &mc, float az, float d2) {
  float dz = fabsf(az) * rsqrtf(d2);
  if (!(dz < 1.0f)) dz = 1.0f; // also d2 = 0 / NaN
  return fmaf(mc.trunc * mc.inv_vs, dz, 2.5f);
}
// A ray the reference would fuse (global_volume.py:56: finite, len >= 1e-4)?
// Certain when 1e-3 < len < 3e18; otherwise the exact ray_length decides.
__device__ __forceinline__ bool surely_a_ray(float d2) { return d2 > 1.0e-6f && d2 < 1.0e37f; }

// multi-GPU, per ray: does the band's z range [zl, zh] (voxel layers, already
// widened by 2) reach a slab of this rank (`any`), lie in ONE owned slab (`all`),
// and which z range holds every owned slab in reach (seam trim)?
__device__ __forceinline__ void slab_band(const VolumeDev &v, float zl, float zh, bool &any,
                                          bool &all, bool &trim, float &zt_lo, float &zt_hi) {
  any = true, all = false, trim = false;
  if (fabsf(zl) < 1.0e9f && fabsf(zh) < 1.0e9f) { // else: keep, per-voxel test decides
    const int s_lo = ((int)floorf(zl) - v.slab_z0) >> v.slab_shift;
    const int s_hi = ((int)floorf(zh) - v.slab_z0) >> v.slab_shift;
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
}

// Cold leg of the fast kernels: a warp in which some ray does not qualify for
// the fast path (band near the bounds, operands out of range) re-does the
// generic set-up of ray.cuh and runs march_warp.  Out of line, so that the fast
// march is compiled without its registers; commits its own tally.
template <bool BOUNDED, bool DIR>
__device__ __noinline__ void march_generic_cold(const VolumeDev *vp, const MarchConst *mcp,
                                                const float *P, bool have, float wx, float wy,
                                                float wz, float c, float cq, float step,
                                                uint32_t rays, uint32_t culled) {
  __shared__ WarpCache wc_cold[DIR ? 1 : 8];
  Ray<float> r = {};
  if (have) {
    float rx, ry, rz;
    have = ray_length<float>(r, wx, wy, wz, P[0], P[1], P[2], rx, ry, rz);
    if (have) ray_finish<float>(r, rx, ry, rz, P[0], P[1], P[2], step, mcp->trunc);
  }
  Tally tl = {rays, 0, 0, 0, 0, 0, culled};
  march_warp<BOUNDED, DIR>(*vp, wc_cold[DIR ? 0 : (threadIdx.x >> 5)], *mcp, have, r, c, cq, true,
                           tl);
  commit_tally(*vp, tl);
}

// One pixel per lane: Sim3 act, slab cull, ray set-up, warp-converged march.
// `take` = the pixel passed the confidence filter (global_manager.py:92).
// The order keeps a rank's cost for rays of OTHER ranks minimal: world point,
// conservative z test, and only for survivors the exact length (sqrt, float64
// sum), the divisions of the ray set-up and the quality-grid lookup.
template <bool BOUNDED, bool DIR>
__device__ __forceinline__ void fuse_pixels_warp(const VolumeDev &v, WarpCache &wc,
                                                 const MarchConst &mc, const float *P, bool take,
                                                 float c, float x0, float x1, float x2, float step,
                                                 const float *__restrict__ qg, long long q_base,
                                                 int gh, int gw, int px, int py, Tally &tl) {
  Ray<float> r = {};
  bool have = false;
  if (take) {
    float wx, wy, wz;
    sim3_act(P, x0, x1, x2, wx, wy, wz);
    const float ax = wx - P[0], ay = wy - P[1], az = wz - P[2];
    const float d2 = fmaf(ax, ax, fmaf(ay, ay, az * az));
    const bool zmiss = BOUNDED && v.slab_count > 1 &&
                       slab_zmiss_vox(v, wz * mc.inv_vs, band_reach_vox(mc, az, d2));
    const bool counted = zmiss && surely_a_ray(d2);
    if (counted) {
      tl.rays += 1;
      tl.culled += 1;
    } else {
      float rx, ry, rz;
      have = ray_length<float>(r, wx, wy, wz, P[0], P[1], P[2], rx, ry, rz);
      if (have) {
        tl.rays += 1;
        if (zmiss) {
          have = false;
          tl.culled += 1;
        } else {
          ray_finish<float>(r, rx, ry, rz, P[0], P[1], P[2], step, mc.trunc);
        }
      }
    }
  }
  // a warp whose rays were all filtered / culled skips the march altogether
  if (__any_sync(0xFFFFFFFFu, have)) {
    float cq = c;
    if (qg) { // nearest-neighbour upsample of the per-patch quality weight
      int gy = mc.cell_h_shift >= 0 ? py >> mc.cell_h_shift : py / mc.cell_h;
      int gx = mc.cell_w_shift >= 0 ? px >> mc.cell_w_shift : px / mc.cell_w;
      gy = min(gy, gh - 1), gx = min(gx, gw - 1);
      cq = c * __ldg(qg + q_base + (long long)gy * gw + gx);
    }
    march_warp<BOUNDED, DIR>(v, wc, mc, have, r, c, cq, true, tl);
  }
}

// The same for a bounded volume with a block directory: the warp takes the fast
// path (fuse_fast.cuh) when every ray of it qualifies, else march_generic_cold.
// Commits the warp's tally itself.
__device__ __forceinline__ void fuse_pixels_warp_fast(const VolumeDev &v, const MarchConst &mc,
                                                      const FastConst &fc, const float *P,
                                                      bool take, float c, float x0, float x1,
                                                      float x2, float step,
                                                      const float *__restrict__ qg, int gh, int gw,
                                                      int px, int py, float *d_cta) {
  FastRay fr = {};
  bool have = false, fast_ok = true;
  float wx = 0.f, wy = 0.f, wz = 0.f;
  Tally tl = {0, 0, 0, 0, 0, 0, 0};
  if (take) {
    const float ox = P[0], oy = P[1], oz = P[2];
    sim3_act(P, x0, x1, x2, wx, wy, wz);
    const float ax = wx - ox, ay = wy - oy, az = wz - oz;
    const float d2 = fmaf(ax, ax, fmaf(ay, ay, az * az));
    const bool zmiss = v.slab_count > 1 &&
                       slab_zmiss_vox(v, wz * mc.inv_vs, band_reach_vox(mc, az, d2));
    const bool counted = zmiss && surely_a_ray(d2);
    if (counted) {
      tl.rays += 1;
      tl.culled += 1;
    } else {
      const int st = fast_ray_setup(fr, wx, wy, wz, ox, oy, oz, step, rcp_newton(step), mc.trunc,
                                    fc.band_k);
      if (st != 0) {
        tl.rays += 1;
        if (zmiss) {
          tl.culled += 1;
        } else {
          have = true;
          // qualifies: operands in range and the whole band inside the shrunk bounds
          const float ux = fmaf(wx, fc.in_sc[0], fc.in_of[0]), uy = fmaf(wy, fc.in_sc[1], fc.in_of[1]),
                      uz = fmaf(wz, fc.in_sc[2], fc.in_of[2]);
          fast_ok = st == 1 && fmaxf(fabsf(ux), fmaxf(fabsf(uy), fabsf(uz))) < 1.0f;
        }
      }
    }
  }
  // a warp whose rays were all filtered / culled skips the march altogether
  if (!__any_sync(0xFFFFFFFFu, have)) {
    commit_tally(v, tl);
    return;
  }
  float cq = c;
  if (qg) { // nearest-neighbour upsample of the per-patch quality weight
    int gy = mc.cell_h_shift >= 0 ? py >> mc.cell_h_shift : py / mc.cell_h;
    int gx = mc.cell_w_shift >= 0 ? px >> mc.cell_w_shift : px / mc.cell_w;
    gy = min(gy, gh - 1), gx = min(gx, gw - 1);
    cq = c * __ldg(qg + (long long)gy * gw + gx);
  }
  if (__builtin_expect(!__all_sync(0xFFFFFFFFu, !have || fast_ok), 0)) {
    march_generic_cold<true, true>(&v, &mc, P, have, wx, wy, wz, c, cq, step, tl.rays, tl.culled);
    return;
  }
  uint32_t cp = (have && c > 0.0f) ? 1u : 0u; // :75  weight <= 0 -> return
  uint32_t cqf = (have && cq > 0.0f) ? 1u : 0u;
  bool live = (cp | cqf) != 0;
  int i0 = 0, cnt = 0;
  bool slab_check = false;
  if (v.slab_count > 1) {
    bool any = true, all = true, trim = false;
    float zt_lo = 0.f, zt_hi = 0.f;
    const float oz = P[2];
    if (live) {
      const float d0 = fmaxf(fr.len - mc.trunc, 0.0f);
      const float az0 = fmaf(d0, fr.dz, oz) * mc.inv_vs, az1 = fmaf(fr.maxd, fr.dz, oz) * mc.inv_vs;
      slab_band(v, fminf(az0, az1) - 2.0f, fmaxf(az0, az1) + 2.0f, any, all, trim, zt_lo, zt_hi);
      if (!any) { // the band misses every slab of this rank
        live = false;
        tl.culled += 1;
      }
    }
    slab_check = !__all_sync(0xFFFFFFFFu, all || !live);
    if (live) {
      cnt = fast_band(fr, mc.trunc, i0);
#if FUSE_SEAM_TRIM
      if (trim && cnt > 0) { // see march_warp: everything cut lies in a foreign slab
        const float g = fr.dz * fr.lstep * mc.inv_vs, zq0 = oz * mc.inv_vs;
        if (fabsf(g) >= 4.0e-3f) {
          const float ig = 1.0f / g;
          const float ia = (zt_lo - zq0) * ig, ib = (zt_hi - zq0) * ig;
          const float flo = fmaxf(fminf(ia, ib), -1.0e9f), fhi = fminf(fmaxf(ia, ib), 1.0e9f);
          const int i1 = i0 + cnt - 1;
          const int j0 = max(i0, (int)floorf(flo) - 1), j1 = min(i1, (int)ceilf(fhi) + 1);
          i0 = j0;
          cnt = j1 >= j0 ? j1 - j0 + 1 : 0;
        }
      }
#endif
    }
    if (!live) cp = cqf = 0u, cnt = 0;
  } else {
    if (live) cnt = fast_band(fr, mc.trunc, i0);
  }
  // rays and enumerated samples are final here: commit them now so that the
  // march carries no tally registers (its drops / overflows are committed after)
  tl.ev_p += (uint32_t)cnt * cp;
  tl.ev_q += (uint32_t)cnt * cqf;
  commit_tally(v, tl);
  const uint32_t lost = march_warp_fast(v, mc, fc, fr, P, i0, cnt, cp ? c : 0.0f,
                                        cqf ? cq : 0.0f, slab_check, d_cta);
  if (__any_sync(0xFFFFFFFFu, lost != 0)) { // drops at slab seams, hash overflow
    Tally t2 = {0, 0, 0, lost & 0xFFFu, (lost >> 12) & 0xFFFu, lost >> 24, 0};
    commit_tally(v, t2);
  }
}

#define TILE_W 32
#define TILE_H 8
#define SX_ROW 104 // floats per staged X row: 96 + 8 pad (bank-conflict-free, 16 B multiple)
#define SC_ROW 40  // floats per staged C row: 32 + 8 pad

#ifndef FUSE_KF_ORDER
#define FUSE_KF_ORDER 1
#endif
#ifndef FUSE_TMA_PAR
#define FUSE_TMA_PAR 0 // 1: the tile's row copies issued by 16 lanes at once (A/B)
#endif
struct KfArgs { // one keyframe batch (device pointers)
  const float *X, *C, *qg, *poses; // bases of segment 0
  // packed multi-GPU ingest: the batch is n segments (one per source rank) of
  // seg_kf keyframes each, seg_stride floats apart (0: one contiguous batch)
  long long seg_stride;
  int seg_kf;
  int H, W, gh, gw;
  float min_conf, step;
  // tsdf_global.max_points_per_kf: per keyframe {hash threshold, salt}; a valid pixel
  // is fused iff sub_hash(pixel, salt) <= threshold (nullptr: every valid pixel)
  const uint2 *sub;
  // peer pull (multi-GPU): pointmaps / confidences of segment s live at segX[s] /
  // segC[s] -- the OWNER rank's HBM, mapped into this process (NVLink) -- instead of
  // X + s * seg_stride; quality grids and poses come from the gathered meta records
  // through qg / poses + s * seg_stride as before.  n_ptr = 0: not used.
  const float *segX[B200TSDF_MAX_SEGMENTS], *segC[B200TSDF_MAX_SEGMENTS];
  int n_ptr;
  // CTA order of the one-CTA-per-tile launch: `order` keyframes of the SAME tile
  // position are adjacent in the block schedule (grid.x = order * tiles_x, grid.z =
  // keyframes / order).  Consecutive keyframes of a trajectory see nearly the same
  // surface through the same tile, so the voxels the resident CTAs reduce into are a
  // few MB instead of a whole view's band and stay in L2 between their updates.
  int order;
};

// Seeded bijective 32-bit hash of the pixel index (murmur3 finaliser: xor-shifts
// and odd multiplications are bijections of Z_2^32, so two pixels of a keyframe
// never tie and "the max_points smallest hashes" is an exact-size choice).
__host__ __device__ __forceinline__ uint32_t sub_fmix(uint32_t x) {
  x ^= x >> 16, x *= 0x85ebca6bu, x ^= x >> 13, x *= 0xc2b2ae35u, x ^= x >> 16;
  return x;
}
__host__ __device__ __forceinline__ uint32_t sub_hash(uint32_t pix, uint32_t salt) {
  return sub_fmix(pix * 0x9E3779B1u + salt);
}
__host__ __device__ __forceinline__ uint32_t sub_salt(uint64_t seed, uint64_t kf_id) {
  return sub_fmix((uint32_t)seed ^ sub_fmix((uint32_t)kf_id + 0x9E3779B9u * (uint32_t)(seed >> 32)) ^
                  sub_fmix((uint32_t)(kf_id >> 32) + 0x7F4A7C15u));
}
__device__ __forceinline__ bool sub_take(const KfArgs &a, int kf, int pix) {
  if (!a.sub) return true;
  const uint2 ts = __ldg(a.sub + kf);
  return sub_hash((uint32_t)pix, ts.y) <= ts.x;
}

struct KfPtrs {
  const float *X, *C, *qg, *P;
};
// Pointers of keyframe kf (pointmap, confidence, quality grid, pose).
__device__ __forceinline__ KfPtrs kf_ptrs(const KfArgs &a, int kf) {
  long long so = 0;
  int loc = kf;
  if (a.seg_stride != 0) {
    const int seg = kf / a.seg_kf;
    loc = kf - seg * a.seg_kf;
    so = (long long)seg * a.seg_stride;
  }
  const long long npix = (long long)a.H * a.W;
  KfPtrs p;
  if (a.n_ptr > 0) {
    const int seg = kf / a.seg_kf;
    p.X = a.segX[seg] + (long long)loc * npix * 3;
    p.C = a.segC[seg] + (long long)loc * npix;
  } else {
    p.X = a.X + so + (long long)loc * npix * 3;
    p.C = a.C + so + (long long)loc * npix;
  }
  p.qg = a.qg ? a.qg + so + (long long)loc * a.gh * a.gw : nullptr;
  p.P = a.poses + so + (long long)loc * 8;
  return p;
}

// thread 0: TMA bulk copies of tile (bx, by) of keyframe kf into shared memory
__device__ __forceinline__ void tile_issue(const KfArgs &a, float *sX, float *sC, uint32_t bar,
                                           int bx, int by, int kf) {
  mbar_expect_tx(bar, TILE_H * (TILE_W * 12 + TILE_W * 4));
  const KfPtrs k = kf_ptrs(a, kf);
  const long long p0 = (long long)(by * TILE_H) * a.W + bx * TILE_W;
#pragma unroll
  for (int row = 0; row < TILE_H; ++row) {
    bulk_g2s(smem_u32(sX + row * SX_ROW), k.X + (p0 + (long long)row * a.W) * 3, TILE_W * 12, bar);
    bulk_g2s(smem_u32(sC + row * SC_ROW), k.C + p0 + (long long)row * a.W, TILE_W * 4, bar);
  }
}

// CTA = 32 x 8 pixel tile, warp = 8 x 4 sub-tile: rays of a warp stay close in
// the volume, so their reductions share 32-byte sectors.
template <bool BOUNDED, bool DIR, bool TMA>
__device__ __forceinline__ void tile_consume(const VolumeDev &v, WarpCache &wc, const KfArgs &a,
                                             const MarchConst &mc, const FastConst &fc,
                                             float *d_sh, const float *sX,
                                             const float *sC, uint32_t bar, uint32_t phase,
                                             int bx, int by, int kf, Tally &tl) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int tx = (warp & 3) * 8 + (lane & 7), ty = (warp >> 2) * 4 + (lane >> 3);
  int px = bx * TILE_W + tx, py = by * TILE_H + ty;
  float c = 0.f, x0 = 0.f, x1 = 0.f, x2 = 0.f;
  bool in_img = true;
  const KfPtrs k = kf_ptrs(a, kf);
  if (TMA) { // host guarantees full tiles and 16-byte aligned rows
    mbar_wait(bar, phase);
    c = sC[ty * SC_ROW + tx];
    x0 = sX[ty * SX_ROW + tx * 3], x1 = sX[ty * SX_ROW + tx * 3 + 1],
    x2 = sX[ty * SX_ROW + tx * 3 + 2];
  } else {
    in_img = px < a.W && py < a.H;
    if (in_img) {
      const long long pix = (long long)py * a.W + px;
      c = k.C[pix];
      if (c > a.min_conf) {
        const float *x = k.X + pix * 3;
        x0 = x[0], x1 = x[1], x2 = x[2];
      }
    }
  }
  bool take = in_img && c > a.min_conf && sub_take(a, kf, py * a.W + px);
  if (a.sub) {
    // tsdf_global.max_points_per_kf is active (launch-uniform): about max_points / (H W)
    // of the tile's pixels survive (20 % at the reference's defaults), which would leave
    // every warp marching with mostly idle lanes.  Compact the survivors onto the first
    // warps of the CTA (order: warp sub-tile, then lane -- neighbours stay neighbours); the
    // other warps find no ray and leave.  Which rays are fused does not change, and the
    // fold is commutative, so the volume is the same up to float32 summation order.
    __shared__ unsigned char s_sel[TILE_W * TILE_H];
    __shared__ int s_cnt[8];
    const unsigned bal = __ballot_sync(0xFFFFFFFFu, take);
    if (lane == 0) s_cnt[warp] = __popc(bal);
    __syncthreads();
    int base = 0, total = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) {
      const int cw = s_cnt[w];
      base += w < warp ? cw : 0;
      total += cw;
    }
    if (take) s_sel[base + __popc(bal & ((1u << lane) - 1u))] = (unsigned char)(ty * TILE_W + tx);
    __syncthreads(); // (the next tile's s_cnt writes come after the caller's loop barrier)
    take = (int)threadIdx.x < total;
    if (take) {
      const int t = s_sel[threadIdx.x];
      tx = t & (TILE_W - 1), ty = t >> 5;
      px = bx * TILE_W + tx, py = by * TILE_H + ty;
      if (TMA) {
        c = sC[ty * SC_ROW + tx];
        x0 = sX[ty * SX_ROW + tx * 3], x1 = sX[ty * SX_ROW + tx * 3 + 1],
        x2 = sX[ty * SX_ROW + tx * 3 + 2];
      } else {
        const long long pix = (long long)py * a.W + px;
        c = k.C[pix];
        const float *x = k.X + pix * 3;
        x0 = x[0], x1 = x[1], x2 = x[2];
      }
    }
  }
  if (BOUNDED && DIR && fc.enabled) { // commits its own tally
    fuse_pixels_warp_fast(v, mc, fc, k.P, take, c, x0, x1, x2, a.step, k.qg, a.gh, a.gw, px, py,
                          d_sh);
  } else {
    fuse_pixels_warp<BOUNDED, DIR>(v, wc, mc, k.P, take, c, x0, x1, x2, a.step, k.qg, 0, a.gh,
                                   a.gw, px, py, tl);
  }
}

// One CTA per tile (grid = tiles_x, tiles_y, keyframes): the single-GPU path.
template <bool BOUNDED, bool DIR, bool TMA>
__global__ void __launch_bounds__(256, FUSE_MINB)
    integrate_keyframes_kernel(const __grid_constant__ VolumeDev v, const __grid_constant__ KfArgs a,
                               const __grid_constant__ MarchConst mc,
                               const __grid_constant__ FastConst fc) {
  __shared__ WarpCache wcs[DIR ? 1 : 8];
  __shared__ float d_shs[(BOUNDED && DIR) ? 8 * 96 : 1];
  float *const d_sh = d_shs;
  __shared__ __align__(128) float sX[TMA ? TILE_H * SX_ROW : 4];
  __shared__ __align__(128) float sC[TMA ? TILE_H * SC_ROW : 4];
  __shared__ __align__(8) unsigned long long s_bar;
  WarpCache &wc = wcs[DIR ? 0 : (threadIdx.x >> 5)];
  const uint32_t bar = smem_u32(&s_bar);
  int bx = blockIdx.x, kf = blockIdx.z;
  if (a.order > 1) kf = kf * a.order + bx % a.order, bx /= a.order;
#if FUSE_TMA_PAR
  if (TMA && threadIdx.x < 32) { // A/B: 16 lanes issue one row copy each instead of thread 0 all 16
    const int lane = threadIdx.x;
    if (lane == 0) {
      mbar_init(bar, 1);
      mbar_expect_tx(bar, TILE_H * (TILE_W * 12 + TILE_W * 4));
    }
    __syncwarp();
    if (lane < 2 * TILE_H) {
      const KfPtrs k = kf_ptrs(a, kf);
      const int row = lane >> 1;
      const long long p = (long long)(blockIdx.y * TILE_H + row) * a.W + bx * TILE_W;
      if (lane & 1) bulk_g2s(smem_u32(sC + row * SC_ROW), k.C + p, TILE_W * 4, bar);
      else bulk_g2s(smem_u32(sX + row * SX_ROW), k.X + p * 3, TILE_W * 12, bar);
    }
  }
#else
  if (TMA && threadIdx.x == 0) {
    mbar_init(bar, 1);
    tile_issue(a, sX, sC, bar, bx, blockIdx.y, kf);
  }
#endif
  if (!DIR) warp_cache_init(wc);
  if (TMA) __syncthreads(); // mbarrier init visible to the waiters
  Tally tl = {0, 0, 0, 0, 0, 0, 0};
  tile_consume<BOUNDED, DIR, TMA>(v, wc, a, mc, fc, d_sh, sX, sC, bar, 0, bx, blockIdx.y, kf, tl);
  commit_tally(v, tl);
}

// ---- tsdf_global.max_points_per_kf on one volume: CTA = CAP_G vertically adjacent tiles
// (a 32 x 32 pixel block).  With the reference's default cap (40 000 of 196 608 pixels)
// a 32 x 8 tile keeps ~52 rays: compacted onto the first warps of a one-tile CTA that
// is two marching warps out of eight, and the SM runs latency-bound (0.32 ms for 22 %
// of the uncapped launch's samples).  Four tiles keep ~208 rays = 6.5 full warps: all
// tiles are staged by TMA at once, the survivors of the block are compacted into one
// list (order: tile, warp sub-tile, lane -- neighbours stay neighbours) and the warps
// take 32-ray chunks of it round robin.  Which rays are fused is decided by sub_take
// alone, so the volume equals the one-tile path's up to float32 summation order.
#ifndef CAP_G
#define CAP_G 4 // tiles per CTA: 1, 2 or 4 (the compaction scans CAP_G * 8 <= 32 counts with one warp)
#endif
#ifndef CAP_WARPS
#define CAP_WARPS 4 // warps per CTA (8, 4 or 2).  4: eight CTAs per SM, a warp takes one or two chunks, and a CTA waits less
                    // for its slowest warp (0.254 -> 0.245 ms at cap 40 000, 0.194 -> 0.144 at 10 000; 2 warps: 0.295)
#endif
#define CAP_THREADS (CAP_WARPS * 32)
template <bool BOUNDED, bool DIR, bool TMA>
__global__ void __launch_bounds__(CAP_THREADS, FUSE_MINB * 8 / CAP_WARPS)
    integrate_keyframes_capped_kernel(const __grid_constant__ VolumeDev v,
                                      const __grid_constant__ KfArgs a,
                                      const __grid_constant__ MarchConst mc,
                                      const __grid_constant__ FastConst fc) {
  __shared__ WarpCache wcs[DIR ? 1 : 8];
  __shared__ float d_shs[(BOUNDED && DIR) ? 8 * 96 : 1];
  __shared__ __align__(128) float sX[TMA ? CAP_G * TILE_H * SX_ROW : 4];
  __shared__ __align__(128) float sC[TMA ? CAP_G * TILE_H * SC_ROW : 4];
  __shared__ __align__(8) unsigned long long s_bar;
  __shared__ unsigned short s_sel[CAP_G * TILE_H * TILE_W];
  __shared__ int s_cnt[CAP_G * 8];
  WarpCache &wc = wcs[DIR ? 0 : (threadIdx.x >> 5)];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int bx = blockIdx.x, kf = blockIdx.z;
  const int row0 = blockIdx.y * (CAP_G * TILE_H);
  const int rows = min(CAP_G * TILE_H, a.H - row0); // TMA: a multiple of TILE_H
  const KfPtrs k = kf_ptrs(a, kf);
  const uint32_t bar = smem_u32(&s_bar);
  if (TMA && warp == 0) {
    // the block's 2 x 32 row copies are issued by the 32 lanes of warp 0 at once (one
    // thread issuing them in turn costs more than the copies take); lane 0 arms the
    // barrier first: the phase cannot complete before its arrival
    if (lane == 0) {
      mbar_init(bar, 1);
      mbar_expect_tx(bar, (uint32_t)rows * (TILE_W * 12 + TILE_W * 4));
    }
    __syncwarp();
    if (lane < rows) {
      const long long p = (long long)(row0 + lane) * a.W + bx * TILE_W;
      bulk_g2s(smem_u32(sX + lane * SX_ROW), k.X + p * 3, TILE_W * 12, bar);
      bulk_g2s(smem_u32(sC + lane * SC_ROW), k.C + p, TILE_W * 4, bar);
    }
  }
  if (!DIR) warp_cache_init(wc);
  if (TMA) {
    __syncthreads(); // mbarrier init visible to the waiters
    mbar_wait(bar, 0);
  }
  // which pixels survive the filter and the cap: warp w tests the 8 x 4 sub-tiles
  // w, w + CAP_WARPS, ... of every tile, one pixel per lane
  constexpr int SUBS = 8 / CAP_WARPS; // sub-tiles of a tile per warp
  unsigned bal[CAP_G * SUBS];
#pragma unroll
  for (int t = 0; t < CAP_G; ++t) {
#pragma unroll
    for (int u = 0; u < SUBS; ++u) {
      const int sub = warp + u * CAP_WARPS;
      const int tx = (sub & 3) * 8 + (lane & 7), ty = (sub >> 2) * 4 + (lane >> 3);
      const int px0 = bx * TILE_W + tx;
      const int r = t * TILE_H + ty, py = row0 + r;
      bool take = false;
      if (r < rows && (TMA || px0 < a.W)) {
        const float c = TMA ? sC[r * SC_ROW + tx] : k.C[(long long)py * a.W + px0];
        take = c > a.min_conf && sub_take(a, kf, py * a.W + px0);
      }
      bal[t * SUBS + u] = __ballot_sync(0xFFFFFFFFu, take);
      if (lane == 0) s_cnt[t * 8 + sub] = __popc(bal[t * SUBS + u]);
    }
  }
  __syncthreads();
  // every warp scans the 32 (tile, warp) counts itself
  const int mine = lane < CAP_G * 8 ? s_cnt[lane] : 0;
  int inc = mine;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int n = __shfl_up_sync(0xFFFFFFFFu, inc, d);
    if (lane >= d) inc += n;
  }
  const int total = __shfl_sync(0xFFFFFFFFu, inc, 31);
#pragma unroll
  for (int t = 0; t < CAP_G; ++t) {
#pragma unroll
    for (int u = 0; u < SUBS; ++u) {
      const int sub = warp + u * CAP_WARPS;
      const int tx = (sub & 3) * 8 + (lane & 7), ty = (sub >> 2) * 4 + (lane >> 3);
      const unsigned bl = bal[t * SUBS + u];
      const int base = __shfl_sync(0xFFFFFFFFu, inc - mine, t * 8 + sub);
      if ((bl >> lane) & 1u)
        s_sel[base + __popc(bl & ((1u << lane) - 1u))] =
            (unsigned short)((t * TILE_H + ty) * TILE_W + tx);
    }
  }
  __syncthreads();
  Tally tl = {0, 0, 0, 0, 0, 0, 0};
#pragma unroll 1
  for (int j0 = warp * 32; j0 < total; j0 += CAP_THREADS) { // warp-uniform
    const bool take = j0 + lane < total;
    float c = 0.f, x0 = 0.f, x1 = 0.f, x2 = 0.f;
    int px = 0, py = 0;
    if (take) {
      const int e = s_sel[j0 + lane], r = e >> 5, ex = e & (TILE_W - 1);
      px = bx * TILE_W + ex, py = row0 + r;
      if (TMA) {
        c = sC[r * SC_ROW + ex];
        x0 = sX[r * SX_ROW + ex * 3], x1 = sX[r * SX_ROW + ex * 3 + 1],
        x2 = sX[r * SX_ROW + ex * 3 + 2];
      } else {
        const long long pix = (long long)py * a.W + px;
        c = k.C[pix];
        const float *x = k.X + pix * 3;
        x0 = x[0], x1 = x[1], x2 = x[2];
      }
    }
    if (BOUNDED && DIR && fc.enabled) { // commits its own tally
      fuse_pixels_warp_fast(v, mc, fc, k.P, take, c, x0, x1, x2, a.step, k.qg, a.gh, a.gw, px, py,
                            d_shs);
    } else {
      fuse_pixels_warp<BOUNDED, DIR>(v, wc, mc, k.P, take, c, x0, x1, x2, a.step, k.qg, 0, a.gh,
                                     a.gw, px, py, tl);
    }
  }
  commit_tally(v, tl);
}

// ---- multi-GPU (z-slab ownership): two launches per batch.
// (1) slab_classify_kernel streams the whole batch once at full occupancy
//     (plain coalesced loads, ~30 registers): a tile with at least one ray that
//     can reach a slab of this rank goes on the tile list; the rays of every
//     other tile are counted (rays, culled) here and never touched again.
// (2) integrate_keyframes_listed_kernel: persistent CTAs take listed tiles
//     through a ticket counter and run the same tile path as above.
// A rank's cost is then one streaming pass plus the march of its own tiles,
// instead of a full-weight CTA per foreign tile.  sched[0] = list length,
// sched[1] = ticket cursor (zeroed on the stream before (1)).
#define CLS_TILES 4 // tiles per classify CTA: 4 x 4 independent loads per thread in flight

__global__ void __launch_bounds__(256)
    slab_classify_kernel(VolumeDev v, KfArgs a, MarchConst mc, int *__restrict__ list,
                         unsigned int *sched, int tiles_x, int tiles_y) {
  // CTA = CLS_TILES consecutive tiles of one tile row; thread = one pixel of each
  const int lx = threadIdx.x & 31, py = blockIdx.y * TILE_H + (threadIdx.x >> 5);
  const int kf = blockIdx.z;
  const KfPtrs k = kf_ptrs(a, kf);
  const float *P = k.P;
  const long long row = (long long)py * a.W;
  float c[CLS_TILES], x0[CLS_TILES], x1[CLS_TILES], x2[CLS_TILES];
#pragma unroll
  for (int t = 0; t < CLS_TILES; ++t) { // all loads first (the pass is latency-bound otherwise)
    const int px = (blockIdx.x * CLS_TILES + t) * TILE_W + lx;
    c[t] = 0.f, x0[t] = x1[t] = x2[t] = 0.f;
    if (px < a.W && py < a.H) {
      c[t] = k.C[row + px];
      const float *x = k.X + (row + px) * 3;
      x0[t] = x[0], x1[t] = x[1], x2[t] = x[2];
    }
  }
  // world z in voxels, z = s * (row 2 of the rotation sim3_act applies) . x + t_z.
  // Not the exact operation order, so the reach is widened by a bound on the
  // rounding difference (+1 voxel): "miss" here implies the exact test misses.
  const float q0 = P[3], q1 = P[4], q2 = P[5], q3 = P[6], sv = P[7] * mc.inv_vs;
  const float k0 = 2.0f * (q0 * q2 - q1 * q3) * sv, k1 = 2.0f * (q1 * q2 + q0 * q3) * sv,
              k2 = (1.0f - 2.0f * (q0 * q0 + q1 * q1)) * sv, tzv = P[2] * mc.inv_vs;
  const float s2 = P[7] * P[7], tv = mc.trunc * mc.inv_vs;
  Tally tl = {0, 0, 0, 0, 0, 0, 0};
#pragma unroll
  for (int t = 0; t < CLS_TILES; ++t) {
    const int bx = blockIdx.x * CLS_TILES + t;
    if (bx >= tiles_x) break; // uniform
    const int px = bx * TILE_W + lx;
    const bool take = px < a.W && py < a.H && c[t] > a.min_conf && sub_take(a, kf, py * a.W + px);
    bool zmiss = false;
    if (take) {
      const float zc = fmaf(x0[t], k0, fmaf(x1[t], k1, fmaf(x2[t], k2, tzv)));
      const float mag = fmaf(fabsf(x0[t]), fabsf(k0),
                             fmaf(fabsf(x1[t]), fabsf(k1), fmaf(fabsf(x2[t]), fabsf(k2), fabsf(tzv))));
      // reach as in band_reach_vox, from len = s * |x|, widened by 1% + 1 voxel
      const float d2 = s2 * fmaf(x0[t], x0[t], fmaf(x1[t], x1[t], x2[t] * x2[t]));
      float dz = fmaf(fabsf(zc - tzv) * mc.vs * rsqrtf(d2), 1.01f, 1.0e-3f);
      if (!(dz < 1.0f)) dz = 1.0f;
      zmiss = slab_zmiss_vox(v, zc, fmaf(tv, dz, fmaf(mag, 4.0e-6f, 3.5f)));
    }
    if (__syncthreads_or(take && !zmiss)) { // the fuse pass counts this tile's rays itself
      if (threadIdx.x == 0)
        list[atomicAdd(sched, 1u)] = (kf * tiles_y + blockIdx.y) * tiles_x + bx;
    } else if (take) { // exactly the accounting of fuse_pixels_warp for a culled ray
      // len = s * |x| up to rounding: certainly a ray when 2e-3 < len < 1e18
      const float d2 = s2 * fmaf(x0[t], x0[t], fmaf(x1[t], x1[t], x2[t] * x2[t]));
      bool ray = d2 > 4.0e-6f && d2 < 1.0e36f;
      if (!ray) {
        float wx, wy, wz, rx, ry, rz;
        sim3_act(P, x0[t], x1[t], x2[t], wx, wy, wz);
        Ray<float> r;
        ray = ray_length<float>(r, wx, wy, wz, P[0], P[1], P[2], rx, ry, rz);
      }
      if (ray) tl.rays += 1, tl.culled += 1;
    }
  }
  commit_tally(v, tl);
}

// Persistent CTAs over the tile list, double-buffered: while tile k is marched,
// the TMA copies of tile k+1 are in flight, the list entry of tile k+2 is being
// read and the ticket after that is being fetched (thread 0 keeps them in
// registers and never waits on a value it requested in the same iteration).
#ifndef FUSE_MINB_LISTED
#define FUSE_MINB_LISTED FUSE_MINB
#endif
template <bool BOUNDED, bool DIR, bool TMA>
__global__ void __launch_bounds__(256, FUSE_MINB_LISTED)
    integrate_keyframes_listed_kernel(const __grid_constant__ VolumeDev v,
                                      const __grid_constant__ KfArgs a,
                                      const __grid_constant__ MarchConst mc,
                                      const __grid_constant__ FastConst fc,
                                      const int *__restrict__ list, unsigned int *sched,
                                      int tiles_x, int tiles_y) {
  __shared__ WarpCache wcs[DIR ? 1 : 8];
  __shared__ float d_shs[(BOUNDED && DIR) ? 8 * 96 : 1];
  float *const d_sh = d_shs;
  __shared__ __align__(128) float sX[2][TMA ? TILE_H * SX_ROW : 4];
  __shared__ __align__(128) float sC[2][TMA ? TILE_H * SC_ROW : 4];
  __shared__ __align__(8) unsigned long long s_bar[2];
  __shared__ int s_tile[2];
  WarpCache &wc = wcs[DIR ? 0 : (threadIdx.x >> 5)];
  const int tiles_kf = tiles_x * tiles_y;
  const unsigned int n_list = *(volatile unsigned int *)sched;
  auto lookup = [&](unsigned int t) { return t < n_list ? list[t] : -1; };
  auto issue = [&](int tile, int b) {
    if (TMA && tile >= 0) {
      const int kf = tile / tiles_kf, rem = tile - kf * tiles_kf;
      const int by = rem / tiles_x;
      tile_issue(a, sX[b], sC[b], smem_u32(&s_bar[b]), rem - by * tiles_x, by, kf);
    }
  };
  int nxt = -1;          // thread 0: tile of the next iteration
  unsigned int pend = 0; // thread 0: ticket of the iteration after that
  if (threadIdx.x == 0) {
    if (TMA) mbar_init(smem_u32(&s_bar[0]), 1), mbar_init(smem_u32(&s_bar[1]), 1);
    const int cur = lookup(atomicAdd(sched + 1, 1u));
    issue(cur, 0);
    s_tile[0] = cur;
    nxt = cur >= 0 ? lookup(atomicAdd(sched + 1, 1u)) : -1;
    pend = nxt >= 0 ? atomicAdd(sched + 1, 1u) : 0xFFFFFFFFu;
  }
  if (!DIR) warp_cache_init(wc);
  for (unsigned int k = 0;; ++k) {
    __syncthreads(); // tile k-1 is consumed: its buffers and s_tile slot are free
    const int b = k & 1;
    const int tile = s_tile[b];
    if (tile < 0) break;
    if (threadIdx.x == 0) {
      issue(nxt, b ^ 1);
      s_tile[b ^ 1] = nxt;
      if (nxt >= 0) {
        nxt = lookup(pend);
        pend = atomicAdd(sched + 1, 1u);
      }
    }
    const int kf = tile / tiles_kf, rem = tile - kf * tiles_kf;
    const int by = rem / tiles_x;
    Tally tl = {0, 0, 0, 0, 0, 0, 0};
    tile_consume<BOUNDED, DIR, TMA>(v, wc, a, mc, fc, d_sh, sX[b], sC[b], smem_u32(&s_bar[b]),
                                    (k >> 1) & 1u, rem - by * tiles_x, by, kf, tl);
    commit_tally(v, tl);
  }
}

// ---- tsdf_global.max_points_per_kf (global_manager.py:94-99): a thread-block cluster
// of SUB_SPLIT CTAs per keyframe finds the hash threshold that keeps exactly
// min(n_valid, max_points) valid pixels -- the max_points smallest hashes.
//   pass 1: every CTA histograms the top 11 hash bits of its share of the pixels in
//     its own shared memory; CTA 0 sums the SUB_SPLIT histograms through distributed
//     shared memory, finds the bin of the need-th hash with one block scan and
//     publishes bin, remaining need and the bin's population to the cluster;
//   pass 2: the bin holds n_valid / 2048 hashes (~96 of a 512x384 keyframe): every
//     CTA appends the hashes of its pixels that fall into it to ONE list in CTA 0's
//     shared memory (remote atomicAdd + remote store), and CTA 0 picks the need-th
//     smallest by rank counting -- exact, hashes are distinct;
//   a bin of more than `list_cap` hashes (images of > 2 M valid pixels, or the test
//     knob) takes two more radix passes (11 + 10 bits) instead.
// C stays L2-resident between the passes; the pass is bound by the hash evaluations.
#define SUB_SPLIT 8
#define SUB_THREADS 512 // 40 keyframes x 8 CTAs = 320 CTAs, four per SM: one wave of clusters
#define SUB_LIST_CAP 1024
__global__ void __cluster_dims__(SUB_SPLIT, 1, 1) __launch_bounds__(SUB_THREADS)
    subsample_threshold_kernel(KfArgs a, int max_points, unsigned long long seed,
                               const long long *__restrict__ kf_ids, uint2 *__restrict__ out,
                               int list_cap) {
  namespace cg = cooperative_groups;
  cg::cluster_group cl = cg::this_cluster();
  __shared__ __align__(16) unsigned int hist[2048];
  __shared__ unsigned int s_warp[SUB_THREADS / 32];
  __shared__ unsigned int s_need, s_bin, s_pop; // CTA 0 publishes, the others read them remotely
  __shared__ unsigned int s_list[SUB_LIST_CAP], s_nlist; // CTA 0: the threshold bin's hashes
  const int kf = blockIdx.y, npix = a.H * a.W;
  const unsigned int part = cl.block_rank();
  const int chunk = (npix + SUB_SPLIT - 1) / SUB_SPLIT;
  const int p_lo = (int)part * chunk, p_hi = min(npix, p_lo + chunk);
  const KfPtrs k = kf_ptrs(a, kf);
  const uint32_t salt = sub_salt(seed, kf_ids ? (unsigned long long)kf_ids[kf] : (unsigned long long)kf);
  const unsigned int *const bin0 = cl.map_shared_rank(&s_bin, 0);
  const unsigned int *const need0 = cl.map_shared_rank(&s_need, 0);
  const unsigned int *const pop0 = cl.map_shared_rank(&s_pop, 0);
  unsigned int prefix = 0, need = (unsigned int)max_points; // hashes < prefix region are all kept
  // pass p fixes the next `bits` bits of the threshold below the already fixed ones
  const int shifts[3] = {21, 10, 0}, nbins[3] = {2048, 2048, 1024};
  unsigned int fixed_mask = 0;
  bool all = false, listed = false;
  if (threadIdx.x == 0) s_nlist = 0;
  for (int p = 0; p < 3; ++p) {
    for (int i = threadIdx.x; i < 2048; i += SUB_THREADS) hist[i] = 0;
    __syncthreads();
    for (int i0 = p_lo + (int)threadIdx.x; i0 < p_hi; i0 += 8 * SUB_THREADS) {
      float cv[8]; // eight coalesced loads in flight per thread: a pass is one or two L2 round trips
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int i = i0 + u * SUB_THREADS;
        cv[u] = i < p_hi ? __ldg(k.C + i) : a.min_conf;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        if (cv[u] > a.min_conf) {
          const uint32_t h = sub_hash((uint32_t)(i0 + u * SUB_THREADS), salt);
          if ((h & fixed_mask) == prefix) atomicAdd(&hist[(h >> shifts[p]) & (nbins[p] - 1)], 1u);
        }
      }
    }
    cl.sync(); // every CTA's histogram is complete
    if (part == 0) {
      // thread t owns bins 4t .. 4t + 3 of the cluster's histogram
      unsigned int b0 = 0, b1 = 0, b2 = 0, b3 = 0;
#pragma unroll
      for (int r = 0; r < SUB_SPLIT; ++r) {
        const uint4 hv = *reinterpret_cast<const uint4 *>(cl.map_shared_rank(hist, r) + 4 * threadIdx.x);
        b0 += hv.x, b1 += hv.y, b2 += hv.z, b3 += hv.w;
      }
      if (4 * (int)threadIdx.x >= nbins[p]) b0 = b1 = b2 = b3 = 0;
      // block-wide exclusive scan of the threads' sums
      const unsigned int v = b0 + b1 + b2 + b3;
      unsigned int inc = v;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const unsigned int n = __shfl_up_sync(0xFFFFFFFFu, inc, d);
        if ((int)(threadIdx.x & 31) >= d) inc += n;
      }
      if ((threadIdx.x & 31) == 31) s_warp[threadIdx.x >> 5] = inc;
      if (threadIdx.x == 0) s_bin = (unsigned int)nbins[p]; // fewer than `need` match: keep them all
      __syncthreads();
      unsigned int ex = inc - v;
      for (int w = 0; w < (int)(threadIdx.x >> 5); ++w) ex += s_warp[w];
      // the one thread whose bins hold the need-th smallest hash of the candidates
      if (ex < need && need <= ex + v) {
        unsigned int bin = 4 * threadIdx.x, acc = ex, pop = b0;
        if (acc + b0 < need) {
          acc += b0, ++bin, pop = b1;
          if (acc + b1 < need) {
            acc += b1, ++bin, pop = b2;
            if (acc + b2 < need) acc += b2, ++bin, pop = b3;
          }
        }
        s_bin = bin;
        s_need = need - acc;
        s_pop = pop;
      }
    }
    cl.sync(); // s_bin / s_need / s_pop of CTA 0 are final; nobody reads a remote histogram any more
    const unsigned int bin = *bin0;
    if (bin >= (unsigned int)nbins[p]) {
      all = true;
      break;
    }
    prefix |= bin << shifts[p];
    fixed_mask |= (unsigned int)(nbins[p] - 1) << shifts[p];
    need = *need0;
    if (p == 0 && *pop0 <= (unsigned int)list_cap) { // cluster-uniform
      listed = true;
      break;
    }
    // (CTA 0 rewrites s_bin only after the next pass's first cluster barrier)
  }
  if (listed) {
    unsigned int *const list0 = cl.map_shared_rank(s_list, 0);
    unsigned int *const nlist0 = cl.map_shared_rank(&s_nlist, 0);
    for (int i0 = p_lo + (int)threadIdx.x; i0 < p_hi; i0 += 8 * SUB_THREADS) {
      float cv[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int i = i0 + u * SUB_THREADS;
        cv[u] = i < p_hi ? __ldg(k.C + i) : a.min_conf;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        if (cv[u] > a.min_conf) {
          const uint32_t h = sub_hash((uint32_t)(i0 + u * SUB_THREADS), salt);
          if ((h & fixed_mask) == prefix) { // s_pop entries, unless the caller rewrites C under us
            const unsigned int at = atomicAdd(nlist0, 1u);
            if (at < SUB_LIST_CAP) list0[at] = h;
          }
        }
      }
    }
    cl.sync(); // the list is complete (and nobody touches CTA 0's memory after this)
    if (part == 0) {
      const unsigned int n = min(s_nlist, (unsigned int)SUB_LIST_CAP); // == s_pop
      for (unsigned int i = threadIdx.x; i < n; i += SUB_THREADS) {
        const unsigned int mine = s_list[i];
        unsigned int below = 0;
        for (unsigned int j = 0; j < n; ++j) below += s_list[j] < mine ? 1u : 0u;
        if (below + 1 == need) out[kf] = make_uint2(mine, salt); // exactly one (distinct hashes)
      }
    }
    return;
  }
  // after three passes `prefix` is the hash of the need-th smallest kept pixel
  // (hashes are distinct), so h <= prefix selects exactly max_points pixels
  if (part == 0 && threadIdx.x == 0) out[kf] = make_uint2(all ? 0xFFFFFFFFu : prefix, salt);
  cl.sync(); // CTA 0's shared memory stays alive until every CTA has read it
}

// ------------------------------------------------------------------ C ABI
static int pow2_shift(int x) {
  if (x <= 0 || (x & (x - 1))) return -1;
  int s = 0;
  while ((1 << s) < x) ++s;
  return s;
}

static bool use_bounded(const b200tsdf *h) { return h->d.bounded || h->d.slab_count > 1; }

static MarchConst make_consts(const b200tsdf *h, int cell_h = 1, int cell_w = 1) {
  MarchConst mc;
  mc.trunc = (float)h->cfg.truncation;
  mc.vs = (float)h->cfg.voxel_size;
  mc.inv_vs = 1.0f / mc.vs;
  mc.inv_trunc = 1.0f / mc.trunc;
  mc.nl2e_inv_trunc = -1.4426950408889634f / mc.trunc;
  float qb = 2.0f;
  for (int a = 0; a < 3; ++a) {
    mc.blo[a] = (float)h->d.kmin[a] + 2.0f;
    mc.bhi[a] = (float)h->d.kmax[a] - 2.0f;
    qb = fmaxf(qb, fmaxf(fabsf((float)h->d.kmin[a]), fabsf((float)h->d.kmax[a])) + 2.0f);
  }
  mc.lim_bounded = (qb < 4.0e6f) ? 0.5f - qb * 2.3841858e-7f : -1.0f;
  if (!h->d.bounded) // slabs on an unbounded volume: no ray is ever "inside"
    mc.blo[0] = mc.blo[1] = mc.blo[2] = 3.0e38f;
  mc.cell_h = cell_h > 0 ? cell_h : 1;
  mc.cell_w = cell_w > 0 ? cell_w : 1;
  mc.cell_h_shift = pow2_shift(mc.cell_h);
  mc.cell_w_shift = pow2_shift(mc.cell_w);
  mc.use_dir = h->d.dir_on;
  mc.negate = 0;
  return mc;
}

// Launch constants of the fast path (fuse_fast.cuh); enabled for bounded volumes
// with a block directory whose slot offsets fit 32 bits.
static FastConst make_fast_consts(const b200tsdf *h, double step_d) {
  FastConst fc;
  memset(&fc, 0, sizeof(fc));
  static const int knob = [] {
    // A/B knob: 1 = always the generic march, 0 = the fast march wherever it applies;
    // unset: the fast march for single volumes only -- on a z-slab rank the generic march is
    // 3.5 % faster (N = 2: 0.481 vs 0.499 ms per launch, profiles/r02z_ab_n2.txt): its
    // per-sample slab test costs more in the fast loop than the one-FMA keys save
    const char *e = getenv("B200TSDF_NO_FAST");
    return e ? atoi(e) : -1;
  }();
  const bool off = knob == 1 || (knob < 0 && h->d.slab_count > 1);
  const VolumeDev &d = h->d;
  const double vs = (double)(float)h->cfg.voxel_size, tr = (double)(float)h->cfg.truncation;
  bool ok = !off && d.bounded && d.dir_on && h->cfg.capacity_log2 <= 23 && step_d > 1.0e-6 &&
            step_d < 1.0e6 && vs > 1.0e-6 && vs < 1.0e6;
  for (int a = 0; a < 3 && ok; ++a) {
    // world coordinates whose band (+- trunc) stays 3 voxels inside the key bounds
    const double lo = ((double)d.kmin[a] + 3.0) * vs + tr, hi = ((double)d.kmax[a] - 3.0) * vs - tr;
    if (!(hi > lo)) {
      ok = false;
      break;
    }
    fc.in_sc[a] = (float)(2.0 / (hi - lo) * (1.0 - 1.0e-6));
    fc.in_of[a] = (float)(-(hi + lo) / (hi - lo));
  }
  fc.band_k = (int)ceil(2.0 * tr / step_d) + 1;
  if (fc.band_k > 1024) ok = false; // per-lane loss counters of the march are 12 bits wide
  fc.nx = (uint32_t)d.dir_n[0], fc.ny = (uint32_t)d.dir_n[1];
  const uint32_t m3 = (uint32_t)MARCH_KEY_MAGIC_BITS >> 3;
  fc.dir_c = ((m3 + (uint32_t)d.dir_b0[2]) * fc.ny + (m3 + (uint32_t)d.dir_b0[1])) * fc.nx +
             (m3 + (uint32_t)d.dir_b0[0]);
  fc.enabled = ok ? 1 : 0;
  return fc;
}

extern "C" int b200tsdf_integrate_points(b200tsdf_t *h, const void *pts_dev, int dtype,
                                         const double *conf_dev, const double *conf_q_dev,
                                         const void *origin_host, int64_t m, double step_scale,
                                         int64_t *rays_fused_host, void *stream) {
  B200_NVTX("b200tsdf_integrate_points");
  if (!h || (m > 0 && (!pts_dev || !conf_dev || !origin_host)) || m < 0 ||
      (dtype != B200TSDF_F32 && dtype != B200TSDF_F64)) {
    b200_set_error("b200tsdf_integrate_points: invalid argument");
    return B200TSDF_E_INVALID;
  }
  cudaStream_t st = (cudaStream_t)stream;
  B200_CUDA(cudaSetDevice(h->cfg.device));
  unsigned long long before = 0;
  if (rays_fused_host) {
    unsigned long long c[CTR_COUNT];
    int rc = b200_read_counters(h, st, c);
    if (rc) return rc;
    before = c[CTR_RAYS];
  }
  if (m > 0) {
    double step_d = h->cfg.voxel_size * step_scale; // global_volume.py:46
    if (step_d < 1.0e-4) step_d = 1.0e-4;
    unsigned grid = (unsigned)((m + 255) / 256);
    if (dtype == B200TSDF_F32) {
      const float *o = (const float *)origin_host;
      const MarchConst mc = make_consts(h);
#define LAUNCH_PTS(BD, DR)                                                                     \
  integrate_points_f32_kernel<BD, DR><<<grid, 256, 0, st>>>(                                    \
      h->d, (const float *)pts_dev, conf_dev, conf_q_dev, o[0], o[1], o[2], (long long)m,       \
      (float)step_d, mc)
      if (!use_bounded(h))
        LAUNCH_PTS(false, false);
      else if (h->d.dir_on)
        LAUNCH_PTS(true, true);
      else
        LAUNCH_PTS(true, false);
#undef LAUNCH_PTS
    } else {
      const double *o = (const double *)origin_host;
      integrate_points_f64_kernel<<<grid, 256, 0, st>>>(
          h->d, (const double *)pts_dev, conf_dev, conf_q_dev, o[0], o[1], o[2], (long long)m,
          step_d, h->cfg.truncation, h->cfg.voxel_size);
    }
    B200_CUDA(cudaGetLastError());
  }
  if (rays_fused_host) {
    unsigned long long after = 0;
    unsigned long long c[CTR_COUNT];
    int rc = b200_read_counters(h, st, c);
    if (rc) return rc;
    after = c[CTR_RAYS];
    *rays_fused_host = (int64_t)(after - before);
  }
  return B200TSDF_OK;
}

static int fuse_keyframes(b200tsdf_t *h, const b200tsdf_kf_batch *kb, void *stream) {
  if (!h || !kb) {
    b200_set_error("b200tsdf_integrate_keyframes: null argument");
    return B200TSDF_E_INVALID;
  }
  const float *X_dev = kb->X_dev, *C_dev = kb->C_dev, *q_grid_dev = kb->q_grid_dev,
              *pose_dev = kb->pose_dev;
  const int32_t B = kb->B, H = kb->H, W = kb->W, gh = kb->gh, gw = kb->gw;
  if (B < 0 || H <= 0 || W <= 0 || (B > 0 && (!X_dev || !C_dev || !pose_dev))) {
    b200_set_error("b200tsdf_integrate_keyframes: invalid argument");
    return B200TSDF_E_INVALID;
  }
  if (q_grid_dev && (gh <= 0 || gw <= 0 || H % gh || W % gw)) {
    b200_set_error("b200tsdf_integrate_keyframes: quality grid %dx%d does not tile %dx%d", gh, gw,
                   H, W);
    return B200TSDF_E_INVALID;
  }
  if (kb->seg_stride != 0 && (kb->seg_kf <= 0 || kb->seg_stride < 0)) {
    b200_set_error("b200tsdf_integrate_keyframes: bad segment layout (seg_kf %d, stride %lld)",
                   kb->seg_kf, (long long)kb->seg_stride);
    return B200TSDF_E_INVALID;
  }
  if (B == 0) return B200TSDF_OK;
  cudaStream_t st = (cudaStream_t)stream;
  B200_CUDA(cudaSetDevice(h->cfg.device));
  double step_d = h->cfg.voxel_size * kb->step_scale;
  if (step_d < 1.0e-4) step_d = 1.0e-4;
  dim3 grid((unsigned)((W + TILE_W - 1) / TILE_W), (unsigned)((H + TILE_H - 1) / TILE_H),
            (unsigned)B);
  MarchConst mc = q_grid_dev ? make_consts(h, H / gh, W / gw) : make_consts(h);
  mc.negate = kb->negate ? 1 : 0;
  const FastConst fc = make_fast_consts(h, step_d);
  // TMA bulk copies need 16-byte aligned global addresses: bases, per-keyframe and
  // per-segment strides (W % 32 == 0 makes every row segment aligned within a keyframe)
  const bool aligned = (((uintptr_t)X_dev & 15) == 0) && (((uintptr_t)C_dev & 15) == 0) &&
                       ((kb->seg_stride & 3) == 0);
  const int mode = !use_bounded(h) ? 0 : (h->d.dir_on ? 2 : 1);
  // TMA staging needs full tiles and 16-byte aligned row segments
  const bool tma = (W % TILE_W == 0) && (H % TILE_H == 0) && aligned;
  KfArgs ka;
  memset(&ka, 0, sizeof(ka));
  ka.X = X_dev, ka.C = C_dev, ka.qg = q_grid_dev, ka.poses = pose_dev;
  ka.seg_stride = (long long)kb->seg_stride, ka.seg_kf = kb->seg_kf > 0 ? kb->seg_kf : 1;
  ka.H = H, ka.W = W, ka.gh = gh, ka.gw = gw, ka.min_conf = kb->min_conf, ka.step = (float)step_d;
  if (kb->max_points > 0 && (long long)kb->max_points < (long long)H * W) {
    // the reference's per-keyframe cap (global_manager.py:94-99), seeded and exact-size
    if ((size_t)B > h->sub_cap) {
      cudaFree(h->sub_thr);
      h->sub_thr = nullptr, h->sub_cap = 0;
      B200_CUDA(cudaMalloc(&h->sub_thr, (size_t)B * sizeof(uint2)));
      h->sub_cap = (size_t)B;
    }
    int list_cap = SUB_LIST_CAP;
    if (const char *e = getenv("B200TSDF_SUB_LIST_CAP")) { // test knob, read per call: 0 = radix passes only
      const int n = atoi(e);
      list_cap = n < 0 ? 0 : (n > SUB_LIST_CAP ? SUB_LIST_CAP : n);
    }
    subsample_threshold_kernel<<<dim3(SUB_SPLIT, (unsigned)B), SUB_THREADS, 0, st>>>(
        ka, kb->max_points, kb->seed, (const long long *)kb->kf_ids_dev, (uint2 *)h->sub_thr, list_cap);
    ka.sub = (const uint2 *)h->sub_thr;
  }
#define DISPATCH_KF(LAUNCH)                                                                    \
  if (tma) {                                                                                   \
    if (mode == 0) LAUNCH(false, false, true);                                                 \
    else if (mode == 1) LAUNCH(true, false, true);                                             \
    else LAUNCH(true, true, true);                                                             \
  } else {                                                                                     \
    if (mode == 0) LAUNCH(false, false, false);                                                \
    else if (mode == 1) LAUNCH(true, false, false);                                            \
    else LAUNCH(true, true, false);                                                            \
  }
  const long long n_tiles = (long long)grid.x * grid.y * grid.z;
  static const int classify_min_ranks = [] {
    const char *e = getenv("B200TSDF_CLASSIFY_MIN_RANKS"); // tuning knob (default 3)
    return e ? atoi(e) : 3;
  }();
  if (h->d.slab_count >= classify_min_ranks && h->d.slab_count > 1 && n_tiles < (1ll << 31)) {
    // multi-GPU: classify pass + persistent fuse over the listed tiles (with two
    // ranks nearly every tile reaches both, so the direct launch is cheaper)
    if ((size_t)n_tiles > h->tile_cap) {
      cudaFree(h->tile_list);
      h->tile_list = nullptr, h->tile_cap = 0;
      B200_CUDA(cudaMalloc(&h->tile_list, (size_t)n_tiles * sizeof(int)));
      h->tile_cap = (size_t)n_tiles;
    }
    B200_CUDA(cudaMemsetAsync(h->sched, 0, 2 * sizeof(unsigned int), st));
    const dim3 cgrid((grid.x + CLS_TILES - 1) / CLS_TILES, grid.y, grid.z);
    slab_classify_kernel<<<cgrid, 256, 0, st>>>(h->d, ka, mc, h->tile_list, h->sched, (int)grid.x,
                                                (int)grid.y);
    const long long cap = (long long)h->sm_count * FUSE_MINB_LISTED;
    const unsigned ctas = (unsigned)(n_tiles < cap ? n_tiles : cap);
#define LAUNCH_LISTED(BD, DR, TM)                                                              \
  integrate_keyframes_listed_kernel<BD, DR, TM><<<ctas, 256, 0, st>>>(                          \
      h->d, ka, mc, fc, h->tile_list, h->sched, (int)grid.x, (int)grid.y)
    DISPATCH_KF(LAUNCH_LISTED)
#undef LAUNCH_LISTED
  } else {
    static const int kf_order = [] {
      const char *e = getenv("B200TSDF_KF_ORDER"); // A/B knob, see KfArgs::order
      return e ? atoi(e) : FUSE_KF_ORDER;
    }();
    int order = kf_order < 1 ? 1 : kf_order;
    while (order > 1 && (B % order) != 0) --order; // largest divisor of the batch <= the knob
    static const int cap_group = [] {
      const char *e = getenv("B200TSDF_CAP_GROUP"); // A/B knob: 0 = one tile per CTA also under the cap
      return e ? atoi(e) : 1;
    }();
    if (ka.sub && cap_group) {
      // max_points_per_kf: CAP_G tiles per CTA, survivors compacted across them
      const dim3 cgrid(grid.x, (grid.y + CAP_G - 1) / CAP_G, (unsigned)B);
#define LAUNCH_CAP(BD, DR, TM) integrate_keyframes_capped_kernel<BD, DR, TM><<<cgrid, CAP_THREADS, 0, st>>>(h->d, ka, mc, fc)
      DISPATCH_KF(LAUNCH_CAP)
#undef LAUNCH_CAP
    } else {
      ka.order = order;
      const dim3 ogrid(grid.x * (unsigned)order, grid.y, (unsigned)(B / order));
#define LAUNCH_KF(BD, DR, TM) integrate_keyframes_kernel<BD, DR, TM><<<ogrid, 256, 0, st>>>(h->d, ka, mc, fc)
      DISPATCH_KF(LAUNCH_KF)
#undef LAUNCH_KF
    }
  }
#undef DISPATCH_KF
  B200_CUDA(cudaGetLastError());
  return B200TSDF_OK;
}

static b200tsdf_kf_batch plain_batch(const float *X_dev, const float *C_dev,
                                     const float *q_grid_dev, const float *pose_dev, int32_t B,
                                     int32_t H, int32_t W, int32_t gh, int32_t gw, float min_conf,
                                     double step_scale, int negate) {
  b200tsdf_kf_batch kb;
  memset(&kb, 0, sizeof(kb));
  kb.X_dev = X_dev, kb.C_dev = C_dev, kb.q_grid_dev = q_grid_dev, kb.pose_dev = pose_dev;
  kb.B = B, kb.H = H, kb.W = W, kb.gh = gh, kb.gw = gw;
  kb.min_conf = min_conf, kb.step_scale = step_scale, kb.negate = negate;
  return kb;
}

extern "C" int b200tsdf_integrate_keyframes(b200tsdf_t *h, const float *X_dev, const float *C_dev,
                                            const float *q_grid_dev, const float *pose_dev,
                                            int32_t B, int32_t H, int32_t W, int32_t gh,
                                            int32_t gw, float min_conf, double step_scale,
                                            void *stream) {
  B200_NVTX("b200tsdf_integrate_keyframes");
  const b200tsdf_kf_batch kb =
      plain_batch(X_dev, C_dev, q_grid_dev, pose_dev, B, H, W, gh, gw, min_conf, step_scale, 0);
  return fuse_keyframes(h, &kb, stream);
}

extern "C" int b200tsdf_deintegrate_keyframes(b200tsdf_t *h, const float *X_dev,
                                              const float *C_dev, const float *q_grid_dev,
                                              const float *pose_dev, int32_t B, int32_t H,
                                              int32_t W, int32_t gh, int32_t gw, float min_conf,
                                              double step_scale, void *stream) {
  B200_NVTX("b200tsdf_deintegrate_keyframes");
  const b200tsdf_kf_batch kb =
      plain_batch(X_dev, C_dev, q_grid_dev, pose_dev, B, H, W, gh, gw, min_conf, step_scale, 1);
  return fuse_keyframes(h, &kb, stream);
}

extern "C" int b200tsdf_integrate_keyframes_ex(b200tsdf_t *h, const b200tsdf_kf_batch *batch,
                                               void *stream) {
  B200_NVTX("b200tsdf_integrate_keyframes_ex");
  return fuse_keyframes(h, batch, stream);
}

int b200_ensure_stage(b200tsdf *h, size_t bytes);

extern "C" int b200tsdf_integrate_keyframes_host(b200tsdf_t *h, const float *X_host,
                                                 const float *C_host, const float *q_grid_host,
                                                 const float *pose_host, int32_t B, int32_t H,
                                                 int32_t W, int32_t gh, int32_t gw,
                                                 float min_conf, double step_scale,
                                                 void *stream) {
  B200_NVTX("b200tsdf_integrate_keyframes_host");
  if (!h || B < 0 || H <= 0 || W <= 0 || (B > 0 && (!X_host || !C_host || !pose_host))) {
    b200_set_error("b200tsdf_integrate_keyframes_host: invalid argument");
    return B200TSDF_E_INVALID;
  }
  if (B == 0) return B200TSDF_OK;
  cudaStream_t st = (cudaStream_t)stream;
  B200_CUDA(cudaSetDevice(h->cfg.device));
  const size_t npix = (size_t)H * W;
  const size_t bq = q_grid_host ? (size_t)B * gh * gw * sizeof(float) : 0;
  const size_t bp = (size_t)B * 8 * sizeof(float);
  auto al = [](size_t x) { return (x + 255) & ~(size_t)255; };
  // Chunked pipeline: the H2D copy of chunk c+1 (copy stream) overlaps the fuse of
  // chunk c (caller's stream); the call stays stream-ordered on `st`.  PCIe pays per
  // copy -- this box moves 55.4 GB/s in 64 MiB copies, 52.6 in 4 MiB ones and loses
  // ~2 us on every tiny one (scripts/h2d_probe.py) -- and per call (the final wait for
  // the copies is a bubble on the link).  So: quality grids and poses of the whole
  // batch first, one copy each; pointmaps / confidences in chunks of up to
  // `chunk_kf` keyframes (one fuse launch each, at least 4 chunks so that the fuse
  // nothing overlaps is a quarter of the batch); and the chunks go through a ring of
  // three staging slots, so a caller may hand over a whole trajectory in one call
  // without the staging area growing with it.
  static const int chunk_kf_knob = [] {
    const char *e = getenv("B200TSDF_H2D_CHUNK_KF"); // A/B knob: keyframes per chunk (default 40)
    const int n = e ? atoi(e) : 40;
    return n < 1 ? 1 : n;
  }();
  const size_t kf_bytes = npix * 4 * sizeof(float);
  int chunk_kf = chunk_kf_knob;
  if ((size_t)chunk_kf * kf_bytes > ((size_t)256 << 20)) // a slot holds at most 256 MiB
    chunk_kf = (int)(((size_t)256 << 20) / kf_bytes) > 1 ? (int)(((size_t)256 << 20) / kf_bytes) : 1;
  int nchunk = (B + chunk_kf - 1) / chunk_kf;
  if (nchunk < 4) nchunk = B < 4 ? B : 4;
  const int slot_kf = (B + nchunk - 1) / nchunk; // chunks differ by at most one keyframe
  const size_t slot_x = al((size_t)slot_kf * npix * 3 * sizeof(float));
  const size_t slot_c = al((size_t)slot_kf * npix * sizeof(float));
  const int nslot = nchunk < 3 ? nchunk : 3;
  const size_t total = al(bq) + al(bp) + (size_t)nslot * (slot_x + slot_c);
  int rc = b200_ensure_stage(h, total);
  if (rc) return rc;
  char *base = (char *)h->stage_dev;
  float *dQ = (float *)base;
  float *dP = (float *)(base + al(bq));
  char *slots = base + al(bq) + al(bp);
  if (!h->copy_stream) {
    B200_CUDA(cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
    for (int i = 0; i <= B200_H2D_CHUNKS; ++i)
      B200_CUDA(cudaEventCreateWithFlags(&h->copy_ev[i], cudaEventDisableTiming));
  }
  // events: [0] "chunk copied" (re-recorded per chunk: a stream wait refers to the record
  // that precedes it), [1..3] "slot fused", [B200_H2D_CHUNKS] "staging area free"
  cudaStream_t cs = h->copy_stream;
  // the staging area may still be read by kernels queued on `st` earlier
  B200_CUDA(cudaEventRecord(h->copy_ev[B200_H2D_CHUNKS], st));
  B200_CUDA(cudaStreamWaitEvent(cs, h->copy_ev[B200_H2D_CHUNKS], 0));
  const size_t qsz = (size_t)gh * gw;
  if (bq) B200_CUDA(cudaMemcpyAsync(dQ, q_grid_host, bq, cudaMemcpyHostToDevice, cs));
  B200_CUDA(cudaMemcpyAsync(dP, pose_host, bp, cudaMemcpyHostToDevice, cs));
  for (int c = 0; c < nchunk; ++c) {
    const int b0 = (int)((long long)B * c / nchunk), b1 = (int)((long long)B * (c + 1) / nchunk);
    const size_t nb = (size_t)(b1 - b0);
    const int slot = c % nslot;
    float *dX = (float *)(slots + (size_t)slot * (slot_x + slot_c));
    float *dC = (float *)(slots + (size_t)slot * (slot_x + slot_c) + slot_x);
    if (c >= nslot) B200_CUDA(cudaStreamWaitEvent(cs, h->copy_ev[1 + slot], 0)); // chunk c - 3 fused
    B200_CUDA(cudaMemcpyAsync(dX, X_host + (size_t)b0 * npix * 3, nb * npix * 3 * sizeof(float),
                              cudaMemcpyHostToDevice, cs));
    B200_CUDA(cudaMemcpyAsync(dC, C_host + (size_t)b0 * npix, nb * npix * sizeof(float),
                              cudaMemcpyHostToDevice, cs));
    B200_CUDA(cudaEventRecord(h->copy_ev[0], cs));
    B200_CUDA(cudaStreamWaitEvent(st, h->copy_ev[0], 0));
    rc = b200tsdf_integrate_keyframes(h, dX, dC, bq ? dQ + (size_t)b0 * qsz : nullptr,
                                      dP + (size_t)b0 * 8, (int32_t)nb, H, W, gh, gw, min_conf,
                                      step_scale, st);
    if (rc) return rc;
    if (c + nslot < nchunk) B200_CUDA(cudaEventRecord(h->copy_ev[1 + slot], st));
  }
  // the host buffers are the caller's again when this returns: wait for the copies
  // (not for the fuse, which stays queued on `stream`)
  B200_CUDA(cudaStreamSynchronize(h->copy_stream));
  return B200TSDF_OK;
}

// ---------------------------------------------------------------- peer pull
// Multi-GPU ingest without broadcasting the pointmaps:
// a z-slab rank needs only the ~1.4 / N of the batch's 32 x 8 pixel tiles whose
// +-truncation band reaches its slabs.  The OWNER of a keyframe (the rank whose HBM
// holds it) runs keyframe_meta_kernel once: per tile the conservative z range of
// the bands of its rays (voxel layers) and the number of rays, plus copies of the
// pose and the quality grid -- ~28 KB per keyframe instead of 3.1 MB.  ONE small
// all-gather moves the meta records; every rank then lists the tiles that reach
// its slabs (tiles_from_meta_kernel, which also does the accounting of the
// skipped tiles' rays) and the persistent fuse kernel TMA-loads exactly those
// tiles straight from the owner's memory over NVLink (KfArgs::segX / segC are
// peer-mapped pointers): the staging of the fuse IS the collective.
struct TileMeta {
  int zlo, zhi;    // voxel layers the tile's bands can touch; zlo > zhi: no ray in the tile
  uint32_t rays;   // rays of the tile the reference would fuse (global_volume.py:56)
  uint32_t pad;
};

static long long meta_words(int seg_kf, int H, int W, int gh, int gw) {
  const long long tiles = (long long)((W + TILE_W - 1) / TILE_W) * ((H + TILE_H - 1) / TILE_H);
  long long w = (long long)seg_kf * tiles * 4 + (long long)seg_kf * 8 + (long long)seg_kf * gh * gw;
  return (w + 63) & ~63ll;
}

// CTA = CLS_TILES consecutive tiles of one tile row of one keyframe (as slab_classify_kernel).
__global__ void __launch_bounds__(256)
    keyframe_meta_kernel(KfArgs a, MarchConst mc, TileMeta *__restrict__ table, int tiles_x,
                         int tiles_y) {
  __shared__ int s_lo[8], s_hi[8];
  __shared__ uint32_t s_rays[8];
  const int lx = threadIdx.x & 31, warp = threadIdx.x >> 5, py = blockIdx.y * TILE_H + warp;
  const int kf = blockIdx.z;
  const KfPtrs k = kf_ptrs(a, kf);
  const float *P = k.P;
  const long long row = (long long)py * a.W;
  float c[CLS_TILES], x0[CLS_TILES], x1[CLS_TILES], x2[CLS_TILES];
#pragma unroll
  for (int t = 0; t < CLS_TILES; ++t) {
    const int px = (blockIdx.x * CLS_TILES + t) * TILE_W + lx;
    c[t] = 0.f, x0[t] = x1[t] = x2[t] = 0.f;
    if (px < a.W && py < a.H) {
      c[t] = k.C[row + px];
      const float *x = k.X + (row + px) * 3;
      x0[t] = x[0], x1[t] = x[1], x2[t] = x[2];
    }
  }
  const float q0 = P[3], q1 = P[4], q2 = P[5], q3 = P[6], sv = P[7] * mc.inv_vs;
  const float k0 = 2.0f * (q0 * q2 - q1 * q3) * sv, k1 = 2.0f * (q1 * q2 + q0 * q3) * sv,
              k2 = (1.0f - 2.0f * (q0 * q0 + q1 * q1)) * sv, tzv = P[2] * mc.inv_vs;
  const float s2 = P[7] * P[7], tv = mc.trunc * mc.inv_vs;
#pragma unroll
  for (int t = 0; t < CLS_TILES; ++t) {
    const int bx = blockIdx.x * CLS_TILES + t;
    if (bx >= tiles_x) break; // uniform
    const int px = bx * TILE_W + lx;
    const bool take = px < a.W && py < a.H && c[t] > a.min_conf;
    int zlo = 0x7FFFFFFF, zhi = (int)0x80000000;
    uint32_t ray = 0;
    if (take) {
      // the conservative z test of slab_classify_kernel, kept as a RANGE
      const float zc = fmaf(x0[t], k0, fmaf(x1[t], k1, fmaf(x2[t], k2, tzv)));
      const float mag = fmaf(fabsf(x0[t]), fabsf(k0),
                             fmaf(fabsf(x1[t]), fabsf(k1), fmaf(fabsf(x2[t]), fabsf(k2), fabsf(tzv))));
      const float d2 = s2 * fmaf(x0[t], x0[t], fmaf(x1[t], x1[t], x2[t] * x2[t]));
      float dz = fmaf(fabsf(zc - tzv) * mc.vs * rsqrtf(d2), 1.01f, 1.0e-3f);
      if (!(dz < 1.0f)) dz = 1.0f;
      const float ext = fmaf(tv, dz, fmaf(mag, 4.0e-6f, 3.5f));
      if (fabsf(zc) < 1.0e9f && ext < 1.0e9f) {
        zlo = (int)floorf(zc - ext), zhi = (int)floorf(zc + ext);
      } else { // the per-voxel test decides: every rank takes the tile
        zlo = (int)0x80000000, zhi = 0x7FFFFFFF;
      }
      bool is_ray = d2 > 4.0e-6f && d2 < 1.0e36f; // certainly a ray when 2e-3 < len < 1e18
      if (!is_ray) {
        float wx, wy, wz, rx, ry, rz;
        sim3_act(P, x0[t], x1[t], x2[t], wx, wy, wz);
        Ray<float> r;
        is_ray = ray_length<float>(r, wx, wy, wz, P[0], P[1], P[2], rx, ry, rz);
      }
      ray = is_ray ? 1u : 0u;
    }
    zlo = __reduce_min_sync(0xFFFFFFFFu, zlo);
    zhi = __reduce_max_sync(0xFFFFFFFFu, zhi);
    ray = __reduce_add_sync(0xFFFFFFFFu, ray);
    __syncthreads(); // the previous tile's totals have been read
    if (lx == 0) s_lo[warp] = zlo, s_hi[warp] = zhi, s_rays[warp] = ray;
    __syncthreads();
    if (threadIdx.x == 0) {
      TileMeta m = {0x7FFFFFFF, (int)0x80000000, 0u, 0u};
      for (int w = 0; w < 8; ++w) {
        m.zlo = min(m.zlo, s_lo[w]), m.zhi = max(m.zhi, s_hi[w]);
        m.rays += s_rays[w];
      }
      table[((long long)kf * tiles_y + blockIdx.y) * tiles_x + bx] = m;
    }
  }
  // the first CTA of a keyframe also copies its pose and quality grid behind the
  // tile table: [tile table | poses | quality grids]
  if (blockIdx.x == 0 && blockIdx.y == 0) {
    float *tail = reinterpret_cast<float *>(table + (long long)gridDim.z * tiles_y * tiles_x);
    if (threadIdx.x < 8) tail[kf * 8 + threadIdx.x] = P[threadIdx.x];
    if (k.qg) {
      float *q = tail + (long long)gridDim.z * 8 + (long long)kf * a.gh * a.gw;
      for (int i = threadIdx.x; i < a.gh * a.gw; i += 256) q[i] = k.qg[i];
    }
  }
}

// One thread per tile of the gathered batch: list it if its z range reaches a slab
// of this rank, else account for its rays (counted, culled) right here.
__global__ void __launch_bounds__(256)
    tiles_from_meta_kernel(VolumeDev v, const uint32_t *__restrict__ meta, long long meta_stride,
                           int seg_kf, int n_kf, int tiles_kf, int *__restrict__ list,
                           unsigned int *sched) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  Tally tl = {0, 0, 0, 0, 0, 0, 0};
  if (i < (long long)n_kf * tiles_kf) {
    const int kf = (int)(i / tiles_kf), tile = (int)(i - (long long)kf * tiles_kf);
    const int seg = kf / seg_kf, loc = kf - seg * seg_kf;
    const TileMeta m = reinterpret_cast<const TileMeta *>(meta + (long long)seg * meta_stride)
        [(long long)loc * tiles_kf + tile];
    if (m.zlo <= m.zhi) {
      bool hit = true;
      if (v.slab_count > 1 && m.zlo > -1000000000 && m.zhi < 1000000000) {
        const int s_lo = (m.zlo - v.slab_z0) >> v.slab_shift, s_hi = (m.zhi - v.slab_z0) >> v.slab_shift;
        int need = v.slab_index - slab_owner_of(v, s_lo);
        if (need < 0) need += v.slab_count;
        hit = need <= s_hi - s_lo;
      }
      if (hit) {
        list[atomicAdd(sched, 1u)] = (int)i;
      } else {
        tl.rays = m.rays, tl.culled = m.rays;
      }
    }
  }
  commit_tally(v, tl);
}

extern "C" int64_t b200tsdf_keyframe_meta_words(int32_t seg_kf, int32_t H, int32_t W, int32_t gh,
                                                int32_t gw) {
  return (int64_t)meta_words(seg_kf, H, W, gh, gw);
}

extern "C" int b200tsdf_keyframe_meta(b200tsdf_t *h, const float *X_dev, const float *C_dev,
                                      const float *q_grid_dev, const float *pose_dev, int32_t B,
                                      int32_t H, int32_t W, int32_t gh, int32_t gw, float min_conf,
                                      uint32_t *meta_dev, void *stream) {
  B200_NVTX("b200tsdf_keyframe_meta");
  if (!h || B < 0 || H <= 0 || W <= 0 || gh < 0 || gw < 0 ||
      (B > 0 && (!X_dev || !C_dev || !pose_dev || !meta_dev)) || (q_grid_dev && (gh == 0 || gw == 0))) {
    b200_set_error("b200tsdf_keyframe_meta: invalid argument");
    return B200TSDF_E_INVALID;
  }
  if (B == 0) return B200TSDF_OK;
  cudaStream_t st = (cudaStream_t)stream;
  B200_CUDA(cudaSetDevice(h->cfg.device));
  const int tiles_x = (W + TILE_W - 1) / TILE_W, tiles_y = (H + TILE_H - 1) / TILE_H;
  KfArgs ka;
  memset(&ka, 0, sizeof(ka));
  ka.X = X_dev, ka.C = C_dev, ka.qg = q_grid_dev, ka.poses = pose_dev;
  ka.seg_kf = B, ka.H = H, ka.W = W, ka.gh = gh, ka.gw = gw, ka.min_conf = min_conf;
  const MarchConst mc = make_consts(h);
  const dim3 grid((unsigned)((tiles_x + CLS_TILES - 1) / CLS_TILES), (unsigned)tiles_y, (unsigned)B);
  keyframe_meta_kernel<<<grid, 256, 0, st>>>(ka, mc, (TileMeta *)meta_dev, tiles_x, tiles_y);
  B200_CUDA(cudaGetLastError());
  return B200TSDF_OK;
}

extern "C" int b200tsdf_integrate_keyframes_pull(b200tsdf_t *h, const b200tsdf_kf_pull *kp,
                                                 void *stream) {
  B200_NVTX("b200tsdf_integrate_keyframes_pull");
  if (!h || !kp || kp->n_seg <= 0 || kp->n_seg > B200TSDF_MAX_SEGMENTS || kp->seg_kf <= 0 ||
      kp->H <= 0 || kp->W <= 0 || !kp->meta_dev || kp->meta_stride <= 0) {
    b200_set_error("b200tsdf_integrate_keyframes_pull: invalid argument");
    return B200TSDF_E_INVALID;
  }
  if (!use_bounded(h) || !h->d.dir_on) {
    b200_set_error("b200tsdf_integrate_keyframes_pull: needs a bounded volume with a block directory");
    return B200TSDF_E_INVALID;
  }
  const int H = kp->H, W = kp->W, gh = kp->gh, gw = kp->gw, B = kp->n_seg * kp->seg_kf;
  if (gh > 0 && (gw <= 0 || H % gh || W % gw)) {
    b200_set_error("b200tsdf_integrate_keyframes_pull: quality grid %dx%d does not tile %dx%d", gh,
                   gw, H, W);
    return B200TSDF_E_INVALID;
  }
  cudaStream_t st = (cudaStream_t)stream;
  B200_CUDA(cudaSetDevice(h->cfg.device));
  double step_d = h->cfg.voxel_size * kp->step_scale;
  if (step_d < 1.0e-4) step_d = 1.0e-4;
  const int tiles_x = (W + TILE_W - 1) / TILE_W, tiles_y = (H + TILE_H - 1) / TILE_H;
  const int tiles_kf = tiles_x * tiles_y;
  const long long n_tiles = (long long)tiles_kf * B;
  if (n_tiles >= (1ll << 31)) {
    b200_set_error("b200tsdf_integrate_keyframes_pull: batch too large");
    return B200TSDF_E_INVALID;
  }
  MarchConst mc = gh > 0 ? make_consts(h, H / gh, W / gw) : make_consts(h);
  mc.negate = kp->negate ? 1 : 0;
  const FastConst fc = make_fast_consts(h, step_d);
  KfArgs ka;
  memset(&ka, 0, sizeof(ka));
  const float *meta_f = reinterpret_cast<const float *>(kp->meta_dev);
  const size_t tab = (size_t)kp->seg_kf * tiles_kf * 4;
  ka.poses = meta_f + tab;
  ka.qg = gh > 0 ? meta_f + tab + (size_t)kp->seg_kf * 8 : nullptr;
  ka.seg_stride = (long long)kp->meta_stride, ka.seg_kf = kp->seg_kf;
  ka.H = H, ka.W = W, ka.gh = gh, ka.gw = gw, ka.min_conf = kp->min_conf, ka.step = (float)step_d;
  bool aligned = true;
  for (int s = 0; s < kp->n_seg; ++s) {
    if (!kp->X_seg_dev[s] || !kp->C_seg_dev[s]) {
      b200_set_error("b200tsdf_integrate_keyframes_pull: segment %d has no pointers", s);
      return B200TSDF_E_INVALID;
    }
    ka.segX[s] = kp->X_seg_dev[s], ka.segC[s] = kp->C_seg_dev[s];
    aligned = aligned && (((uintptr_t)kp->X_seg_dev[s] & 15) == 0) &&
              (((uintptr_t)kp->C_seg_dev[s] & 15) == 0);
  }
  ka.n_ptr = kp->n_seg;
  ka.X = ka.segX[0], ka.C = ka.segC[0];
  const bool tma = (W % TILE_W == 0) && (H % TILE_H == 0) && aligned && ((H * (long long)W) % 4 == 0);
  if ((size_t)n_tiles > h->tile_cap) {
    cudaFree(h->tile_list);
    h->tile_list = nullptr, h->tile_cap = 0;
    B200_CUDA(cudaMalloc(&h->tile_list, (size_t)n_tiles * sizeof(int)));
    h->tile_cap = (size_t)n_tiles;
  }
  B200_CUDA(cudaMemsetAsync(h->sched, 0, 2 * sizeof(unsigned int), st));
  tiles_from_meta_kernel<<<(unsigned)((n_tiles + 255) / 256), 256, 0, st>>>(
      h->d, kp->meta_dev, (long long)kp->meta_stride, kp->seg_kf, B, tiles_kf, h->tile_list, h->sched);
  const long long cap = (long long)h->sm_count * FUSE_MINB_LISTED;
  const unsigned ctas = (unsigned)(n_tiles < cap ? n_tiles : cap);
  if (tma)
    integrate_keyframes_listed_kernel<true, true, true><<<ctas, 256, 0, st>>>(
        h->d, ka, mc, fc, h->tile_list, h->sched, tiles_x, tiles_y);
  else
    integrate_keyframes_listed_kernel<true, true, false><<<ctas, 256, 0, st>>>(
        h->d, ka, mc, fc, h->tile_list, h->sched, tiles_x, tiles_y);
  B200_CUDA(cudaGetLastError());
  return B200TSDF_OK;
}

// ---- CUDA IPC: map another rank's keyframe buffer into this process (one process
// per GPU; NVSwitch gives every peer full NVLink bandwidth).
// The shareable buffer itself: its own cudaMalloc allocation (offset 0 in its IPC
// handle), so nothing depends on how a framework's caching allocator carves blocks.
extern "C" int b200tsdf_peer_alloc(int32_t device, int64_t bytes, void **ptr_out, void *handle64_out) {
  B200_NVTX("b200tsdf_peer_alloc");
  if (bytes <= 0 || !ptr_out || !handle64_out) {
    b200_set_error("b200tsdf_peer_alloc: invalid argument");
    return B200TSDF_E_INVALID;
  }
  B200_CUDA(cudaSetDevice(device));
  void *p = nullptr;
  if (cudaMalloc(&p, (size_t)bytes) != cudaSuccess) {
    (void)cudaGetLastError();
    b200_set_error("b200tsdf_peer_alloc: cudaMalloc of %lld bytes failed", (long long)bytes);
    return B200TSDF_E_NOMEM;
  }
  cudaIpcMemHandle_t hnd;
  cudaError_t e = cudaMemset(p, 0, (size_t)bytes);
  if (e == cudaSuccess) e = cudaIpcGetMemHandle(&hnd, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    b200_set_error("b200tsdf_peer_alloc: %s", cudaGetErrorString(e));
    return B200TSDF_E_CUDA;
  }
  memcpy(handle64_out, &hnd, sizeof(hnd));
  *ptr_out = p;
  return B200TSDF_OK;
}

extern "C" int b200tsdf_peer_free(int32_t device, void *ptr) {
  B200_NVTX("b200tsdf_peer_free");
  if (!ptr) return B200TSDF_OK;
  B200_CUDA(cudaSetDevice(device));
  B200_CUDA(cudaFree(ptr));
  return B200TSDF_OK;
}

extern "C" int b200tsdf_ipc_open(int32_t device, const void *handle64, int64_t offset_bytes,
                                 void **ptr_out) {
  B200_NVTX("b200tsdf_ipc_open");
  if (!handle64 || !ptr_out || offset_bytes < 0) {
    b200_set_error("b200tsdf_ipc_open: invalid argument");
    return B200TSDF_E_INVALID;
  }
  B200_CUDA(cudaSetDevice(device));
  cudaIpcMemHandle_t hnd;
  memcpy(&hnd, handle64, sizeof(hnd));
  void *base = nullptr;
  B200_CUDA(cudaIpcOpenMemHandle(&base, hnd, cudaIpcMemLazyEnablePeerAccess));
  *ptr_out = (char *)base + offset_bytes;
  return B200TSDF_OK;
}

extern "C" int b200tsdf_ipc_close(int32_t device, void *ptr, int64_t offset_bytes) {
  B200_NVTX("b200tsdf_ipc_close");
  if (!ptr) return B200TSDF_OK;
  B200_CUDA(cudaSetDevice(device));
  B200_CUDA(cudaIpcCloseMemHandle((char *)ptr - offset_bytes));
  return B200TSDF_OK;
}

// ---------------------------------------------------------------- self-test
// The shared-reciprocal division of fuse_fast.cuh must give the IEEE quotient
// for every operand pair in its documented range; this kernel compares it with
// __fdiv_rn on pseudo-random operands shaped like the three uses in
// fast_ray_setup (component / length, maxd / step, maxd / (n - 1)).
__device__ __forceinline__ uint64_t st_mix(uint64_t x) {
  x ^= x >> 33, x *= 0xff51afd7ed558ccdull, x ^= x >> 33, x *= 0xc4ceb9fe1a85ec53ull, x ^= x >> 33;
  return x;
}
__device__ __forceinline__ float st_logu(uint64_t h, float lo_exp, float hi_exp) {
  // 2^U(lo, hi) with a random 23-bit mantissa
  const float e = lo_exp + (hi_exp - lo_exp) * (float)((h >> 40) & 0xFFFF) * (1.0f / 65536.0f);
  const float m = 1.0f + (float)(h & 0x7FFFFF) * (1.0f / 8388608.0f);
  return exp2f(floorf(e)) * m;
}
__global__ void selftest_div_kernel(uint64_t seed, long long n, unsigned long long *bad) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t h0 = st_mix(seed + 3 * (uint64_t)i), h1 = st_mix(h0 + 1), h2 = st_mix(h1 + 2);
  float a, b;
  switch (i % 3) {
    case 0: // direction component / length: |a| <= b
      b = st_logu(h0, -13.0f, 56.0f);
      a = b * (float)((h1 >> 11) * (1.0 / 9007199254740992.0));
      if (h2 & 1) a = -a;
      if (fabsf(a) < 1.0e-18f) a = 1.0e-18f * 1.5f;
      break;
    case 1: // maxd / step
      b = st_logu(h0, -13.0f, 0.0f);
      a = st_logu(h1, -13.0f, 56.0f);
      break;
    default: // maxd / (n - 1): integer divisor
      b = (float)(1 + (h0 % 1073741823ull));
      a = st_logu(h1, -13.0f, 56.0f);
      break;
  }
  const float q = div_rn_r1(a, b, rcp_newton(b));
  const float ref = __fdiv_rn(a, b);
  if (__float_as_uint(q) != __float_as_uint(ref) && !(q == 0.0f && ref == 0.0f))
    atomicAdd(bad, 1ull);
}

extern "C" int b200tsdf_selftest_div(int32_t device, int64_t n, uint64_t seed,
                                     int64_t *mismatches_host, void *stream) {
  B200_NVTX("b200tsdf_selftest_div");
  if (n < 0 || !mismatches_host) {
    b200_set_error("b200tsdf_selftest_div: invalid argument");
    return B200TSDF_E_INVALID;
  }
  cudaStream_t st = (cudaStream_t)stream;
  B200_CUDA(cudaSetDevice(device));
  unsigned long long *bad = nullptr;
  B200_CUDA(cudaMalloc(&bad, sizeof(*bad)));
  B200_CUDA(cudaMemsetAsync(bad, 0, sizeof(*bad), st));
  if (n > 0) selftest_div_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(seed, (long long)n, bad);
  unsigned long long hbad = 0;
  cudaError_t e = cudaMemcpyAsync(&hbad, bad, sizeof(hbad), cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  cudaFree(bad);
  if (e != cudaSuccess) {
    b200_set_error("b200tsdf_selftest_div: %s", cudaGetErrorString(e));
    return B200TSDF_E_CUDA;
  }
  *mismatches_host = (int64_t)hbad;
  return B200TSDF_OK;
}
