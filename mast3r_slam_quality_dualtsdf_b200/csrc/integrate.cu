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
#include <stdlib.h>

#include "common.cuh"
#include "ray.cuh"
#include "march_fast.cuh"
#include "fuse.cuh"

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
                                            uint32_t &dropped, uint32_t &overflow) {
  if (!a.live) return;
  if (v.bounded && (a.kx < v.kmin[0] || a.kx >= v.kmax[0] || a.ky < v.kmin[1] ||
                    a.ky >= v.kmax[1] || a.kz < v.kmin[2] || a.kz >= v.kmax[2])) {
    dropped += a.np + a.nq;
    return;
  }
  int bx = a.kx >> B200_BLOCK_LOG2, by = a.ky >> B200_BLOCK_LOG2, bz = a.kz >> B200_BLOCK_LOG2;
  if (!block_in_range(bx, by, bz)) {
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
      flush_voxel(v, acc, bc, dropped, overflow);
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
  flush_voxel(v, acc, bc, dropped, overflow);
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
    if (rays) atomicAdd(v.ctr + CTR_RAYS, (unsigned long long)rays);
    if (ev_p) atomicAdd(v.ctr + CTR_EV_P, (unsigned long long)ev_p);
    if (ev_q) atomicAdd(v.ctr + CTR_EV_Q, (unsigned long long)ev_q);
    if (dropped) atomicAdd(v.ctr + CTR_DROPPED, (unsigned long long)dropped);
    if (overflow) atomicAdd(v.ctr + CTR_OVERFLOW, (unsigned long long)overflow);
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

template <bool BOUNDED>
__global__ void __launch_bounds__(256)
    integrate_points_f32_kernel(VolumeDev v, const float *__restrict__ pts,
                                const double *__restrict__ conf,
                                const double *__restrict__ conf_q, float ox, float oy, float oz,
                                long long m, float step, MarchConst mc) {
  __shared__ WarpCache wcs[8];
  WarpCache &wc = wcs[threadIdx.x >> 5];
  warp_cache_init(wc);
  long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  Tally tl = {0, 0, 0, 0, 0};
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
  march_warp<BOUNDED>(v, wc, mc, have, r, c, cq, conf_q != nullptr, tl);
  commit_tally(v, tl);
}

// ---------------------------------------------------------- keyframe entry
// Sim3 act = s * R(q) x + t with pose = [t(3), q(xyzw), s]; float32 operation
// order of oracle/tsdf_oracle.c:364-376 (restating the reference's own
// actSO3/actSim3, backend/src/gn_kernels.cu:195-221).
__device__ __forceinline__ void sim3_act(const float *__restrict__ P, float x0, float x1, float x2,
                                         float &y0, float &y1, float &y2) {
  const float q0 = P[3], q1 = P[4], q2 = P[5], q3 = P[6], s = P[7];
  float uv0 = __fmul_rn(2.0f, __fsub_rn(__fmul_rn(q1, x2), __fmul_rn(q2, x1)));
  float uv1 = __fmul_rn(2.0f, __fsub_rn(__fmul_rn(q2, x0), __fmul_rn(q0, x2)));
  float uv2 = __fmul_rn(2.0f, __fsub_rn(__fmul_rn(q0, x1), __fmul_rn(q1, x0)));
  float a0 = __fadd_rn(__fadd_rn(x0, __fmul_rn(q3, uv0)),
                       __fsub_rn(__fmul_rn(q1, uv2), __fmul_rn(q2, uv1)));
  float a1 = __fadd_rn(__fadd_rn(x1, __fmul_rn(q3, uv1)),
                       __fsub_rn(__fmul_rn(q2, uv0), __fmul_rn(q0, uv2)));
  float a2 = __fadd_rn(__fadd_rn(x2, __fmul_rn(q3, uv2)),
                       __fsub_rn(__fmul_rn(q0, uv1), __fmul_rn(q1, uv0)));
  y0 = __fadd_rn(__fmul_rn(a0, s), P[0]);
  y1 = __fadd_rn(__fmul_rn(a1, s), P[1]);
  y2 = __fadd_rn(__fmul_rn(a2, s), P[2]);
}

template <bool BOUNDED>
__global__ void __launch_bounds__(256)
    integrate_keyframes_kernel(VolumeDev v, const float *__restrict__ X,
                               const float *__restrict__ C, const float *__restrict__ qg,
                               const float *__restrict__ poses, int H, int W, int gh, int gw,
                               float min_conf, float step, MarchConst mc) {
  // CTA = 32 x 8 pixel tile, warp = 8 x 4 sub-tile: rays of a warp stay close
  // in the volume, so their reductions share 32-byte sectors.
  __shared__ WarpCache wcs[8];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  WarpCache &wc = wcs[warp];
  warp_cache_init(wc);
  const int kf = blockIdx.z;
  const int px = blockIdx.x * 32 + (warp & 3) * 8 + (lane & 7);
  const int py = blockIdx.y * 8 + (warp >> 2) * 4 + (lane >> 3);
  const long long npix = (long long)H * W;
  Tally tl = {0, 0, 0, 0, 0};
  Ray<float> r = {};
  bool have = false;
  float c = 0.f, cq = 0.f;
  if (px < W && py < H) {
    const long long pix = (long long)py * W + px;
    const float *P = poses + 8 * kf;
    c = C[kf * npix + pix];
    if (c > min_conf) { // global_manager.py:92
      const float *x = X + (kf * npix + pix) * 3;
      float wx, wy, wz;
      sim3_act(P, x[0], x[1], x[2], wx, wy, wz);
      have = ray_setup<float>(r, wx, wy, wz, P[0], P[1], P[2], step, mc.trunc);
      if (have) {
        tl.rays = 1;
        float q = 1.0f;
        if (qg) {
          const int ch = H / gh, cw = W / gw; // uniform; power-of-two cells avoid the division
          int gy = (ch & (ch - 1)) ? py / ch : py >> (31 - __clz(ch));
          int gx = (cw & (cw - 1)) ? px / cw : px >> (31 - __clz(cw));
          gy = min(gy, gh - 1), gx = min(gx, gw - 1);
          q = qg[((long long)kf * gh + gy) * gw + gx];
        }
        cq = c * q;
      }
    }
  }
  march_warp<BOUNDED>(v, wc, mc, have, r, c, cq, true, tl);
  commit_tally(v, tl);
}

// ------------------------------------------------------------------ C ABI
static MarchConst make_consts(const b200tsdf *h) {
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
  const char *e = getenv("B200TSDF_DEBUG");
  mc.dbg = e ? atoi(e) : 0;
  return mc;
}

extern "C" int b200tsdf_integrate_points(b200tsdf_t *h, const void *pts_dev, int dtype,
                                         const double *conf_dev, const double *conf_q_dev,
                                         const void *origin_host, int64_t m, double step_scale,
                                         int64_t *rays_fused_host, void *stream) {
  if (!h || (m > 0 && (!pts_dev || !conf_dev || !origin_host)) || m < 0 ||
      (dtype != B200TSDF_F32 && dtype != B200TSDF_F64)) {
    b200_set_error("b200tsdf_integrate_points: invalid argument");
    return B200TSDF_E_INVALID;
  }
  cudaStream_t st = (cudaStream_t)stream;
  B200_CUDA(cudaSetDevice(h->cfg.device));
  unsigned long long before = 0;
  if (rays_fused_host) {
    B200_CUDA(cudaMemcpyAsync(&before, h->d.ctr + CTR_RAYS, 8, cudaMemcpyDeviceToHost, st));
    B200_CUDA(cudaStreamSynchronize(st));
  }
  if (m > 0) {
    double step_d = h->cfg.voxel_size * step_scale; // global_volume.py:46
    if (step_d < 1.0e-4) step_d = 1.0e-4;
    unsigned grid = (unsigned)((m + 255) / 256);
    if (dtype == B200TSDF_F32) {
      const float *o = (const float *)origin_host;
      if (h->d.bounded)
        integrate_points_f32_kernel<true><<<grid, 256, 0, st>>>(
            h->d, (const float *)pts_dev, conf_dev, conf_q_dev, o[0], o[1], o[2], (long long)m,
            (float)step_d, make_consts(h));
      else
        integrate_points_f32_kernel<false><<<grid, 256, 0, st>>>(
            h->d, (const float *)pts_dev, conf_dev, conf_q_dev, o[0], o[1], o[2], (long long)m,
            (float)step_d, make_consts(h));
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
    B200_CUDA(cudaMemcpyAsync(&after, h->d.ctr + CTR_RAYS, 8, cudaMemcpyDeviceToHost, st));
    B200_CUDA(cudaStreamSynchronize(st));
    *rays_fused_host = (int64_t)(after - before);
  }
  return B200TSDF_OK;
}

extern "C" int b200tsdf_integrate_keyframes(b200tsdf_t *h, const float *X_dev, const float *C_dev,
                                            const float *q_grid_dev, const float *pose_dev,
                                            int32_t B, int32_t H, int32_t W, int32_t gh,
                                            int32_t gw, float min_conf, double step_scale,
                                            void *stream) {
  if (!h || B < 0 || H <= 0 || W <= 0 || (B > 0 && (!X_dev || !C_dev || !pose_dev))) {
    b200_set_error("b200tsdf_integrate_keyframes: invalid argument");
    return B200TSDF_E_INVALID;
  }
  if (q_grid_dev && (gh <= 0 || gw <= 0 || H % gh || W % gw)) {
    b200_set_error("b200tsdf_integrate_keyframes: quality grid %dx%d does not tile %dx%d", gh, gw,
                   H, W);
    return B200TSDF_E_INVALID;
  }
  if (B == 0) return B200TSDF_OK;
  cudaStream_t st = (cudaStream_t)stream;
  B200_CUDA(cudaSetDevice(h->cfg.device));
  double step_d = h->cfg.voxel_size * step_scale;
  if (step_d < 1.0e-4) step_d = 1.0e-4;
  long long npix = (long long)H * W;
  dim3 grid((unsigned)((W + 31) / 32), (unsigned)((H + 7) / 8), (unsigned)B);
  if (h->d.bounded)
    integrate_keyframes_kernel<true><<<grid, 256, 0, st>>>(
        h->d, X_dev, C_dev, q_grid_dev, pose_dev, H, W, gh, gw, min_conf, (float)step_d,
        make_consts(h));
  else
    integrate_keyframes_kernel<false><<<grid, 256, 0, st>>>(
        h->d, X_dev, C_dev, q_grid_dev, pose_dev, H, W, gh, gw, min_conf, (float)step_d,
        make_consts(h));
  B200_CUDA(cudaGetLastError());
  return B200TSDF_OK;
}

int b200_ensure_stage(b200tsdf *h, size_t bytes);

extern "C" int b200tsdf_integrate_keyframes_host(b200tsdf_t *h, const float *X_host,
                                                 const float *C_host, const float *q_grid_host,
                                                 const float *pose_host, int32_t B, int32_t H,
                                                 int32_t W, int32_t gh, int32_t gw,
                                                 float min_conf, double step_scale,
                                                 void *stream) {
  if (!h || B < 0 || H <= 0 || W <= 0 || (B > 0 && (!X_host || !C_host || !pose_host))) {
    b200_set_error("b200tsdf_integrate_keyframes_host: invalid argument");
    return B200TSDF_E_INVALID;
  }
  if (B == 0) return B200TSDF_OK;
  cudaStream_t st = (cudaStream_t)stream;
  B200_CUDA(cudaSetDevice(h->cfg.device));
  size_t npix = (size_t)H * W;
  size_t bx = (size_t)B * npix * 3 * sizeof(float), bc = (size_t)B * npix * sizeof(float);
  size_t bq = q_grid_host ? (size_t)B * gh * gw * sizeof(float) : 0;
  size_t bp = (size_t)B * 8 * sizeof(float);
  auto al = [](size_t x) { return (x + 255) & ~(size_t)255; };
  size_t total = al(bx) + al(bc) + al(bq) + al(bp);
  int rc = b200_ensure_stage(h, total);
  if (rc) return rc;
  char *base = (char *)h->stage_dev;
  float *dX = (float *)base;
  float *dC = (float *)(base + al(bx));
  float *dQ = (float *)(base + al(bx) + al(bc));
  float *dP = (float *)(base + al(bx) + al(bc) + al(bq));
  B200_CUDA(cudaMemcpyAsync(dX, X_host, bx, cudaMemcpyHostToDevice, st));
  B200_CUDA(cudaMemcpyAsync(dC, C_host, bc, cudaMemcpyHostToDevice, st));
  if (bq) B200_CUDA(cudaMemcpyAsync(dQ, q_grid_host, bq, cudaMemcpyHostToDevice, st));
  B200_CUDA(cudaMemcpyAsync(dP, pose_host, bp, cudaMemcpyHostToDevice, st));
  return b200tsdf_integrate_keyframes(h, dX, dC, bq ? dQ : nullptr, dP, B, H, W, gh, gw, min_conf,
                                      step_scale, st);
}
