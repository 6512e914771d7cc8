/*
 * oracle/tsdf_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C CPU restatement of the reference's global TSDF path
 * (mast3r_slam/tsdf/global_volume.py).  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load this file's
 * library; the product (libb200tsdf.so) never does.
 *
 * Parity status: PINNED for the global volume -- tests/test_oracle_golden.py
 * checks this file bit-for-bit (keys, per-voxel update counts, tsdf, weight)
 * against fixtures produced by running the unmodified reference
 * TSDFVolume here (tests/golden/make_golden.py).
 *
 * Every function cites the reference lines it follows.  The arithmetic
 * types follow what numpy >= 2 (NEP 50) does with the dtypes the SLAM loop
 * feeds (global_manager.py:99-106: float32 points / origin, float64
 * confidences): all ray / sample / key math is float32, only the weight and
 * the running average are float64.  A second entry point restates the
 * all-float64 case.
 *
 * Build: gcc -O2 -ffp-contract=off -fno-fast-math (no FMA contraction, so the
 * float32 sequence is exactly the one numpy executes).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
  int64_t k[3];
  double tsdf;
  double weight;
  uint32_t count; /* number of _update_voxel hits (not in the reference; an
                     integer the parity tests compare bit-exactly) */
  uint8_t used;
  uint8_t t_f32; /* tsdf still holds the np.float32 scalar of the first insert
                    (global_volume.py:80): numpy then does float32 arithmetic
                    on it in _estimate_gradient (:122) */
} orc_voxel;

typedef struct {
  double voxel_size, truncation, max_weight, min_weight;
  orc_voxel *tab;
  uint64_t cap; /* power of two */
  uint64_t n;
  uint64_t events; /* total _update_voxel calls that passed weight > 0 */
} orc_volume;

static uint64_t orc_hash(const int64_t k[3]) {
  uint64_t h = (uint64_t)k[0] * 0x9E3779B97F4A7C15ull;
  h ^= ((uint64_t)k[1] + 0x7F4A7C15ull) * 0xC2B2AE3D27D4EB4Full;
  h = (h << 31) | (h >> 33);
  h ^= ((uint64_t)k[2] + 0x165667B1ull) * 0xD6E8FEB86659FD93ull;
  h ^= h >> 29;
  h *= 0xBF58476D1CE4E5B9ull;
  h ^= h >> 32;
  return h;
}

/* global_volume.py:18-30  TSDFVolume.__init__ */
orc_volume *orc_create(double voxel_size, double truncation, double max_weight,
                       double min_weight) {
  orc_volume *v = (orc_volume *)calloc(1, sizeof(orc_volume));
  v->voxel_size = voxel_size;
  v->truncation = truncation;
  v->max_weight = max_weight;
  v->min_weight = min_weight;
  v->cap = 1u << 16;
  v->tab = (orc_voxel *)calloc(v->cap, sizeof(orc_voxel));
  return v;
}

void orc_destroy(orc_volume *v) {
  if (!v) return;
  free(v->tab);
  free(v);
}

static orc_voxel *orc_find(orc_volume *v, const int64_t k[3], int insert);

static void orc_grow(orc_volume *v) {
  orc_voxel *old = v->tab;
  uint64_t oldcap = v->cap;
  v->cap *= 2;
  v->tab = (orc_voxel *)calloc(v->cap, sizeof(orc_voxel));
  v->n = 0;
  for (uint64_t i = 0; i < oldcap; ++i)
    if (old[i].used) {
      orc_voxel *d = orc_find(v, old[i].k, 1);
      *d = old[i];
    }
  free(old);
}

static orc_voxel *orc_find(orc_volume *v, const int64_t k[3], int insert) {
  uint64_t m = v->cap - 1, i = orc_hash(k) & m;
  for (;;) {
    orc_voxel *e = &v->tab[i];
    if (!e->used) {
      if (!insert) return NULL;
      e->used = 1;
      e->k[0] = k[0];
      e->k[1] = k[1];
      e->k[2] = k[2];
      e->count = 0;
      v->n++;
      return e;
    }
    if (e->k[0] == k[0] && e->k[1] == k[1] && e->k[2] == k[2]) return e;
    i = (i + 1) & m;
  }
}

/* global_volume.py:74-88  TSDFVolume._update_voxel (key already computed by
 * the caller per :133-134).  Running average divides by the CLAMPED total
 * weight; the first insert stores tsdf / weight verbatim (no clamp). */
static void orc_update_voxel(orc_volume *v, const int64_t key[3], double tsdf,
                             double weight, int t_is_f32) {
  if (weight <= 0.0) return; /* :75 */
  v->events++;
  if ((v->n + 1) * 10 > v->cap * 7) orc_grow(v);
  uint64_t before = v->n;
  orc_voxel *e = orc_find(v, key, 1);
  if (v->n != before) { /* :79-82 first touch */
    e->tsdf = tsdf;
    e->weight = weight;
    e->count = 1;
    e->t_f32 = (uint8_t)t_is_f32;
    return;
  }
  double total = e->weight + weight; /* :84 */
  if (total > v->max_weight) total = v->max_weight;
  double den = total > 1.0e-9 ? total : 1.0e-9; /* :85-87 */
  e->tsdf = (e->tsdf * e->weight + tsdf * weight) / den;
  e->weight = total; /* :88 */
  e->t_f32 = 0;      /* float32 * float64 -> float64 */
  e->count++;
}

/* np.linalg.norm of a float32 3-vector = sqrt(x.dot(x)); x86-64 OpenBLAS sdot
 * rounds each product to float32 and accumulates them in a double, rounding
 * the sum back to float32 (probed here against numpy 2.3.5 / OpenBLAS 0.3.30:
 * 0 mismatches in 20 000 random vectors; plain float32 accumulation differs in
 * ~19 % of them). */
static float orc_norm3f(float x, float y, float z) {
  float px = x * x, py = y * y, pz = z * z;
  double acc = (double)px;
  acc += (double)py;
  acc += (double)pz;
  return sqrtf((float)acc);
}

/* global_volume.py:35-72  TSDFVolume.integrate with float32 points/origin
 * and float64 confidences (the dtypes global_manager.py:99-106 passes), as
 * numpy >= 2 evaluates it: python-float constants are "weak" and take the
 * float32 type of the array operand.  Returns the number of rays fused. */
int64_t orc_integrate_f32(orc_volume *v, const float *pts, const double *conf,
                          const float *origin, int64_t m, double step_scale) {
  if (m == 0) return 0; /* :42-43 */
  double step_d = v->voxel_size * step_scale; /* :46 */
  if (step_d < 1.0e-4) step_d = 1.0e-4;
  const float step = (float)step_d;
  const float trunc = (float)v->truncation; /* :47 */
  const float vs = (float)v->voxel_size;
  const float ox = origin[0], oy = origin[1], oz = origin[2];
  int64_t fused = 0;
  for (int64_t p = 0; p < m; ++p) { /* :53 */
    float rx = pts[3 * p + 0] - ox, ry = pts[3 * p + 1] - oy,
          rz = pts[3 * p + 2] - oz;      /* :54 */
    float len = orc_norm3f(rx, ry, rz); /* :55 */
    if (!isfinite(len) || len < (float)1.0e-4) continue; /* :56-57 */
    float dx = rx / len, dy = ry / len, dz = rz / len; /* :58 */
    float maxd = len + trunc;                          /* :59 */
    float q = maxd / step;                             /* :60 */
    int64_t n = (int64_t)q; /* int() truncates toward zero */
    if (n < 1) n = 1;
    /* :61 np.linspace(0.0, maxd, n) in float32:
     * y[i] = float32(i) * (maxd / float32(n-1)); y[n-1] = maxd; n==1 -> [0] */
    float lstep = 0.0f;
    if (n > 1) lstep = maxd / (float)(n - 1);
    for (int64_t i = 0; i < n; ++i) { /* :63 */
      float dist;
      if (n == 1)
        dist = 0.0f;
      else if (i == n - 1)
        dist = maxd;
      else
        dist = (float)i * lstep;
      float mx = dist * dx, my = dist * dy, mz = dist * dz; /* :64 */
      float sx = ox + mx, sy = oy + my, sz = oz + mz;
      float sdf = len - dist;           /* :65 */
      if (fabsf(sdf) > trunc) continue; /* :66-67 */
      float t = sdf / trunc;            /* :68 */
      if (t < -1.0f) t = -1.0f;
      if (t > 1.0f) t = 1.0f;
      float a = -fabsf(sdf) / trunc; /* :69, float32 quotient */
      double w = conf[p] * exp((double)a);
      int64_t key[3]; /* :133-134 floor(sample / voxel_size) */
      key[0] = (int64_t)floorf(sx / vs);
      key[1] = (int64_t)floorf(sy / vs);
      key[2] = (int64_t)floorf(sz / vs);
      orc_update_voxel(v, key, (double)t, w, 1); /* :70 */
    }
    fused++; /* :71 */
  }
  return fused;
}

/* np.linalg.norm of a float64 3-vector = sqrt(ddot): the x86-64 OpenBLAS
 * ddot tail loop is FMA-contracted, dot = fma(z,z, fma(y,y, x*x)) (probed:
 * 0 mismatches in 3000 random vectors; every non-FMA order mismatches). */
static double orc_norm3d(double x, double y, double z) {
  double acc = x * x;
  acc = fma(y, y, acc);
  acc = fma(z, z, acc);
  return sqrt(acc);
}

/* Same function with every input float64: all intermediates are float64. */
int64_t orc_integrate_f64(orc_volume *v, const double *pts, const double *conf,
                          const double *origin, int64_t m, double step_scale) {
  if (m == 0) return 0;
  double step = v->voxel_size * step_scale;
  if (step < 1.0e-4) step = 1.0e-4;
  const double trunc = v->truncation, vs = v->voxel_size;
  const double ox = origin[0], oy = origin[1], oz = origin[2];
  int64_t fused = 0;
  for (int64_t p = 0; p < m; ++p) {
    double rx = pts[3 * p + 0] - ox, ry = pts[3 * p + 1] - oy,
           rz = pts[3 * p + 2] - oz;
    double len = orc_norm3d(rx, ry, rz);
    if (!isfinite(len) || len < 1.0e-4) continue;
    double dx = rx / len, dy = ry / len, dz = rz / len;
    double maxd = len + trunc;
    int64_t n = (int64_t)(maxd / step);
    if (n < 1) n = 1;
    double lstep = 0.0;
    if (n > 1) lstep = maxd / (double)(n - 1);
    for (int64_t i = 0; i < n; ++i) {
      double dist;
      if (n == 1)
        dist = 0.0;
      else if (i == n - 1)
        dist = maxd;
      else
        dist = (double)i * lstep;
      double mx = dist * dx, my = dist * dy, mz = dist * dz;
      double sx = ox + mx, sy = oy + my, sz = oz + mz;
      double sdf = len - dist;
      if (fabs(sdf) > trunc) continue;
      double t = sdf / trunc;
      if (t < -1.0) t = -1.0;
      if (t > 1.0) t = 1.0;
      double w = conf[p] * exp(-fabs(sdf) / trunc);
      int64_t key[3];
      key[0] = (int64_t)floor(sx / vs);
      key[1] = (int64_t)floor(sy / vs);
      key[2] = (int64_t)floor(sz / vs);
      orc_update_voxel(v, key, t, w, 0);
    }
    fused++;
  }
  return fused;
}

/* ------------------------------------------------------------------ */
/* Inspection                                                          */
/* ------------------------------------------------------------------ */
uint64_t orc_num_voxels(const orc_volume *v) { return v->n; }
uint64_t orc_num_events(const orc_volume *v) { return v->events; }

static int orc_cmp_key(const void *a, const void *b) {
  const orc_voxel *x = *(const orc_voxel *const *)a;
  const orc_voxel *y = *(const orc_voxel *const *)b;
  for (int d = 0; d < 3; ++d) {
    if (x->k[d] < y->k[d]) return -1;
    if (x->k[d] > y->k[d]) return 1;
  }
  return 0;
}

/* Dump every voxel, sorted lexicographically by (ix, iy, iz).  Arrays must
 * hold orc_num_voxels() entries.  Returns the count written. */
uint64_t orc_export(const orc_volume *v, int64_t *keys, double *tsdf,
                    double *weight, uint32_t *count) {
  const orc_voxel **p =
      (const orc_voxel **)malloc(sizeof(orc_voxel *) * (v->n ? v->n : 1));
  uint64_t j = 0;
  for (uint64_t i = 0; i < v->cap; ++i)
    if (v->tab[i].used) p[j++] = &v->tab[i];
  qsort(p, j, sizeof(*p), orc_cmp_key);
  for (uint64_t i = 0; i < j; ++i) {
    keys[3 * i + 0] = p[i]->k[0];
    keys[3 * i + 1] = p[i]->k[1];
    keys[3 * i + 2] = p[i]->k[2];
    tsdf[i] = p[i]->tsdf;
    weight[i] = p[i]->weight;
    count[i] = p[i]->count;
  }
  free(p);
  return j;
}

/* global_volume.py:136-140  TSDFVolume.stats */
void orc_stats(const orc_volume *v, int64_t *valid, int64_t *total) {
  int64_t nv = 0;
  for (uint64_t i = 0; i < v->cap; ++i)
    if (v->tab[i].used && v->tab[i].weight >= v->min_weight) nv++;
  *valid = nv;
  *total = (int64_t)v->n;
}

/* global_volume.py:93-128  TSDFVolume.query + _estimate_gradient.
 * pts are float32 (dtype_f64 == 0) or float64 (== 1); the key uses the same
 * dtype division as _world_to_voxel.  flags bit0 = value present,
 * bit1 = gradient present.  tsdf/grad are float64 like the reference. */
void orc_query(orc_volume *v, const void *pts, int dtype_f64, int64_t m,
               double *tsdf_out, double *grad_out, uint8_t *flags) {
  for (int64_t p = 0; p < m; ++p) {
    int64_t key[3];
    if (dtype_f64) {
      const double *q = (const double *)pts + 3 * p;
      for (int d = 0; d < 3; ++d)
        key[d] = (int64_t)floor(q[d] / v->voxel_size);
    } else {
      const float *q = (const float *)pts + 3 * p;
      const float vs = (float)v->voxel_size;
      for (int d = 0; d < 3; ++d) key[d] = (int64_t)floorf(q[d] / vs);
    }
    flags[p] = 0;
    tsdf_out[p] = 0.0;
    grad_out[3 * p] = grad_out[3 * p + 1] = grad_out[3 * p + 2] = 0.0;
    orc_voxel *c = orc_find(v, key, 0);
    if (!c || c->weight < v->min_weight) continue; /* :98-100 */
    tsdf_out[p] = c->tsdf;
    flags[p] = 1;
    double g[3] = {0, 0, 0};
    int denom = 0;
    for (int ax = 0; ax < 3; ++ax) { /* :112-123 */
      int64_t kp[3] = {key[0], key[1], key[2]}, kn[3] = {key[0], key[1], key[2]};
      kp[ax] += 1;
      kn[ax] -= 1;
      orc_voxel *vp = orc_find(v, kp, 0), *vn = orc_find(v, kn, 0);
      if (!vp || !vn) continue;
      if (vp->weight < v->min_weight || vn->weight < v->min_weight) continue;
      if (vp->t_f32 && vn->t_f32) { /* np.float32 - np.float32, / weak python float */
        float d32 = (float)vp->tsdf - (float)vn->tsdf;
        g[ax] = (double)(d32 / (float)(2.0 * v->voxel_size));
      } else {
        g[ax] = (vp->tsdf - vn->tsdf) / (2.0 * v->voxel_size);
      }
      denom++;
    }
    if (!denom) continue; /* :124-125 */
    /* np.linalg.norm(float64[3]) -> sqrt(ddot), see orc_norm3d */
    double nrm = orc_norm3d(g[0], g[1], g[2]);
    if (nrm < 1.0e-9) continue; /* :126-128 */
    grad_out[3 * p + 0] = g[0] / nrm;
    grad_out[3 * p + 1] = g[1] / nrm;
    grad_out[3 * p + 2] = g[2] / nrm;
    flags[p] = 3;
  }
}

/* ------------------------------------------------------------------ */
/* Keyframe staging: global_manager.py:81-106 minus the unseeded random  */
/* subsample (every pixel with C > min_conf is kept, in pixel order).    */
/* Sim3 act = s * R(q) p + t with pose = [t(3), q(xyzw), s]; the         */
/* quaternion form restates the reference's own actSO3 / actSim3         */
/* (backend/src/gn_kernels.cu:195-221); lietorch itself is absent, so    */
/* the float32 operation order below is this repo's definition           */
/* ("parity unpinned at the lietorch boundary").              */
/* ------------------------------------------------------------------ */
void orc_sim3_act_f32(const float *pose, const float *x, float *y) {
  const float *t = pose, *q = pose + 3;
  const float s = pose[7];
  float uv0 = 2.0f * (q[1] * x[2] - q[2] * x[1]);
  float uv1 = 2.0f * (q[2] * x[0] - q[0] * x[2]);
  float uv2 = 2.0f * (q[0] * x[1] - q[1] * x[0]);
  float y0 = (x[0] + q[3] * uv0) + (q[1] * uv2 - q[2] * uv1);
  float y1 = (x[1] + q[3] * uv1) + (q[2] * uv0 - q[0] * uv2);
  float y2 = (x[2] + q[3] * uv2) + (q[0] * uv1 - q[1] * uv0);
  y[0] = y0 * s + t[0];
  y[1] = y1 * s + t[1];
  y[2] = y2 * s + t[2];
}

/* Stage one keyframe: returns the number of kept points; writes world
 * points (float32), confidences (float64) and, when q_grid != NULL, the
 * quality-scaled confidences conf_q = (double)C * (double)q with q the
 * nearest-neighbour upsample of the gh x gw patch grid.
 * cam_origin = act(0) = t. */
int64_t orc_stage_keyframe(const float *X, const float *C, const float *pose,
                           const float *q_grid, int gh, int gw, int H, int W,
                           float min_conf, float *pts_world, double *conf,
                           double *conf_q, float *cam_origin) {
  int64_t n = 0;
  const int ph = gh > 0 ? H / gh : 1, pw = gw > 0 ? W / gw : 1;
  for (int64_t i = 0; i < (int64_t)H * W; ++i) {
    if (!(C[i] > min_conf)) continue; /* global_manager.py:92 */
    orc_sim3_act_f32(pose, X + 3 * i, pts_world + 3 * n);
    conf[n] = (double)C[i];
    if (conf_q) {
      float q = 1.0f;
      if (q_grid) {
        int y = (int)(i / W), x = (int)(i % W);
        int gy = y / ph, gx = x / pw;
        if (gy >= gh) gy = gh - 1;
        if (gx >= gw) gx = gw - 1;
        q = q_grid[gy * gw + gx];
      }
      conf_q[n] = (double)C[i] * (double)q;
    }
    n++;
  }
  cam_origin[0] = pose[0];
  cam_origin[1] = pose[1];
  cam_origin[2] = pose[2];
  return n;
}
