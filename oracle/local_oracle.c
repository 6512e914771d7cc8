/*
 * oracle/local_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C CPU restatement of the reference's LOCAL (per-ROI, dense) TSDF:
 *   TSDFRefiner._build_tsdf_robust      mast3r_slam/tsdf_refine.py:837-940
 *   TSDFRefiner._extract_surface_safe   mast3r_slam/tsdf_refine.py:942-1021
 *   TSDFRefiner._sample_tsdf_trilinear  mast3r_slam/tsdf_refine.py:1023-1064
 * as torch 2.x evaluates them ON THE CPU (float32 tensor ops, Python float64
 * scalars from .item()).  Probed here against torch 2.11:
 *   - torch.linspace(float32): y[i] = fma(step, i, start) for i < n/2, else
 *     fma(-step, n-1-i, end), step = (end - start) / (n - 1) in float32;
 *   - Tensor.norm() of a float32 3-vector: s = x0*x0; s = fma(x1,x1,s);
 *     s = fma(x2,x2,s); sqrtf(s);
 *   - tensor / python_float and tensor * python_float: the scalar is cast to
 *     float32, true IEEE division / multiplication.
 * Parity status: pinned by tests/golden/local_*.npz, produced by running the
 * unmodified reference methods on CPU tensors (tests/golden/make_golden_local.py)
 * with the ~30-line lietorch.Sim3 stand-in of oracle/ref_harness.py (lietorch
 * itself is absent from the reference tree).
 *
 * Build: gcc -O2 -ffp-contract=off (explicit fmaf where torch fuses).
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

void orc_sim3_act_f32(const float *pose, const float *x, float *y); /* tsdf_oracle.c */

static float norm3_torch(float x, float y, float z) {
  float s = x * x;
  s = fmaf(y, y, s);
  s = fmaf(z, z, s);
  return sqrtf(s);
}

/* torch.linspace(start, end, n) float32, element i */
static float linspace_torch(float start, float end, int n, int i) {
  if (n == 1) return start;
  float step = (end - start) / (float)(n - 1);
  if (i < n / 2) return fmaf(step, (float)i, start);
  return fmaf(-step, (float)(n - 1 - i), end);
}

/* tsdf_refine.py:844-846  grid_dims = clamp(ceil(roi / voxel).long(), max=max_grid_dim) */
void orc_local_dims(const float *xyz_min, const float *xyz_max, double voxel_size, int max_grid_dim,
                    int dims[3]) {
  for (int a = 0; a < 3; ++a) {
    float roi = xyz_max[a] - xyz_min[a];
    long long g = (long long)ceilf(roi / (float)voxel_size);
    if (g > max_grid_dim) g = max_grid_dim;
    dims[a] = (int)g;
  }
}

/* tsdf_refine.py:837-940.  tsdf / weights are (nz, ny, nx) float32, caller
 * allocated with the dims of orc_local_dims.  Returns the number of points
 * fused (fusion_count), or -1 when fewer than 5 points are valid (:877-878,
 * grids stay at their initial 1 / 0). */
int64_t orc_local_build(const float *X_canon, const float *C, const float *pose,
                        const float *xyz_min, const float *xyz_max, int64_t n_pts,
                        double voxel_size, double trunc_dist, int max_grid_dim,
                        double min_confidence, float *tsdf, float *weights, int dims[3]) {
  orc_local_dims(xyz_min, xyz_max, voxel_size, max_grid_dim, dims);
  const int nx = dims[0], ny = dims[1], nz = dims[2];
  const int64_t nvox = (int64_t)nx * ny * nz;
  for (int64_t i = 0; i < nvox; ++i) tsdf[i] = 1.0f, weights[i] = 0.0f; /* :850-851 */
  const float o[3] = {pose[0], pose[1], pose[2]}; /* :859-860 matrix()[:3,3] = t */
  float roi[3], av[3];
  const int nd[3] = {nx, ny, nz};
  for (int a = 0; a < 3; ++a) {
    roi[a] = xyz_max[a] - xyz_min[a];
    av[a] = roi[a] / (float)nd[a]; /* :880 actual_voxel */
  }
  /* :866-876 count valid points first (the < 5 early return) */
  int64_t n_valid = 0;
  for (int pass = 0; pass < 2; ++pass) {
    int64_t fused = 0;
    for (int64_t p = 0; p < n_pts; ++p) {
      float w[3];
      orc_sim3_act_f32(pose, X_canon + 3 * p, w); /* :863 */
      int ok = isfinite(w[0]) && isfinite(w[1]) && isfinite(w[2]);
      for (int a = 0; a < 3 && ok; ++a) ok = w[a] >= xyz_min[a] && w[a] <= xyz_max[a];
      ok = ok && (C[p] > (float)min_confidence);
      if (!ok) continue;
      if (pass == 0) {
        n_valid++;
        continue;
      }
      const double conf = (double)C[p];
      float rv[3] = {w[0] - o[0], w[1] - o[1], w[2] - o[2]}; /* :889 */
      const float len_f = norm3_torch(rv[0], rv[1], rv[2]);  /* :890 */
      const double len = (double)len_f;
      const float den = (float)(len + 1.0e-8); /* :891 */
      float dir[3] = {rv[0] / den, rv[1] / den, rv[2] / den};
      if (len < 0.05) continue; /* :893-894 */
      double t_start = len - trunc_dist * 2.0; /* :897-899 */
      if (t_start < 0.05) t_start = 0.05;
      const double t_end = len + trunc_dist * 2.0;
      int ns = (int)((t_end - t_start) / voxel_size) + 1;
      if (ns > 32) ns = 32;
      for (int i = 0; i < ns; ++i) { /* :901 */
        const float t = linspace_torch((float)t_start, (float)t_end, ns, i);
        float s[3];
        int out = 0;
        for (int a = 0; a < 3; ++a) {
          float m = dir[a] * t;
          s[a] = o[a] + m; /* :903 */
          if (s[a] < xyz_min[a] || s[a] > xyz_max[a]) out = 1; /* :906 */
        }
        if (out) continue;
        int g[3];
        for (int a = 0; a < 3; ++a) { /* :910-913 */
          float gp = (s[a] - xyz_min[a]) / av[a];
          float hi = (float)(nd[a] - 1);
          gp = gp < 0.0f ? 0.0f : (gp > hi ? hi : gp);
          g[a] = (int)gp;
        }
        double sdf = (len - (double)t) / trunc_dist; /* :916-917 */
        if (sdf < -1.0) sdf = -1.0;
        if (sdf > 1.0) sdf = 1.0;
        double dw = 1.0 - fabs(sdf); /* :920-921 */
        if (dw < 0.0) dw = 0.0;
        const double weight = conf * dw;
        const int64_t idx = ((int64_t)g[2] * ny + g[1]) * nx + g[0];
        const double old_w = (double)weights[idx]; /* :924-925 */
        const double new_w = old_w + weight;
        if (new_w > 1.0e-6) { /* :927-930 */
          const double old_sdf = (double)tsdf[idx];
          tsdf[idx] = (float)((old_sdf * old_w + sdf * weight) / new_w);
          weights[idx] = (float)new_w;
        }
      }
      fused++;
    }
    if (pass == 0 && n_valid < 5) return -1; /* :877-878 */
    if (pass == 1) return fused;
  }
  return 0;
}

/* tsdf_refine.py:1023-1064 */
static double trilinear(const float *vol, int nx, int ny, int nz, float x, float y, float z) {
  float hx = (float)(nx - 1), hy = (float)(ny - 1), hz = (float)(nz - 1);
  x = x < 0.0f ? 0.0f : (x > hx ? hx : x);
  y = y < 0.0f ? 0.0f : (y > hy ? hy : y);
  z = z < 0.0f ? 0.0f : (z > hz ? hz : z);
  int x0 = (int)floorf(x), y0 = (int)floorf(y), z0 = (int)floorf(z);
  int x1 = x0 + 1 < nx - 1 ? x0 + 1 : nx - 1;
  int y1 = y0 + 1 < ny - 1 ? y0 + 1 : ny - 1;
  int z1 = z0 + 1 < nz - 1 ? z0 + 1 : nz - 1;
  double xd = (double)(x - (float)x0), yd = (double)(y - (float)y0), zd = (double)(z - (float)z0);
#define V(zz, yy, xx) ((double)vol[((int64_t)(zz) * ny + (yy)) * nx + (xx)])
  double c00 = V(z0, y0, x0) * (1 - xd) + V(z0, y0, x1) * xd;
  double c01 = V(z0, y1, x0) * (1 - xd) + V(z0, y1, x1) * xd;
  double c10 = V(z1, y0, x0) * (1 - xd) + V(z1, y0, x1) * xd;
  double c11 = V(z1, y1, x0) * (1 - xd) + V(z1, y1, x1) * xd;
#undef V
  double c0 = c00 * (1 - yd) + c01 * yd;
  double c1 = c10 * (1 - yd) + c11 * yd;
  return c0 * (1 - zd) + c1 * zd;
}

/* tsdf_refine.py:942-1021 for an explicit, ordered pixel list (the
 * reference's torch.randperm choice is made by the caller).  X_refined must
 * be a copy of X_original; hits has m entries (one per listed pixel).
 * Returns the number of hits. */
int64_t orc_local_raycast(const float *tsdf, const int dims[3], const float *xyz_min,
                          const float *xyz_max, const float *X_original, const int64_t *pix,
                          int64_t m, int n_samples, double max_displacement, float *X_refined,
                          uint8_t *hits) {
  const int nx = dims[0], ny = dims[1], nz = dims[2];
  const int nd[3] = {nx, ny, nz};
  float av[3];
  for (int a = 0; a < 3; ++a) av[a] = (xyz_max[a] - xyz_min[a]) / (float)nd[a]; /* :958 */
  int64_t n_hits = 0;
  for (int64_t q = 0; q < m; ++q) {
    hits[q] = 0;
    const float *P = X_original + 3 * pix[q];
    const double depth = (double)P[2]; /* :970 */
    if (depth < 0.05) continue;
    double t_start = depth - 0.1; /* :976-978 */
    if (t_start < 0.05) t_start = 0.05;
    const double t_end = depth + 0.1;
    int have_prev = 0;
    double prev_sdf = 0.0;
    float prev_t = 0.0f;
    for (int j = 0; j < n_samples; ++j) {
      const float t = linspace_torch((float)t_start, (float)t_end, n_samples, j);
      const float ratio = t / (float)depth; /* :986 */
      float s[3];
      int out = 0;
      for (int a = 0; a < 3; ++a) {
        s[a] = P[a] * ratio;
        if (s[a] < xyz_min[a] || s[a] > xyz_max[a]) out = 1;
      }
      if (out) { /* :989-991 */
        have_prev = 0;
        continue;
      }
      float gp[3];
      for (int a = 0; a < 3; ++a) gp[a] = (s[a] - xyz_min[a]) / av[a]; /* :994 */
      const double sdf = trilinear(tsdf, nx, ny, nz, gp[0], gp[1], gp[2]);
      if (have_prev && prev_sdf * sdf < 0) { /* :998-1000 */
        const double alpha = fabs(prev_sdf) / (fabs(prev_sdf) + fabs(sdf) + 1.0e-8);
        const float dt = t - prev_t;
        const float ts = prev_t + (float)alpha * dt; /* :1002 */
        const float rs = ts / (float)depth;
        float sp[3] = {P[0] * rs, P[1] * rs, P[2] * rs}; /* :1004 */
        const double disp = (double)norm3_torch(sp[0] - P[0], sp[1] - P[1], sp[2] - P[2]);
        if (disp <= max_displacement) { /* :1008-1011 */
          memcpy(X_refined + 3 * pix[q], sp, sizeof(sp));
          hits[q] = 1;
          n_hits++;
        }
        break; /* :1013 */
      }
      prev_sdf = sdf; /* :1015-1016 */
      prev_t = t;
      have_prev = 1;
    }
  }
  return n_hits;
}
