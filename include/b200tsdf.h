/*
 * b200tsdf.h -- C ABI of libb200tsdf.so, the sm_100a replacement for the
 * dual-TSDF fusion hot path of MASt3R-SLAM-Quality-dualTSDF.
 *
 * This is the drop-in boundary: plain pointers and sizes, no torch types.
 * Every entry point cites the reference interface it replaces (paths relative
 * to the reference tree).  The reference-side binding (ctypes) is shown in
 * INTEGRATION.md and implemented in mast3r_slam_quality_dualtsdf_b200/_ffi.py.
 *
 * Conventions
 *  - every function returns 0 on success, <0 on error (B200TSDF_E_*); the
 *    message of the last error of the calling thread: b200tsdf_last_error().
 *  - pointers named *_dev are DEVICE pointers borrowed for the duration of the
 *    stream-ordered call; *_host are host pointers.  `stream` is a
 *    cudaStream_t passed as void* (NULL = legacy default stream).
 *  - calls are asynchronous w.r.t. the host unless they return host data.
 *  - the handle owns all volume memory (voxel-block hash + block payload).
 *  - there is NO CPU fallback: every compute call needs a CUDA device.
 */
#ifndef B200TSDF_H_
#define B200TSDF_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200TSDF_ABI_VERSION 4

enum {
  B200TSDF_OK = 0,
  B200TSDF_E_INVALID = -1,  /* bad argument */
  B200TSDF_E_CUDA = -2,     /* CUDA runtime error (message has the detail) */
  B200TSDF_E_CAPACITY = -3, /* voxel-block hash full: raise capacity_log2 */
  B200TSDF_E_NOMEM = -4
};

enum { B200TSDF_F32 = 0, B200TSDF_F64 = 1 };
#define B200TSDF_MAX_SEGMENTS 16 /* source ranks of one keyframe batch (one NVSwitch box: 8) */

typedef struct b200tsdf b200tsdf_t;

/* Replaces TSDFVolume.__init__ (mast3r_slam/tsdf/global_volume.py:18-30) and
 * the tsdf_global config keys (config/base.yaml:110-125). */
typedef struct {
  double voxel_size;  /* tsdf_global.voxel_size   */
  double truncation;  /* tsdf_global.trunc_dist   */
  double max_weight;  /* tsdf_global.max_weight   */
  double min_weight;  /* tsdf_global.min_tsdf_weight */
  /* voxel-block hash: 2^capacity_log2 slots, each slot owns one 8x8x8 voxel
   * block (payload indexed by slot).  Keep load factor <= ~0.6. */
  int32_t capacity_log2;
  /* optional bound in integer VOXEL keys (key = floor(p / voxel_size)):
   * updates whose key is outside [key_min, key_max) are dropped.  The z range
   * doubles as the z-slab of a rank in multi-GPU runs.  bounded = 0: the
   * volume is unbounded like the reference dict. */
  int32_t bounded;
  int64_t key_min[3];
  int64_t key_max[3];
  int32_t device; /* CUDA device ordinal */
  int32_t reserved;
  /* multi-GPU partition: the world-z axis is cut into slabs of
   * slab_width voxel layers (a power of two >= 8, measured from key_min[2] for a
   * bounded volume, from 0 otherwise); slab s belongs to rank s mod
   * slab_count.  This handle fuses only the voxels of rank slab_index and
   * culls rays whose +-truncation band misses its slabs.  slab_count <= 1:
   * the handle owns everything.  slab_count = R with slab_width = dim/R gives
   * R contiguous z-slabs; a smaller width deals slabs round-robin (balanced
   * when the surface is not uniform in z). */
  int32_t slab_count;
  int32_t slab_index;
  int64_t slab_width;
} b200tsdf_config;

const char *b200tsdf_last_error(void);
int b200tsdf_abi_version(void);

int b200tsdf_create(const b200tsdf_config *cfg, b200tsdf_t **out);
int b200tsdf_destroy(b200tsdf_t *h);
/* Forget every voxel (keeps the allocation). */
int b200tsdf_reset(b200tsdf_t *h, void *stream);

/* Replaces TSDFVolume.integrate (global_volume.py:35-72) with the
 * _update_voxel fold (:74-88) for BOTH volumes in one pass.
 *  pts_dev    M x 3 world points, dtype B200TSDF_F32 or _F64 (the arithmetic
 *             follows the dtype exactly as numpy >= 2 does in the reference)
 *  conf_dev   M float64 confidences of the primary volume
 *  conf_q_dev M float64 confidences of the quality-weighted volume (C*q), or
 *             NULL: the second volume is then left untouched
 *  origin_host 3 values of pts dtype (camera centre)
 *  rays_fused_host (nullable) receives the reference's return value; passing
 *             it makes the call synchronise the stream. */
int b200tsdf_integrate_points(b200tsdf_t *h, const void *pts_dev, int dtype,
                              const double *conf_dev, const double *conf_q_dev,
                              const void *origin_host, int64_t m,
                              double step_scale, int64_t *rays_fused_host,
                              void *stream);

/* Replaces TSDFGlobalIntegrator._integrate_single
 * (mast3r_slam/tsdf/global_manager.py:81-106) for a batch of B keyframes:
 * staging (confidence filter, Sim3 act, quality upsample), voxel-block
 * allocation and the dual fold in one launch.  No random subsample: every
 * pixel with C > min_conf is fused, in pixel order.
 *  X_dev     B x H*W x 3 float32 camera-frame pointmaps (frame.py:242)
 *  C_dev     B x H*W     float32 summed confidences     (frame.py:243)
 *  q_grid_dev B x gh x gw float32 per-patch quality weights, or NULL (q = 1)
 *  pose_dev  B x 8 float32 Sim3 [t(3), q(xyzw), s]       (frame.py:241)
 *  H % gh == 0 and W % gw == 0 are required when q_grid_dev != NULL. */
int b200tsdf_integrate_keyframes(b200tsdf_t *h, const float *X_dev,
                                 const float *C_dev, const float *q_grid_dev,
                                 const float *pose_dev, int32_t B, int32_t H,
                                 int32_t W, int32_t gh, int32_t gw,
                                 float min_conf, double step_scale,
                                 void *stream);

/* Exact inverse of b200tsdf_integrate_keyframes for the same inputs (no
 * reference counterpart: the reference re-fuses a keyframe at its new pose
 * WITHOUT removing the old contribution, global_manager.py:71-79; SURVEY 8(f)
 * rank 2).  Sums are subtracted, hit counters decremented exactly; a voxel whose
 * counter returns to 0 reads as absent again.  The keyframe must have been
 * integrated with these very tensors and pose. */
int b200tsdf_deintegrate_keyframes(b200tsdf_t *h, const float *X_dev,
                                   const float *C_dev, const float *q_grid_dev,
                                   const float *pose_dev, int32_t B, int32_t H,
                                   int32_t W, int32_t gh, int32_t gw,
                                   float min_conf, double step_scale,
                                   void *stream);

/* General form of the two calls above (same reference interface,
 * global_manager.py:81-106) for a batch that is NOT one contiguous array:
 *  - packed multi-GPU ingest: the batch is n = ceil(B / seg_kf)
 *    segments, one per source rank, laid out exactly as the all-gather of the
 *    ranks' packed shares delivers them: segment s starts seg_stride floats
 *    after segment s-1 for ALL FOUR arrays (X_dev, C_dev, q_grid_dev, pose_dev
 *    point into segment 0; inside a segment each array is contiguous over its
 *    seg_kf keyframes).  seg_stride = 0: one contiguous batch.
 *  - negate = 1: b200tsdf_deintegrate_keyframes.
 *  - max_points > 0: tsdf_global.max_points_per_kf (global_manager.py:94-99,
 *    config/base.yaml:116): of the pixels with C > min_conf at most max_points
 *    per keyframe are fused.  The reference draws them with an unseeded
 *    torch.randperm; here the choice is the max_points pixels with the smallest
 *    value of a seeded bijective 32-bit hash of the pixel index (salted with
 *    `seed` and the keyframe's id from kf_ids_dev, or its batch position when
 *    kf_ids_dev is NULL), so it is reproducible and has exactly the reference's
 *    sample count min(n_valid, max_points).  max_points <= 0: every valid pixel.
 */
typedef struct {
  const float *X_dev, *C_dev, *q_grid_dev, *pose_dev;
  int32_t B, H, W, gh, gw;
  int32_t seg_kf;      /* keyframes per segment (ignored when seg_stride == 0) */
  int64_t seg_stride;  /* floats between consecutive segments; 0 = contiguous  */
  float min_conf;
  int32_t negate;
  double step_scale;
  int32_t max_points;
  int32_t reserved;
  uint64_t seed;
  const int64_t *kf_ids_dev; /* nullable, B entries */
} b200tsdf_kf_batch;
int b200tsdf_integrate_keyframes_ex(b200tsdf_t *h, const b200tsdf_kf_batch *batch,
                                    void *stream);

/* ---- multi-GPU peer pull (no reference counterpart) ------------
 * A z-slab rank needs only the 32 x 8 pixel tiles of a batch whose +-truncation
 * band reaches its slabs (~1.4 / N of them).  Instead of broadcasting 3.1 MB per
 * keyframe to every rank, the rank that holds a keyframe (its OWNER) describes it
 * by a meta record -- per tile the conservative z range of its bands in voxel
 * layers and its ray count, the pose, the quality grid: ~28 KB -- and the ranks
 * all-gather the records (one small collective).  b200tsdf_integrate_keyframes_pull
 * then lists the tiles that reach this handle's slabs, accounts for the rays of the
 * others (counted + culled, exactly as the broadcast path does), and its fuse
 * kernel TMA-loads the listed tiles straight from the owner's HBM over NVLink:
 * X_seg_dev[s] / C_seg_dev[s] are the bases of source rank s's seg_kf pointmaps /
 * confidences in THIS process's address space (own memory, or a peer buffer mapped
 * with b200tsdf_ipc_open).  meta_dev: n_seg gathered records, meta_stride 32-bit
 * words apart, each of b200tsdf_keyframe_meta_words(seg_kf, H, W, gh, gw) words as
 * b200tsdf_keyframe_meta writes them.  Bounded volumes with a block directory
 * only; no max_points (the owner would have to ship its selection). */
typedef struct {
  const float *X_seg_dev[B200TSDF_MAX_SEGMENTS];
  const float *C_seg_dev[B200TSDF_MAX_SEGMENTS];
  const uint32_t *meta_dev;
  int64_t meta_stride;
  int32_t n_seg, seg_kf, H, W, gh, gw;
  float min_conf;
  int32_t negate;
  double step_scale;
} b200tsdf_kf_pull;
int64_t b200tsdf_keyframe_meta_words(int32_t seg_kf, int32_t H, int32_t W, int32_t gh, int32_t gw);
int b200tsdf_keyframe_meta(b200tsdf_t *h, const float *X_dev, const float *C_dev,
                           const float *q_grid_dev, const float *pose_dev, int32_t B, int32_t H,
                           int32_t W, int32_t gh, int32_t gw, float min_conf, uint32_t *meta_dev,
                           void *stream);
int b200tsdf_integrate_keyframes_pull(b200tsdf_t *h, const b200tsdf_kf_pull *pull, void *stream);
/* CUDA IPC plumbing for the peer buffers (one process per GPU).  peer_alloc: a
 * zeroed device buffer as its own allocation plus the 64 bytes of its
 * cudaIpcMemHandle_t (ship them to the other ranks); ipc_open maps a buffer from
 * those bytes (offset_bytes = offset of the data inside the exported allocation, 0
 * for peer_alloc buffers) with peer access enabled; ipc_close takes the pointer
 * ipc_open returned. */
int b200tsdf_peer_alloc(int32_t device, int64_t bytes, void **ptr_out, void *handle64_out);
int b200tsdf_peer_free(int32_t device, void *ptr);
int b200tsdf_ipc_open(int32_t device, const void *handle64, int64_t offset_bytes, void **ptr_out);
int b200tsdf_ipc_close(int32_t device, void *ptr, int64_t offset_bytes);

/* Same call with HOST buffers (pinned or pageable): copies H2D in chunks on an
 * internal copy stream into a staging area while the chunks already there are
 * fused on `stream`.  Returns once every copy has completed -- the host buffers
 * may be reused or freed right away -- while the fuse itself stays queued on
 * `stream`.  This is the entry the reference-facing plugin uses when keyframes
 * live on the host.  B may be a whole trajectory: quality grids and poses cross
 * in one copy each, pointmaps / confidences in chunks of 40 keyframes (env
 * B200TSDF_H2D_CHUNK_KF; at least 4 chunks) through a ring of three staging
 * slots, one fuse launch per chunk, so the staging area does not grow with B.
 * One call per pass keeps the link busy (54 GB/s of this box's 55.4); a call per
 * 40 keyframes idles it during each call's final wait (50 GB/s). */
int b200tsdf_integrate_keyframes_host(b200tsdf_t *h, const float *X_host,
                                      const float *C_host,
                                      const float *q_grid_host,
                                      const float *pose_host, int32_t B,
                                      int32_t H, int32_t W, int32_t gh,
                                      int32_t gw, float min_conf,
                                      double step_scale, void *stream);

/* Counters since create/reset (synchronises the stream).  Returns
 * B200TSDF_E_CAPACITY (with out_host filled) when the voxel-block hash
 * overflowed, i.e. updates were lost: out[5] > 0.
 *  out[0] rays fused        out[1] voxel updates, primary volume
 *  out[2] voxel updates, quality volume   out[3] allocated voxel blocks
 *  out[4] updates dropped (outside bounds / foreign slab)
 *  out[5] hash overflow events   out[6] rays culled (band misses this
 *  rank's slabs)   out[7] reserved
 * out[1], out[2] count APPLIED updates only, so the sum over the ranks of a
 * sharded volume equals the single-volume count. */
int b200tsdf_counters(b200tsdf_t *h, int64_t out_host[8], void *stream);

/* Replaces TSDFVolume.stats (global_volume.py:136-140); which = 0 primary,
 * 1 quality volume.  Synchronises. */
int b200tsdf_stats(b200tsdf_t *h, int which, int64_t *valid_host,
                   int64_t *total_host, void *stream);

/* Replaces TSDFVolume.query + _estimate_gradient (global_volume.py:93-128),
 * batched.  flags bit0 = value present, bit1 = gradient present. */
int b200tsdf_query(b200tsdf_t *h, int which, const void *pts_dev, int dtype,
                   int64_t m, double *tsdf_dev, double *grad_dev,
                   uint8_t *flags_dev, void *stream);

/* Dump every voxel of one volume (parity tests, checkpoints): two-call
 * protocol -- cap = 0 returns the count in *n_host.  Order is unspecified;
 * keys are integer voxel coordinates like the reference dict keys.
 * tsdf/weight are what the reference VoxelData holds, count = number of
 * _update_voxel hits.  Synchronises. */
int b200tsdf_export_voxels(b200tsdf_t *h, int which, int64_t *keys_dev,
                           float *tsdf_dev, float *weight_dev,
                           uint32_t *count_dev, int64_t cap, int64_t *n_host,
                           void *stream);

/* Block keys (integer block coordinates = voxel key >> 3) of all allocated
 * voxel blocks; same two-call protocol. */
int b200tsdf_export_blocks(b200tsdf_t *h, int64_t *bkeys_dev, int64_t cap,
                           int64_t *n_host, void *stream);

/* Marching cubes over the voxel blocks (no reference counterpart; SURVEY 8(a)
 * A11).  A cube is emitted only if all 8 corners hold weight >= min_weight.
 * Two-call protocol: verts_dev == NULL returns the triangle count only.
 * verts_dev receives 3 vertices (9 floats) per triangle. */
int b200tsdf_extract_mesh(b200tsdf_t *h, int which, float *verts_dev,
                          int64_t tri_cap, int64_t *n_tris_host, void *stream);

/* The same pass without the host round trip: the triangle count stays on the
 * device (n_tris_dev, 8 bytes, zeroed by the call), nothing is synchronised, so
 * extractions can be queued back to back or feed a device consumer (the weld).
 * The count can exceed tri_cap: triangles beyond it were not written. */
int b200tsdf_extract_mesh_async(b200tsdf_t *h, int which, float *verts_dev,
                                int64_t tri_cap, uint64_t *n_tris_dev, void *stream);

/* b200tsdf_extract_mesh_async for callers that may not allocate (CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc).  It has NO
 * counterpart in the reference.
 * The neighbour table and the block ticket of the pass live in a
 * file-scope device array of the library (2 MB + 256 B per GPU, held until the
 * process exits) instead of the handle's lazily allocated table, so the call
 * performs no cudaMalloc / cudaFree.  Limit: volumes of at most
 * B200TSDF_STATIC_MESH_MAX_SLOTS hash slots (capacity_log2 <= 16); larger
 * volumes are rejected with B200TSDF_E_INVALID.  All calls on a GPU share the
 * array: calls that overlap on the GPU (other streams, other volumes) race.
 * Same arguments and results as b200tsdf_extract_mesh_async. */
#define B200TSDF_STATIC_MESH_MAX_SLOTS 65536
int b200tsdf_extract_mesh_static(b200tsdf_t *h, int which, float *verts_dev,
                                 int64_t tri_cap, uint64_t *n_tris_dev, void *stream);

/* The same mesh INDEXED: unique vertices + int32 faces.  Vertices
 * that marching cubes computes on a shared cube edge are bit-identical, so the
 * weld is an exact de-duplication, done on the device (csrc/weld.cu: hash table
 * keyed by the 96-bit position, smallest soup index wins, exclusive scan).
 * verts_dev: vert_cap x 3 float32, faces_dev: tri_cap x 3 int32 (either may be
 * NULL / too small: the counts are still returned, so the two-call protocol of
 * b200tsdf_extract_mesh applies to both).  Vertex order = first occurrence in
 * the soup.  Synchronises. */
int b200tsdf_extract_mesh_indexed(b200tsdf_t *h, int which, float *verts_dev, int64_t vert_cap,
                                  int32_t *faces_dev, int64_t tri_cap, int64_t *n_verts_host,
                                  int64_t *n_tris_host, void *stream);
/* The weld alone, for a soup the caller holds (e.g. the gathered per-rank meshes
 * of a sharded volume, whose seam vertices agree bit for bit): soup_dev
 * n_tris x 9 float32; faces_dev n_tris x 3 int32 or NULL. */
int b200tsdf_weld_mesh(int32_t device, const float *soup_dev, int64_t n_tris, float *verts_dev,
                       int64_t vert_cap, int32_t *faces_dev, int64_t *n_verts_host, void *stream);

/* ---- multi-GPU seams (no reference counterpart) ---------------
 * Marching cubes over the top voxel layer of a z-slab needs the first layer of
 * the slab above, which another rank owns.  export_seam_blocks dumps the voxel
 * blocks of the first block layer of every slab this handle owns (two-call
 * protocol: cap = 0 returns the count): bkeys n x 3 int64 block coordinates,
 * vox n x 512 x 4 float32 packed voxels {S_p, W_p, S_q, W_q}, cnt n x 512
 * uint64 hit counters.  The rank that owns the slabs below passes them to
 * import_halo_blocks: the blocks become read-only neighbours for
 * b200tsdf_extract_mesh (found through the block hash) but are never fused
 * into, meshed, counted or exported by this handle.  Re-importing overwrites. */
int b200tsdf_export_seam_blocks(b200tsdf_t *h, int64_t *bkeys_dev, float *vox_dev,
                                uint64_t *cnt_dev, int64_t cap, int64_t *n_host,
                                void *stream);
int b200tsdf_import_halo_blocks(b200tsdf_t *h, const int64_t *bkeys_dev,
                                const float *vox_dev, const uint64_t *cnt_dev,
                                int64_t n, void *stream);

/* ---- volume state in / out (checkpoint / resume; no
 * reference counterpart -- the reference keeps its dict in process memory).
 * export_block_data: every voxel block this handle OWNS, same layout and
 * two-call protocol as export_seam_blocks.  import_blocks ADDS the blocks into
 * the volume as owned blocks (allocating them as needed): the stored state
 * (sums S, W and hit counters) is additive, so importing into an empty volume
 * restores a checkpoint and importing into a filled one merges two partial
 * fusions of disjoint keyframe sets into the volume a single fusion of all of
 * them would give. */
int b200tsdf_export_block_data(b200tsdf_t *h, int64_t *bkeys_dev, float *vox_dev,
                               uint64_t *cnt_dev, int64_t cap, int64_t *n_host, void *stream);
int b200tsdf_import_blocks(b200tsdf_t *h, const int64_t *bkeys_dev, const float *vox_dev,
                           const uint64_t *cnt_dev, int64_t n, void *stream);

/* ---- local (per-ROI, dense) TSDF of the refiner: stateless calls ---------- */

/* Replaces the config keys read by TSDFRefiner (mast3r_slam/tsdf_refine.py
 * :839-841, :872, :953-955; config/base.yaml:84-101). */
typedef struct {
  double voxel_size;      /* tsdf_refine.voxel_size      */
  double trunc_dist;      /* tsdf_refine.trunc_dist      */
  double min_confidence;  /* tsdf_refine.min_confidence  */
  double max_displacement;/* tsdf_refine.max_displacement*/
  int32_t max_grid_dim;   /* tsdf_refine.max_grid_dim    */
  int32_t ray_samples;    /* tsdf_refine.ray_samples     */
} b200tsdf_local_cfg;

/* Grid dims (nx, ny, nz) of an ROI box [min(3), max(3)] (tsdf_refine.py:844-846). */
int b200tsdf_local_dims(const float *box6_host, const b200tsdf_local_cfg *cfg,
                        int32_t dims3_host[3]);

/* Replaces TSDFRefiner._build_tsdf_robust (tsdf_refine.py:837-940).
 *  X_dev (N,3) f32 camera-frame pointmap, C_dev (N) f32, pose_dev 8 f32 Sim3,
 *  tsdf_dev / weights_dev: (nz,ny,nx) f32 grids sized by b200tsdf_local_dims.
 *  fused_host (nullable): number of points fused, -1 when fewer than 5 points
 *  are valid (grids are then the initial 1 / 0); passing it synchronises. */
int b200tsdf_local_build(const float *X_dev, const float *C_dev, const float *pose_dev,
                         const float *box6_host, int64_t n_pts, const b200tsdf_local_cfg *cfg,
                         float *tsdf_dev, float *weights_dev, int64_t *fused_host,
                         void *stream);

/* Replaces TSDFRefiner._extract_surface_safe + _sample_tsdf_trilinear
 * (tsdf_refine.py:942-1064) for an explicit ordered pixel list (the caller
 * makes the reference's random choice).  X_refined_dev must already hold a
 * copy of X_dev; hits_dev has m bytes. */
int b200tsdf_local_raycast(const float *tsdf_dev, const int32_t dims3_host[3],
                           const float *box6_host, const float *X_dev, const int64_t *pix_dev,
                           int64_t m, const b200tsdf_local_cfg *cfg, float *X_refined_dev,
                           uint8_t *hits_dev, void *stream);

/* ---- pose optimiser normal equations ---------
 * One Gauss-Newton linearisation of TSDFPoseOptimizer._optimize_single
 * (mast3r_slam/tsdf/tsdf_optimizer.py:72-116) in one launch: world points =
 * Sim3(pose) . pts_cam (float32, :76), r = tsdf(p), J = [g, -p x g, p . g]
 * (:107-110), w = max(lambda * conf, 1e-6) (:113), H = sum w J J^T (7 x 7
 * row-major), b = sum w r J, used = points with value and gradient.  pts_cam:
 * n x 3 float32, conf: n float64 (device); pose: 8 float32 [t, q(xyzw), s],
 * H, b, used: host.  Synchronises the stream (the caller solves on the host). */
int b200tsdf_pose_normal_equations(b200tsdf_t *h, int which, const float *pts_cam_dev,
                                   const double *conf_dev, int64_t n, const float *pose_host,
                                   double lambda, double *H_host, double *b_host,
                                   int64_t *used_host, void *stream);

/* The whole of TSDFPoseOptimizer._optimize_single's loop (tsdf_optimizer.py
 * :72-92) on the device: `iterations` launches, each one Gauss-Newton iteration
 * -- the linearisation above, then, in the last block to finish, the 7 x 7 solve
 * (H + damping I) delta = -b in float64 (:82), the float32 cast of delta (:84)
 * and the retraction pose <- exp(delta) * pose (:85-86) -- with no host round
 * trip in between; ONE synchronisation at the end returns the pose.  The loop
 * stops like the reference when fewer than 6 points are usable (:78-79) or the
 * system is singular (later launches idle).  pose_inout_host: 8 float64
 * [t, q_xyzw, s]; used_host: points used by the last linearisation;
 * iterations_done_host: retractions applied.  Single volume only (a sharded
 * volume merges taps between linearisation and solve, see below). */
int b200tsdf_pose_optimize(b200tsdf_t *h, int which, const float *pts_cam_dev,
                           const double *conf_dev, int64_t n, double *pose_inout_host,
                           double lambda, double damping, int32_t iterations,
                           int64_t *used_host, int32_t *iterations_done_host, void *stream);

/* ---- sharded query (multi-GPU) ---------------------------------
 * TSDFVolume.query / _estimate_gradient (global_volume.py:93-128) read the
 * voxel of the point and its six axis neighbours; with z-slab ownership those
 * 7 voxels can live on different ranks.  query_taps writes, per point, a
 * 64-byte record of the taps THIS handle owns (16 float32: tsdf and weight of
 * the centre, +x, -x, +y, -y, +z, -z voxel, then the presence mask as a float
 * and a zero; absent taps are zero; halo blocks are not reported).  Summing
 * the records over the ranks (one float all-reduce) merges them bit-exactly,
 * because every voxel has exactly one owner.  pose_host != NULL: the points
 * are camera-frame float32 and are first moved by Sim3(pose)
 * (tsdf_optimizer.py:76).  taps_dev must be 16-byte aligned.
 * query_finish evaluates the reference arithmetic on merged records -- the
 * same code b200tsdf_query runs, so a sharded query equals the single-volume
 * query bit for bit.  pose_normal_equations_taps is
 * b200tsdf_pose_normal_equations with the query replaced by merged records. */
int b200tsdf_query_taps(b200tsdf_t *h, int which, const void *pts_dev, int dtype, int64_t m,
                        const float *pose_host, float *taps_dev, void *stream);
int b200tsdf_query_finish(b200tsdf_t *h, const float *taps_dev, int64_t m, double *tsdf_dev,
                          double *grad_dev, uint8_t *flags_dev, void *stream);
int b200tsdf_pose_normal_equations_taps(b200tsdf_t *h, const float *taps_dev,
                                        const float *pts_cam_dev, const double *conf_dev,
                                        int64_t n, const float *pose_host, double lambda,
                                        double *H_host, double *b_host, int64_t *used_host,
                                        void *stream);

/* ---- quality maps on the device -----------------------
 * Replaces the per-keyframe torch/numpy round trip of
 * mast3r_slam/quality_core.py compute_batch (:127-140).
 * quality_patch_grids: per-patch (patch_size^2 pixels, row-major H x W maps in
 * device memory) lower medians -- u = median(1 - sqrt(clamp(C/(C_thr+1e-8)) *
 * clamp(Q/(Q_thr+1e-8)))) (u_from_CQ :43-49; NaN when the patch holds no
 * finite value), r = median of r_pix over the valid pixels, 0 when there is
 * none (r_from_scalar :51-52 / reduce_grid :15-27), valid = 1 if any pixel of
 * the patch is valid (valid_grid :54-56).  Grids are (H/ps) x (W/ps) float32.
 * quality_classify: classify (:63-109) on flattened grids of n patches
 * (n <= 3072): class_id int64, priority float32 and, if q_grid_dev is not
 * NULL, q = 1 - clamp(priority, 0, 1) -- the per-patch weight
 * b200tsdf_integrate_keyframes takes as q_grid_dev.  Both calls are
 * stream-ordered and never synchronise. */
int b200tsdf_quality_patch_grids(const float *C_dev, const float *Q_dev, const float *r_pix_dev,
                                 const uint8_t *valid_dev, int32_t H, int32_t W,
                                 int32_t patch_size, double C_thr, double Q_thr,
                                 float *u_grid_dev, float *r_grid_dev, float *valid_grid_dev,
                                 int32_t device, void *stream);
int b200tsdf_quality_classify(const float *delta_cov_dev, const float *r_dev, const float *u_dev,
                              int32_t n, double thr_zr, double thr_zu, double thr_dc,
                              int64_t *class_dev, float *priority_dev, float *q_grid_dev,
                              int32_t device, void *stream);

/* compute_batch (quality_core.py:127-140) for B keyframes of one size in two
 * launches: the per-patch medians of every keyframe (maps B x H*W), then one
 * CTA per keyframe for inc = valid_grid * view_w[b] (:131; view_w = the
 * view_weight scalar of :34-37), cov_ewma = alpha * prev + (1 - alpha) * inc
 * and delta_cov = cov_ewma - prev (ema_delta :39-41, float32 like torch;
 * prev_ewma_dev NULL = zeros) and classify.  All grids are B x n float32
 * (class_id int64), n = (H/ps) * (W/ps) <= 3072.  Stream-ordered. */
int b200tsdf_quality_batch(const float *C_dev, const float *Q_dev, const float *r_pix_dev,
                           const uint8_t *valid_dev, const float *view_w_dev,
                           const float *prev_ewma_dev, int32_t B, int32_t H, int32_t W,
                           int32_t patch_size, double C_thr, double Q_thr, double alpha,
                           double thr_zr, double thr_zu, double thr_dc, float *u_grid_dev,
                           float *r_grid_dev, float *valid_grid_dev, float *ewma_dev,
                           float *delta_cov_dev, int64_t *class_dev, float *priority_dev,
                           float *q_grid_dev, int32_t device, void *stream);

/* ---- self-test -------------------------------------------------------------
 * The fast keyframe path divides with one shared reciprocal per ray instead of
 * three IEEE divisions (csrc/fuse_fast.cuh).  This entry runs that sequence
 * against the IEEE division on n pseudo-random operand pairs shaped like its
 * three uses and returns the number of quotients that differ (must be 0). */
int b200tsdf_selftest_div(int32_t device, int64_t n, uint64_t seed, int64_t *mismatches_host,
                          void *stream);

#ifdef __cplusplus
}
#endif
#endif /* B200TSDF_H_ */
