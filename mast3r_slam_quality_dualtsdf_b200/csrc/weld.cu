// weld.cu -- indexed mesh on the device (b200tsdf_extract_mesh with
// `verts` + `faces i32`; 8(a) A11).
//
// The marching-cubes kernel (mc.cu) emits a triangle soup.  Two triangles that
// share a cube edge compute their common vertex with the same float32 operation
// order from the same two corners, so shared vertices are BIT-identical -- also
// across voxel-block and z-slab seams.  Welding is therefore an exact
// de-duplication on the 96-bit pattern (x, y, z), done here with an
// open-addressing table instead of a host-side sort (mesh_io.weld):
//   1. weld_insert_kernel: every soup vertex finds / claims the table slot of its
//      position; the slot keeps the SMALLEST soup index with that position
//      (atomicMin), so the result does not depend on the thread schedule;
//   2. weld_flag_kernel marks the representatives, an exclusive scan numbers them
//      in soup order;
//   3. weld_emit_kernel writes the unique vertices and the int32 face indices.
// No reference counterpart (the reference has no mesh path; its PLY writer,
// mast3r_slam/evaluate.py:88-106, stores points).
#include <stdint.h>

#include "common.cuh"

#define WELD_EMPTY 0xFFFFFFFFu
#define SCAN_THREADS 1024
#define SCAN_ITEMS 4

__device__ __forceinline__ uint32_t weld_hash(uint32_t a, uint32_t b, uint32_t c) {
  uint32_t h = a * 0x9E3779B1u;
  h = (h ^ (h >> 15)) + b * 0x85EBCA77u;
  h = (h ^ (h >> 13)) + c * 0xC2B2AE3Du;
  h ^= h >> 16, h *= 0x7FEB352Du, h ^= h >> 15;
  return h;
}

__global__ void __launch_bounds__(256)
    weld_insert_kernel(const uint32_t *__restrict__ v, long long n, uint32_t *table, uint32_t mask,
                       uint32_t *__restrict__ slot_of) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t a = v[3 * i], b = v[3 * i + 1], c = v[3 * i + 2];
  uint32_t s = weld_hash(a, b, c) & mask;
  for (;;) {
    uint32_t e = __ldcg(table + s);
    if (e == WELD_EMPTY) {
      e = atomicCAS(table + s, WELD_EMPTY, (uint32_t)i);
      if (e == WELD_EMPTY) break; // claimed
    }
    // e: some soup index stored in this slot (every index a slot ever holds has the same position)
    if (v[3 * (size_t)e] == a && v[3 * (size_t)e + 1] == b && v[3 * (size_t)e + 2] == c) {
      atomicMin(table + s, (uint32_t)i);
      break;
    }
    s = (s + 1) & mask;
  }
  slot_of[i] = s;
}

__global__ void __launch_bounds__(256)
    weld_flag_kernel(const uint32_t *__restrict__ table, const uint32_t *__restrict__ slot_of,
                     long long n, uint32_t *__restrict__ flag) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) flag[i] = table[slot_of[i]] == (uint32_t)i ? 1u : 0u;
}

__global__ void __launch_bounds__(256)
    weld_emit_kernel(const float *__restrict__ v, const uint32_t *__restrict__ table,
                     const uint32_t *__restrict__ slot_of, const uint32_t *__restrict__ pos,
                     long long n, float *__restrict__ verts_out, long long vert_cap,
                     int32_t *__restrict__ faces_out) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t r = table[slot_of[i]];
  const uint32_t idx = pos[r];
  if (faces_out) faces_out[i] = (int32_t)idx;
  if (r == (uint32_t)i && verts_out && (long long)idx < vert_cap) {
    verts_out[3 * (size_t)idx] = v[3 * i];
    verts_out[3 * (size_t)idx + 1] = v[3 * i + 1];
    verts_out[3 * (size_t)idx + 2] = v[3 * i + 2];
  }
}

// ---- exclusive scan of uint32 (block scan + recursive scan of the block sums)
__global__ void __launch_bounds__(SCAN_THREADS)
    scan_block_kernel(const uint32_t *__restrict__ in, uint32_t *__restrict__ out, long long n,
                      uint32_t *__restrict__ block_sums) {
  __shared__ uint32_t warp_tot[SCAN_THREADS / 32];
  const long long base = ((long long)blockIdx.x * SCAN_THREADS + threadIdx.x) * SCAN_ITEMS;
  uint32_t x[SCAN_ITEMS], t = 0;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k) {
    x[k] = base + k < n ? in[base + k] : 0u;
    t += x[k];
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t inc = t;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, inc, d);
    if (lane >= d) inc += y;
  }
  if (lane == 31) warp_tot[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    uint32_t w = warp_tot[lane];
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t y = __shfl_up_sync(0xFFFFFFFFu, w, d);
      if (lane >= d) w += y;
    }
    warp_tot[lane] = w; // inclusive over warps
  }
  __syncthreads();
  uint32_t excl = inc - t + (warp > 0 ? warp_tot[warp - 1] : 0u);
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k) {
    if (base + k < n) out[base + k] = excl;
    excl += x[k];
  }
  if (block_sums && threadIdx.x == SCAN_THREADS - 1) block_sums[blockIdx.x] = warp_tot[31];
}

__global__ void __launch_bounds__(SCAN_THREADS)
    scan_add_kernel(uint32_t *__restrict__ out, long long n, const uint32_t *__restrict__ offs) {
  const long long base = ((long long)blockIdx.x * SCAN_THREADS + threadIdx.x) * SCAN_ITEMS;
  const uint32_t o = offs[blockIdx.x];
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k)
    if (base + k < n) out[base + k] += o;
}

// out[i] = sum(in[0..i)); `tmp` holds the block sums of every level (>= n / 2048 + 64 entries).
static cudaError_t exclusive_scan_u32(const uint32_t *in, uint32_t *out, long long n, uint32_t *tmp,
                                      cudaStream_t st) {
  const long long per = (long long)SCAN_THREADS * SCAN_ITEMS;
  const long long nb = (n + per - 1) / per;
  if (nb <= 1) {
    scan_block_kernel<<<1, SCAN_THREADS, 0, st>>>(in, out, n, nullptr);
    return cudaGetLastError();
  }
  scan_block_kernel<<<(unsigned)nb, SCAN_THREADS, 0, st>>>(in, out, n, tmp);
  cudaError_t e = exclusive_scan_u32(tmp, tmp, nb, tmp + nb, st); // in place: block-local reads first
  if (e != cudaSuccess) return e;
  scan_add_kernel<<<(unsigned)nb, SCAN_THREADS, 0, st>>>(out, n, tmp);
  return cudaGetLastError();
}

static int weld_impl(int device, const float *soup_dev, int64_t n_tris, float *verts_dev,
                     int64_t vert_cap, int32_t *faces_dev, int64_t *n_verts_host, cudaStream_t st) {
  const long long n = 3 * (long long)n_tris;
  *n_verts_host = 0;
  if (n == 0) return B200TSDF_OK;
  if (n >= 0xFFFFFFF0ll) {
    b200_set_error("b200tsdf_weld_mesh: %lld triangles exceed the 32-bit vertex index", (long long)n_tris);
    return B200TSDF_E_INVALID;
  }
  B200_CUDA(cudaSetDevice(device));
  unsigned long long tsize = 1024;
  while (tsize < 2ull * (unsigned long long)n) tsize <<= 1;
  const long long scan_tmp = n / 2048 + 4096;
  uint32_t *buf = nullptr;
  // table | slot_of | flag -> pos (scan in place is not possible: separate) | scan tmp
  const size_t words = (size_t)tsize + 3 * (size_t)n + (size_t)scan_tmp;
  B200_CUDA(cudaMalloc(&buf, words * sizeof(uint32_t)));
  uint32_t *table = buf, *slot_of = table + tsize, *flag = slot_of + n, *pos = flag + n,
           *tmp = pos + n;
  cudaError_t e = cudaMemsetAsync(table, 0xFF, (size_t)tsize * sizeof(uint32_t), st);
  const unsigned grid = (unsigned)((n + 255) / 256);
  if (e == cudaSuccess) {
    weld_insert_kernel<<<grid, 256, 0, st>>>((const uint32_t *)soup_dev, n, table,
                                             (uint32_t)(tsize - 1), slot_of);
    weld_flag_kernel<<<grid, 256, 0, st>>>(table, slot_of, n, flag);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = exclusive_scan_u32(flag, pos, n, tmp, st);
  uint32_t last_pos = 0, last_flag = 0;
  if (e == cudaSuccess) {
    weld_emit_kernel<<<grid, 256, 0, st>>>(soup_dev, table, slot_of, pos, n, verts_dev,
                                           (long long)vert_cap, faces_dev);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess)
    e = cudaMemcpyAsync(&last_pos, pos + n - 1, 4, cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess)
    e = cudaMemcpyAsync(&last_flag, flag + n - 1, 4, cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  cudaFree(buf);
  if (e != cudaSuccess) {
    b200_set_error("b200tsdf_weld_mesh: %s", cudaGetErrorString(e));
    return B200TSDF_E_CUDA;
  }
  *n_verts_host = (int64_t)last_pos + (int64_t)last_flag;
  return B200TSDF_OK;
}

extern "C" int b200tsdf_weld_mesh(int32_t device, const float *soup_dev, int64_t n_tris,
                                  float *verts_dev, int64_t vert_cap, int32_t *faces_dev,
                                  int64_t *n_verts_host, void *stream) {
  B200_NVTX("b200tsdf_weld_mesh");
  if (n_tris < 0 || !n_verts_host || (n_tris > 0 && !soup_dev) || vert_cap < 0) {
    b200_set_error("b200tsdf_weld_mesh: invalid argument");
    return B200TSDF_E_INVALID;
  }
  return weld_impl(device, soup_dev, n_tris, verts_dev, vert_cap, faces_dev, n_verts_host,
                   (cudaStream_t)stream);
}

extern "C" int b200tsdf_extract_mesh_indexed(b200tsdf_t *h, int which, float *verts_dev,
                                             int64_t vert_cap, int32_t *faces_dev,
                                             int64_t tri_cap, int64_t *n_verts_host,
                                             int64_t *n_tris_host, void *stream) {
  B200_NVTX("b200tsdf_extract_mesh_indexed");
  if (!h || !n_verts_host || !n_tris_host || vert_cap < 0 || tri_cap < 0) {
    b200_set_error("b200tsdf_extract_mesh_indexed: invalid argument");
    return B200TSDF_E_INVALID;
  }
  *n_verts_host = 0;
  // triangle count first (the soup is internal scratch sized exactly)
  int rc = b200tsdf_extract_mesh(h, which, nullptr, 0, n_tris_host, stream);
  if (rc) return rc;
  const int64_t nt = *n_tris_host;
  if (nt == 0) return B200TSDF_OK;
  float *soup = nullptr;
  B200_CUDA(cudaMalloc(&soup, (size_t)nt * 9 * sizeof(float)));
  int64_t nt2 = 0;
  rc = b200tsdf_extract_mesh(h, which, soup, nt, &nt2, stream);
  if (rc == 0 && nt2 != nt) {
    b200_set_error("b200tsdf_extract_mesh_indexed: the volume changed between the two passes");
    rc = B200TSDF_E_INVALID;
  }
  // faces are written only when the caller's buffer holds all of them
  if (rc == 0)
    rc = weld_impl(h->cfg.device, soup, nt, verts_dev, vert_cap, tri_cap >= nt ? faces_dev : nullptr,
                   n_verts_host, (cudaStream_t)stream);
  cudaFree(soup);
  return rc;
}
