// mc3.cuh -- mc_kernel3: marching cubes with ONE WARP per voxel block (included by mc.cu).
//
// mc_kernel2's profile (profiles/r02i_mc2_ncu_full.txt) is a barrier profile: 7.9
// stalled warps per issue at `bar.sync`, 54 % issue utilisation, and a deeper
// prefetch does not help (3 stages: 0.371 ms, 2 stages: 0.350 ms).  A block's work
// is too small and too uneven for 8 warps: the row pass occupies 2 of them, the
// emission of a typical block (~100 triangles) fills 3 of 8 warps.  Here every
// warp owns a block from the first load to the last store, so the only
// synchronisation is __syncwarp and DRAM latency is hidden the ordinary way, by
// the 32 independent warps of an SM:
//   load     23 lattice points per lane straight into registers, 8 + 8 + 7 at a
//            time (8-byte sums + 4-byte hit counter of the selected volume).  The
//            lattice is split so that every address is a per-block base plus an
//            immediate: 16 steps of the block itself (voxel 32 j + lane), two
//            steps per +x / +y / +z face, one step for the edges and the corner;
//            a missing neighbour block is a warp-uniform skip.
//   convert  VoxelData view -> tv[] (per-warp shared), sign / validity ballots;
//            81 nine-bit lattice rows are then assembled from the ballot words
//   rows     2 cube rows per lane on the packed rows, warp scan of (cubes,
//            triangles), active-cube list; ONE atomicAdd per block
//   emit     one triangle per lane and step; its cube comes from a triangle ->
//            cube table written with the list (binary search over the scan only
//            for halves with more than MC3_SRC_CAP triangles); the 32 triangles
//            of a step are staged in shared memory and stored as coalesced rows
// Blocks are drawn from a global ticket in list order (no static stride).
// Loads are plain read-only loads: with `.L1::no_allocate` the halo faces (re-read
// by up to 7 neighbour blocks) stopped hitting in L2 -- 1.06 GB of DRAM reads for
// a 594 MB volume (profiles/r02r_mc3_na_ncu_full.txt).
#pragma once

#define MC3_WARPS 8
#ifndef MC3_CTAS_PER_SM
#define MC3_CTAS_PER_SM 3 // 80 registers, 24 warps per SM (4 x 64 registers: 0.240 vs 0.227 ms)
#endif
#define MC3_SRC_CAP 256
struct Mc3Warp {
  float tv[736];
  // ballot words of the 23 load steps: [0..15] main block (word j: z = j >> 1, four y rows
  // of 8 x bits), [16, 17] +x face, [18, 19] +y face, [20, 21] +z face, [22] edges + corner
  uint32_t p_sgn[24], p_val[24];
  uint32_t row[84];   // index z * 9 + y: 9-bit sign row | 9-bit validity row << 16
  // lists of ONE HALF of the block (32 cube rows, <= 256 cubes): the two halves are emitted
  // one after the other, which halves the shared memory of the lists
  uint16_t cube[256]; // active cubes: cube id
  uint16_t scan[260]; // exclusive scan of their triangle counts (+ total)
  unsigned char kase[256];
  uint16_t src[MC3_SRC_CAP]; // triangle -> entry of the cube list
  float stg[32 * 9];         // the 32 triangles of one emission step, stored coalesced
};
struct Mc3Smem {
  signed char tri[256 * 16];
  unsigned char num[256];
  unsigned char ec[16]; // edge -> corner a | corner b << 4
  Mc3Warp w[MC3_WARPS];
};

// Where the 9 bits of lattice row (y, z) sit in the ballot words: word / shift of
// its x = 0..7 byte, word / bit of its x = 8 bit (packed 5 bits each).
__device__ __forceinline__ uint32_t mc3_row_desc(int r) {
  const int z = r / 9, y = r - 9 * z;
  int wl, sl, w8, s8;
  if (y < 8 && z < 8) wl = 2 * z + (y >> 2), sl = 8 * (y & 3), w8 = 16 + (z >> 2), s8 = 8 * (z & 3) + y;
  else if (z < 8) wl = 18 + (z >> 2), sl = 8 * (z & 3), w8 = 22, s8 = z;      // y == 8: +y face, xy edge
  else if (y < 8) wl = 20 + (y >> 2), sl = 8 * (y & 3), w8 = 22, s8 = 8 + y;  // z == 8: +z face, xz edge
  else wl = 22, sl = 16, w8 = 22, s8 = 24;                                    // yz edge, corner
  return (uint32_t)(wl | (sl << 5) | (w8 << 10) | (s8 << 15));
}

template <bool WRITE>
__global__ void __launch_bounds__(MC3_WARPS * 32, MC3_CTAS_PER_SM)
    mc_kernel3(const __grid_constant__ VolumeDev v, const uint32_t *__restrict__ nbr, int which,
               float vs, float *__restrict__ verts, long long tri_cap, unsigned long long *n_tris,
               uint32_t *__restrict__ sched) {
  extern __shared__ __align__(16) unsigned char mc3_raw[];
  Mc3Smem &sm = *reinterpret_cast<Mc3Smem *>(mc3_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  reinterpret_cast<int4 *>(sm.tri)[tid] = reinterpret_cast<const int4 *>(MC_TRI_TABLE)[tid];
  sm.num[tid] = MC_NUM_TRIS[tid];
  if (tid < 12) sm.ec[tid] = (unsigned char)(MC_EDGE_CORNERS[tid][0] | (MC_EDGE_CORNERS[tid][1] << 4));
  __syncthreads(); // the only CTA barrier: tables ready
  Mc3Warp &w = sm.w[warp];
  // block-invariant lane geometry (see the load steps below)
  const int l7 = lane & 7, l3 = lane >> 3;
  const int tv0 = l3 * 9 + l7, tv1 = l3 * 81 + l7 * 9 + 8, tv2 = l3 * 81 + 72 + l7;
  const int tvE = lane < 8 ? lane * 81 + 80 : lane < 16 ? 648 + (lane - 8) * 9 + 8 : lane < 24 ? 704 + lane : 728;
  const int locE = lane < 8 ? lane * 64 : lane < 16 ? (lane - 8) * 8 : lane < 24 ? lane - 16 : 0;
  const int nbE = lane < 8 ? 3 : lane < 16 ? 5 : lane < 24 ? 6 : 7;
  const uint32_t rd0 = mc3_row_desc(lane), rd1 = mc3_row_desc(lane + 32),
                 rd2 = mc3_row_desc(lane + 64 < 81 ? lane + 64 : 80);
  const char *vox_b = reinterpret_cast<const char *>(v.vox) + which * 8;
  const char *cnt_b = reinterpret_cast<const char *>(v.cnt) + which * 4;
  const float max_w = v.max_weight, min_w = v.min_weight;
  const uint32_t nb = (uint32_t)v.ctr[CTR_BLOCKS];
  // blocks are handed out in list order by a global ticket (zeroed by the neighbour pass):
  // no tail behind the warp that drew the blocks with the most triangles, and the blocks in
  // flight are a contiguous window of the list (neighbours are mostly allocated together, so
  // their halo reads hit in L2).  The ticket and the neighbour slots of the next block are
  // fetched one block ahead.
  uint32_t b = 0, bn_raw = 0, bn = 0;
  if (lane == 0) b = atomicAdd(sched, 1u);
  b = __shfl_sync(0xFFFFFFFFu, b, 0);
  if (b >= nb) return;
  uint32_t myslot = lane < 8 ? nbr[(size_t)b * 8 + lane] : B200_NO_SLOT, slot_n = B200_NO_SLOT;
  if (lane == 0) bn_raw = atomicAdd(sched, 1u);

  // VoxelData view of one lattice point (voxel_view, common.cuh) -> tv, sign / validity ballots
  auto conv = [&](float2 sw, uint32_t hh, int tvi, int word, bool st) {
    float val = NAN;
    if (hh != 0u) {
      const float wt = (hh == 1u) ? sw.y : fminf(sw.y, max_w);
      if (wt >= min_w) val = sw.x / fmaxf(1.0e-9f, wt);
    }
    if (WRITE && st) w.tv[tvi] = val;
    const unsigned bs = __ballot_sync(0xFFFFFFFFu, val < 0.0f);
    const unsigned bv = __ballot_sync(0xFFFFFFFFu, !isnan(val));
    if (lane == 0) w.p_sgn[word] = bs, w.p_val[word] = bv;
  };

  for (; b < nb; b = bn, myslot = slot_n) { // myslot, lanes 0..7: the 8 block slots of this lattice
    bn = __shfl_sync(0xFFFFFFFFu, bn_raw, 0);
    slot_n = (bn < nb && lane < 8) ? nbr[(size_t)bn * 8 + lane] : B200_NO_SLOT;
    if (lane == 0 && bn < nb) bn_raw = atomicAdd(sched, 1u);
    const uint32_t s0 = __shfl_sync(0xFFFFFFFFu, myslot, 0), s1 = __shfl_sync(0xFFFFFFFFu, myslot, 1),
                   s2 = __shfl_sync(0xFFFFFFFFu, myslot, 2), s4 = __shfl_sync(0xFFFFFFFFu, myslot, 4),
                   sE = __shfl_sync(0xFFFFFFFFu, myslot, nbE);
    unsigned long long key = 0;
    if (lane == 0) key = v.keys[s0];
    // ---- main block, steps j = 0..15: voxel index 32 j + lane (x = lane & 7, y = 4 (j & 1) +
    // (lane >> 3), z = j >> 1): every address is one base + an immediate
    const char *pv = vox_b + ((size_t)s0 * B200_BLOCK_VOX + lane) * sizeof(float4);
    const char *pc = cnt_b + ((size_t)s0 * B200_BLOCK_VOX + lane) * sizeof(unsigned long long);
#pragma unroll 1
    for (int j0 = 0; j0 < 16; j0 += 8) {
      float2 sw[8];
      uint32_t hh[8];
      const char *qv = pv + (size_t)j0 * 32 * sizeof(float4);
      const char *qc = pc + (size_t)j0 * 32 * sizeof(unsigned long long);
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        sw[u] = __ldg(reinterpret_cast<const float2 *>(qv + u * 32 * sizeof(float4)));
        hh[u] = __ldg(reinterpret_cast<const uint32_t *>(qc + u * 32 * sizeof(unsigned long long)));
      }
      const int tvj = tv0 + (j0 >> 1) * 81;
#pragma unroll
      for (int u = 0; u < 8; ++u) conv(sw[u], hh[u], tvj + (u >> 1) * 81 + (u & 1) * 36, j0 + u, true);
    }
    {
      // the halo: +x face (block 1: x = 0, y = lane & 7, z = (lane >> 3) + 4 jj), +y face
      // (block 2: y = 0), +z face (block 4: z = 0), and one step for the three edges
      // (blocks 3, 5, 6) and the corner (block 7)
      float2 sw[7];
      uint32_t hh[7];
#pragma unroll
      for (int u = 0; u < 7; ++u) sw[u] = make_float2(0.f, 0.f), hh[u] = 0u;
      auto face = [&](uint32_t s, int loc, int step, int u) { // two steps, `step` voxels apart
        if (s == B200_NO_SLOT) return;                        // (warp-uniform)
        const size_t off = (size_t)s * B200_BLOCK_VOX + loc;
        const char *fv = vox_b + off * sizeof(float4);
        const char *fc = cnt_b + off * sizeof(unsigned long long);
        sw[u] = __ldg(reinterpret_cast<const float2 *>(fv));
        hh[u] = __ldg(reinterpret_cast<const uint32_t *>(fc));
        sw[u + 1] = __ldg(reinterpret_cast<const float2 *>(fv + (size_t)step * sizeof(float4)));
        hh[u + 1] = __ldg(reinterpret_cast<const uint32_t *>(fc + (size_t)step * sizeof(unsigned long long)));
      };
      face(s1, l3 * 64 + l7 * 8, 256, 0);
      face(s2, l3 * 64 + l7, 256, 2);
      face(s4, lane, 32, 4);
      if (lane < 25 && sE != B200_NO_SLOT) {
        const size_t off = (size_t)sE * B200_BLOCK_VOX + locE;
        sw[6] = __ldg(reinterpret_cast<const float2 *>(vox_b + off * sizeof(float4)));
        hh[6] = __ldg(reinterpret_cast<const uint32_t *>(cnt_b + off * sizeof(unsigned long long)));
      }
      conv(sw[0], hh[0], tv1, 16, true);
      conv(sw[1], hh[1], tv1 + 324, 17, true);
      conv(sw[2], hh[2], tv2, 18, true);
      conv(sw[3], hh[3], tv2 + 324, 19, true);
      conv(sw[4], hh[4], tv0 + 648, 20, true);
      conv(sw[5], hh[5], tv0 + 648 + 36, 21, true);
      conv(sw[6], hh[6], tvE, 22, lane < 25);
    }
    __syncwarp();
    // ---- the 81 lattice rows (9 bits each) from the ballot words
    {
      auto build = [&](uint32_t d, int r) {
        const uint32_t wl = d & 31u, sl = (d >> 5) & 31u, w8 = (d >> 10) & 31u, s8 = d >> 15;
        const uint32_t sg = ((w.p_sgn[wl] >> sl) & 0xFFu) | (((w.p_sgn[w8] >> s8) & 1u) << 8);
        const uint32_t va = ((w.p_val[wl] >> sl) & 0xFFu) | (((w.p_val[w8] >> s8) & 1u) << 8);
        w.row[r] = sg | (va << 16);
      };
      build(rd0, lane);
      build(rd1, lane + 32);
      if (lane + 64 < 81) build(rd2, lane + 64);
    }
    __syncwarp();
    // ---- cube rows: lane owns rows r = lane and lane + 32 (y = r & 7, z = r >> 3)
    uint32_t act[2], sg[2][4], ta[2];
#pragma unroll
    for (int h2 = 0; h2 < 2; ++h2) {
      const int r = lane + 32 * h2, idx = 9 * (r >> 3) + (r & 7);
      const uint32_t R0 = w.row[idx], R1 = w.row[idx + 1], R2 = w.row[idx + 9], R3 = w.row[idx + 10];
      // x and x + 1 of all four rows at once, sign half and validity half of the packed word
      const uint32_t all1 = (R0 & (R0 >> 1)) & (R1 & (R1 >> 1)) & (R2 & (R2 >> 1)) & (R3 & (R3 >> 1));
      const uint32_t N0 = ~R0, N1 = ~R1, N2 = ~R2, N3 = ~R3;
      const uint32_t all0 = (N0 & (N0 >> 1)) & (N1 & (N1 >> 1)) & (N2 & (N2 >> 1)) & (N3 & (N3 >> 1));
      act[h2] = (all1 >> 16) & ~(all1 | all0) & 0xFFu; // all 8 corners observed, signs mixed
      const uint32_t q0 = R0 & 0x1FFu, q1 = R1 & 0x1FFu, q2 = R2 & 0x1FFu, q3 = R3 & 0x1FFu;
      sg[h2][0] = q0, sg[h2][1] = q1, sg[h2][2] = q2, sg[h2][3] = q3;
      ta[h2] = 0;
      for (uint32_t a = act[h2]; a; a &= a - 1) {
        const int x = __ffs(a) - 1;
        const uint32_t m = ((q0 >> x) & 3u) | (((q1 >> x) & 3u) << 2) | (((q2 >> x) & 3u) << 4) |
                           (((q3 >> x) & 3u) << 6);
        ta[h2] += sm.num[m];
      }
    }
    // triangles of the lower half (rows 0..31) and of the block: ONE atomicAdd per block
    const uint32_t T01 = __reduce_add_sync(0xFFFFFFFFu, ta[0] | (ta[1] << 16));
    const uint32_t T0 = T01 & 0xFFFFu, T = T0 + (T01 >> 16);
    if (T == 0) continue; // (warp-uniform) most blocks of a band have no sign change
    unsigned long long base_o = 0;
    if (lane == 0) base_o = atomicAdd(n_tris, (unsigned long long)T);
    if (!WRITE) continue;
    base_o = __shfl_sync(0xFFFFFFFFu, base_o, 0);
    key = __shfl_sync(0xFFFFFFFFu, key, 0);
    int bx, by, bz;
    unpack_block_key(key, bx, by, bz);
#pragma unroll 1
    for (int h2 = 0; h2 < 2; ++h2) {
      const uint32_t Th = h2 ? T - T0 : T0;
      if (Th == 0) continue;
      const uint32_t ac = h2 ? act[1] : act[0];
      const uint32_t mine = ((uint32_t)__popc(ac) << 16) | (h2 ? ta[1] : ta[0]);
      uint32_t incl = mine;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(0xFFFFFFFFu, incl, d);
        if (lane >= d) incl += o;
      }
      const uint32_t A = __shfl_sync(0xFFFFFFFFu, incl, 31) >> 16;
      const bool table = Th <= MC3_SRC_CAP;
      uint32_t co = (incl - mine) >> 16, to = (incl - mine) & 0xFFFFu;
      const uint32_t g0 = h2 ? sg[1][0] : sg[0][0], g1 = h2 ? sg[1][1] : sg[0][1],
                     g2 = h2 ? sg[1][2] : sg[0][2], g3 = h2 ? sg[1][3] : sg[0][3];
      const int r = lane + 32 * h2;
      for (uint32_t a = ac; a; a &= a - 1) {
        const int x = __ffs(a) - 1;
        const uint32_t m = ((g0 >> x) & 3u) | (((g1 >> x) & 3u) << 2) | (((g2 >> x) & 3u) << 4) |
                           (((g3 >> x) & 3u) << 6);
        const uint32_t n = sm.num[m];
        w.cube[co] = (uint16_t)(r * 8 + x);
        w.kase[co] = (unsigned char)m;
        w.scan[co] = (uint16_t)to;
        if (table)
          for (uint32_t q = 0; q < n; ++q) w.src[to + q] = (uint16_t)co;
        ++co;
        to += n;
      }
      if (lane == 31) w.scan[A] = (uint16_t)Th;
      __syncwarp();
      const unsigned long long base_h = base_o + (h2 ? T0 : 0u);
      // ---- emission: one triangle per lane and step, staged in shared memory and stored
      // as 9 coalesced rows (a lane's own 36 bytes would be 9 stores of stride 36)
      for (uint32_t k0 = 0; k0 < Th; k0 += 32) {
        const uint32_t k = k0 + lane;
        if (k < Th) {
          uint32_t lo = 0;
          if (table) {
            lo = w.src[k];
          } else {
            uint32_t hi = A; // largest c with scan[c] <= k
            while (hi - lo > 1) {
              const uint32_t mid = (lo + hi) >> 1;
              if (w.scan[mid] <= k) lo = mid; else hi = mid;
            }
          }
          const int q = (int)(k - w.scan[lo]), c = (int)w.cube[lo], m = (int)w.kase[lo];
          const int x = c & 7, y = (c >> 3) & 7, z = c >> 6;
          float *out = w.stg + lane * 9;
#pragma unroll
          for (int vtx = 0; vtx < 3; ++vtx) {
            const int e = sm.tri[m * 16 + 3 * q + vtx];
            const int ab = sm.ec[e], a = ab & 7, bb = ab >> 4;
            const int ax = x + (a & 1), ay = y + ((a >> 1) & 1), az = z + (a >> 2);
            const int cx = x + (bb & 1), cy = y + ((bb >> 1) & 1), cz = z + (bb >> 2);
            const float fa = w.tv[az * 81 + ay * 9 + ax], fb = w.tv[cz * 81 + cy * 9 + cx];
            const float mu = __fdiv_rn(fa, __fsub_rn(fa, fb));
            const float pa[3] = {__fmul_rn((float)(bx * 8 + ax), vs), __fmul_rn((float)(by * 8 + ay), vs),
                                 __fmul_rn((float)(bz * 8 + az), vs)};
            const float pb[3] = {__fmul_rn((float)(bx * 8 + cx), vs), __fmul_rn((float)(by * 8 + cy), vs),
                                 __fmul_rn((float)(bz * 8 + cz), vs)};
#pragma unroll
            for (int d = 0; d < 3; ++d)
              out[vtx * 3 + d] = __fadd_rn(pa[d], __fmul_rn(__fsub_rn(pb[d], pa[d]), mu));
          }
        }
        __syncwarp();
        const long long f0 = ((long long)base_h + k0) * 9;             // first float of the step
        const long long fend = min((long long)base_h + Th, tri_cap) * 9; // triangles beyond the
        const int nf = (int)min((long long)min(32u, Th - k0) * 9, fend - f0); // capacity are dropped
#pragma unroll
        for (int i = 0; i < 9; ++i) {
          const int f = i * 32 + lane;
          if (f < nf) verts[f0 + f] = w.stg[f];
        }
        __syncwarp(); // stg is rewritten by the next step
      }
      __syncwarp(); // the lists are rewritten by the other half
    }
    __syncwarp(); // tv / lists are rewritten by the next block
  }
}
