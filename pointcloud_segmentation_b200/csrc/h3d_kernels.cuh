// h3d_kernels.cuh -- the Hough kernels (sm_100a).
//   K1 vote_kernel      Hough::add  -> pointVote   lib/hough.cpp:50-55, 67-100
//   K2 argmax kernels   Hough::getLine             lib/hough.cpp:103-112
//   K4 unvote_kernel    Hough::subtract            lib/hough.cpp:58-63
// K3 (inliers + orthogonal_LSQ) lives in h3d_fit.cuh, the fused loop in h3d_loop.cuh.
#ifndef H3D_KERNELS_CUH_
#define H3D_KERNELS_CUH_

#include "h3d_device.cuh"

namespace h3d {

constexpr int kVoteThreads = 512;
constexpr int kVoteWarps = kVoteThreads / 32;

// scan one plane for (max, first max cell), zeroing it for its next use; optional dense
// write-back in the reference layout.  Returns the warp-combined key in lane 0..31
// (votes << 32 | ~cell).  THREADS threads take part, thread `t` of them.
template <int THREADS>
__device__ __forceinline__ unsigned long long scan_plane(const HoughDev &h, unsigned *plane, int t, int j,
                                                         unsigned *__restrict__ dense) {
  uint4 *p4 = reinterpret_cast<uint4 *>(plane);
  unsigned best = 0, bcell = 0;
  for (unsigned q = t; q < h.plane_stride / 4; q += THREADS) {
    uint4 v = p4[q];
    if ((v.x | v.y | v.z | v.w) == 0) continue;
    p4[q] = make_uint4(0, 0, 0, 0);
    const unsigned c0 = 4 * q;
    if (c0 + 3 >= h.plane_cells) {   // slack cells of the BOUNDED vote forms: zeroed, never counted
      if (c0 >= h.plane_cells) v.x = 0;
      if (c0 + 1 >= h.plane_cells) v.y = 0;
      if (c0 + 2 >= h.plane_cells) v.z = 0;
      v.w = 0;
    }
    if (v.x > best) { best = v.x; bcell = c0; }
    if (v.y > best) { best = v.y; bcell = c0 + 1; }
    if (v.z > best) { best = v.z; bcell = c0 + 2; }
    if (v.w > best) { best = v.w; bcell = c0 + 3; }
    if (dense) {
      const size_t nb = (size_t)h.num_b;
      if (v.x) dense[(size_t)c0 * nb + j] = v.x;
      if (v.y) dense[(size_t)(c0 + 1) * nb + j] = v.y;
      if (v.z) dense[(size_t)(c0 + 2) * nb + j] = v.z;
      if (v.w) dense[(size_t)(c0 + 3) * nb + j] = v.w;
    }
  }
  // strided threads see interleaved cells: combine on (votes desc, cell asc)
  const unsigned m = __reduce_max_sync(0xFFFFFFFFu, best);
  const unsigned cm = __reduce_min_sync(0xFFFFFFFFu, best == m ? bcell : 0xFFFFFFFFu);
  return ((unsigned long long)m << 32) | (0xFFFFFFFFu - cm);
}

__device__ __forceinline__ void store_record(const HoughDev &h, int l, unsigned long long k) {
  const unsigned votes = (unsigned)(k >> 32);
  h.dvotes[l] = votes;
  h.dcell[l] = votes ? 0xFFFFFFFFu - (unsigned)k : 0u;
  h.dstale[l] = 0;
}

// ---------------------------------------------------------------------------
// K1 (streaming form, any N): build the planes of a list of directions in shared
// memory.
//
// A work item is a group of D consecutive entries of `list` (local direction
// numbers; identity when list == nullptr).  The CTA zeroes D planes once, then
// per item: every thread projects its points on the D directions (scaled
// coefficients in registers, points streamed coalesced from the SoA arrays, which
// stay in L1/L2), increments shared-memory cells, and after one barrier the planes
// are scanned with uint4 loads for (max, first max cell) and zeroed in the same
// pass for the next item.  Nothing but 9 bytes per direction goes to HBM unless
// `dense` is given (Hough::VotingSpace materialisation, reference layout).
//
// ABLATE (measurement only): 1 = projection arithmetic without shared atomics,
// 2 = shared atomics on hashed cells without the FP64 projection.
// ---------------------------------------------------------------------------
template <int D, int ABLATE>
__device__ __forceinline__ void vote_items(const HoughDev &h, const int *__restrict__ list, int count,
                                           unsigned *__restrict__ dense, unsigned *planes,
                                           unsigned long long (*warp_best)[kVoteWarps], int cta, int n_cta) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const unsigned stride = h.plane_stride;
  const int n = point_count(h);
  unsigned sink = 0;

  for (int item = cta; item * D < count; item += n_cta) {
    int dl[D];
    FastCoef f[D];
#pragma unroll
    for (int d = 0; d < D; ++d) {
      const int e = item * D + d;
      dl[d] = e < count ? (list ? list[e] : e) : -1;
      f[d] = make_fast(h.coef[h.dir_begin + (dl[d] < 0 ? 0 : dl[d])], h.inv_s);
    }

    constexpr int U = 2;   // points in flight per thread: U*D independent vote chains
    const double kmagic = h.kmagic;
    const unsigned num_x = (unsigned)h.num_x, plane_cells = h.plane_cells;
    const bool no_fast = !h.fast_ok;
    for (int i0 = tid; i0 < n; i0 += U * kVoteThreads) {
      double x[U], y[U], z[U];
      int wi[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int i = i0 + u * kVoteThreads;
        const bool in = i < n;
        wi[u] = in ? h.w[i] : 0;
        x[u] = in ? h.px[i] : 0.0; y[u] = in ? h.py[i] : 0.0; z[u] = in ? h.pz[i] : 0.0;
      }
      unsigned cell[U][D];
      unsigned unsure = 0;
#pragma unroll
      for (int u = 0; u < U; ++u) {
#pragma unroll
        for (int d = 0; d < D; ++d) {
          bool sure = true;
          if (ABLATE == 2) {
            cell[u][d] = ((unsigned)(i0 + u * kVoteThreads) * 2654435761u + (unsigned)dl[d] * 40503u) % plane_cells;
          } else {
            cell[u][d] = vote_cell_fast(f[d], kmagic, num_x, x[u], y[u], z[u], &sure);
          }
          if (wi[u] == 0 || dl[d] < 0) { cell[u][d] = 0xFFFFFFFFu; sure = true; }
          unsure |= (!sure | no_fast) ? (1u << (u * D + d)) : 0u;
        }
      }
      if (ABLATE != 2 && unsure) {   // rare: a projection within 8e-6 cell of a boundary
#pragma unroll
        for (int u = 0; u < U; ++u) {
#pragma unroll
          for (int d = 0; d < D; ++d) {
            if (unsure & (1u << (u * D + d)) && wi[u] != 0 && dl[d] >= 0)
              cell[u][d] = vote_cell_exact(h.coef + h.dir_begin + dl[d], h.max_x, h.dx, h.num_x, x[u], y[u], z[u]);
          }
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
#pragma unroll
        for (int d = 0; d < D; ++d) {
          if (ABLATE == 1) sink ^= cell[u][d];
          else if (cell[u][d] < plane_cells) atomicAdd(&planes[d * stride + cell[u][d]], (unsigned)wi[u]);
        }
      }
    }
    __syncthreads();

#pragma unroll
    for (int d = 0; d < D; ++d) {
      if (dl[d] < 0) continue;   // uniform across the CTA
      const unsigned long long k =
          scan_plane<kVoteThreads>(h, planes + d * stride, tid, h.dir_begin + dl[d], dense);
      if (lane == 0) warp_best[d][warp] = k;
    }
    __syncthreads();
    if (warp < D && dl[warp] >= 0) {
      unsigned long long k = lane < kVoteWarps ? warp_best[warp][lane] : 0ull;
      k = warp_max_u64(k);
      if (lane == 0) store_record(h, dl[warp], k);
    }
    // the next item's shared atomics may start at once: planes are zero again and
    // warp_best is rewritten only after the next post-vote barrier
  }
  if (ABLATE == 1 && sink == 0x9E3779B9u) h.dvotes[0] = sink;   // keep the arithmetic alive
}

template <int D, int ABLATE>
__global__ void __launch_bounds__(kVoteThreads, 2)
vote_kernel(const HoughDev h, const int *__restrict__ list, const int *__restrict__ count_ptr, int count_imm,
            unsigned *__restrict__ dense) {
  extern __shared__ __align__(16) unsigned planes[];
  __shared__ unsigned long long warp_best[D][kVoteWarps];
  const int count = count_ptr ? *count_ptr : count_imm;
  uint4 *p4 = reinterpret_cast<uint4 *>(planes);
  for (unsigned q = threadIdx.x; q < D * h.plane_stride / 4; q += kVoteThreads) p4[q] = make_uint4(0, 0, 0, 0);
  __syncthreads();
  vote_items<D, ABLATE>(h, list, count, dense, planes, warp_best, blockIdx.x, gridDim.x);
}

// ---------------------------------------------------------------------------
// K1 (TMA-staged streaming form, any N, unit weights): the 1 M-point / direction-sharded kernel.
//
// The D = 2 streaming form above re-reads every point once per PAIR of directions with four LDGs per point
// (ncu, round 1: 39.7 warp-instructions per warp-vote, LSU pipe 37.6 %, spills at D = 4).  Here
//   * a work item is D <= 8 directions (planes in shared memory), so a point is fetched once per EIGHT directions;
//   * the points arrive as TILES through the TMA: one elected thread issues three 1-D bulk copies
//     (cp.async.bulk.shared.global, SASS UBLKCP) of kStreamTile x/y/z doubles per stage into a ring of
//     kStreamStages stages, completion is counted on an mbarrier per stage; no thread spends an LDG or an
//     address computation on a point;
//   * a thread lifts its kStreamP points of the tile from shared memory into registers once (LDS.64, conflict
//     free), the CTA barrier that follows frees the stage for the next bulk copy, and the D directions are then
//     voted from registers with the scaled coefficients of one direction broadcast from shared memory;
//   * all weights are +1 (compact live points), so the increment is the merging red.shared.add form.
// TMA = false keeps the same loop with plain coalesced LDGs (the A/B of DESIGN.md).
// ---------------------------------------------------------------------------
constexpr int kStreamThreads = 512, kStreamWarps = kStreamThreads / 32;
constexpr int kStreamP = 4;                                  // points per thread per tile
constexpr int kStreamTile = kStreamThreads * kStreamP;       // 2048 points = 48 KB per stage
constexpr int kStreamStages = 2;
constexpr int kStreamMaxD = 8;

__device__ __forceinline__ void mbar_init(unsigned bar_sa, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(bar_sa), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned bar_sa, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar_sa), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar_sa, unsigned parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}"
      :: "r"(bar_sa), "r"(parity) : "memory");
}
// 1-D bulk copy global -> shared through the TMA unit; bytes % 16 == 0, both addresses 16-byte aligned
__device__ __forceinline__ void tma_load_1d(unsigned dst_sa, const void *src, unsigned bytes, unsigned bar_sa) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               :: "r"(dst_sa), "l"(src), "r"(bytes), "r"(bar_sa) : "memory");
}

struct StreamShared {
  unsigned long long full[kStreamStages];            // mbarriers: the stage's three bulk copies have landed
  unsigned long long empty[kStreamStages];           // mbarriers: every consumer warp has lifted its points out of the stage
  unsigned long long warp_best[kStreamMaxD][kStreamWarps];
  double coef[kStreamMaxD][6];                       // scaled coefficients (FastCoef) of the item's directions
  int dir[kStreamMaxD];                              // local direction numbers of the item (-1: none)
};

__device__ __forceinline__ void mbar_arrive(unsigned bar_sa) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(bar_sa) : "memory");
}
__device__ __forceinline__ void consumer_bar() {     // the 512 voting threads only (the producer warp runs ahead)
  asm volatile("bar.sync 1, %0;" :: "n"(kStreamThreads) : "memory");
}

// TMA = true: launched with kStreamThreads + 32 threads; warp 16 is the PRODUCER (one lane issues the bulk
// copies, paced by the `empty` barriers), warps 0-15 vote and never wait for each other inside an item --
// they meet only through the `full` barriers (the first version put a CTA barrier behind every tile to free
// the stage and lost 25 % against plain LDGs: profiles/r02_stream_vote_ab_v1.txt).
template <int D, bool BOUNDED, bool TMA>
__global__ void __launch_bounds__(TMA ? kStreamThreads + 32 : kStreamThreads, 1)
vote_stream_kernel(const __grid_constant__ HoughDev h, const int *__restrict__ list, const int *__restrict__ count_ptr,
                   int count_imm, unsigned *__restrict__ dense, unsigned *__restrict__ gplanes, int gplanes_cap,
                   int *__restrict__ sliced_flag) {
  extern __shared__ __align__(128) unsigned char stream_smem[];
  __shared__ __align__(16) StreamShared sh;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int count = count_ptr ? *count_ptr : count_imm;
  const int n = point_count(h);
  // POINT SLICES (LDG form, re-vote lists): a list of a few hundred stale planes gives fewer work items than the GPU
  // has SMs (40 planes of a direction shard: 6 items, each streaming the whole cloud).  Then an item becomes (group of
  // D directions, slice of the point tiles): the CTA votes its slice, adds the non-zero cells of its planes to the
  // group's planes in global memory, and stream_merge_kernel reduces those to the records.  Decided on the device
  // (the host does not know the list length); *sliced_flag tells the merge kernel how many global planes to reduce.
  const int n_groups_all = (count + D - 1) / D;
  const int n_tiles_all = (n + kStreamTile - 1) / kStreamTile;
  int slices = 1;
  if (!TMA && gplanes && !dense && count > 0 && count <= gplanes_cap && n_groups_all < (int)gridDim.x && n_tiles_all >= 8) {
    slices = (2 * (int)gridDim.x + n_groups_all - 1) / n_groups_all;
    slices = min(slices, n_tiles_all / 4);             // at least 4 tiles (8192 points) per slice
  }
  if (sliced_flag && blockIdx.x == 0 && tid == 0) *sliced_flag = slices > 1 ? count : 0;
  const unsigned stride = h.plane_stride;
  // layout: [ring: stages x {x, y, z} x kStreamTile doubles][planes: D x stride u32][pad word]
  double *ring = reinterpret_cast<double *>(stream_smem);
  unsigned *planes = reinterpret_cast<unsigned *>(stream_smem + (size_t)kStreamStages * 3 * kStreamTile * sizeof(double));
  const unsigned planes_sa = (unsigned)__cvta_generic_to_shared(planes);
  const unsigned dummy_sa = planes_sa + (unsigned)D * stride * 4u;
  const unsigned ring_sa = (unsigned)__cvta_generic_to_shared(ring);
  const unsigned full_sa = (unsigned)__cvta_generic_to_shared(&sh.full[0]);
  const unsigned empty_sa = (unsigned)__cvta_generic_to_shared(&sh.empty[0]);
  const int n_items = n_groups_all * slices;
  const int my_items = blockIdx.x < n_items ? (n_items - blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  const int n_tiles = n_tiles_all;
  const long long total_fills = (long long)my_items * n_tiles;   // TMA form only (never sliced)

  if (tid < kStreamThreads) {
    uint4 *p4 = reinterpret_cast<uint4 *>(planes);
    for (unsigned q = tid; q < (unsigned)D * stride / 4 + 1; q += kStreamThreads) p4[q] = make_uint4(0, 0, 0, 0);
  }
  if (TMA && tid == 0) {
    for (int s = 0; s < kStreamStages; ++s) { mbar_init(full_sa + 8u * s, 1); mbar_init(empty_sa + 8u * s, kStreamWarps); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();
  if (TMA && warp == kStreamWarps) {
    // ---- producer: fill k = tile (k mod n_tiles) of the cloud into stage (k mod kStreamStages) ----------
    if (lane == 0) {
      for (long long k = 0; k < total_fills; ++k) {
        const int st = (int)(k % kStreamStages);
        if (k >= kStreamStages) mbar_wait(empty_sa + 8u * st, (unsigned)(((k / kStreamStages) - 1) & 1));
        const int i0 = (int)(k % n_tiles) * kStreamTile;
        const int cnt = min(kStreamTile, n - i0);
        const unsigned bytes = (unsigned)(((cnt + 1) & ~1) * 8);   // 16-byte granules: one double past an odd end is read, never used
        const unsigned bar = full_sa + 8u * st;
        const unsigned dst = ring_sa + (unsigned)st * 3u * kStreamTile * 8u;
        mbar_expect_tx(bar, 3u * bytes);
        tma_load_1d(dst, h.px + i0, bytes, bar);
        tma_load_1d(dst + kStreamTile * 8u, h.py + i0, bytes, bar);
        tma_load_1d(dst + 2u * kStreamTile * 8u, h.pz + i0, bytes, bar);
      }
    }
    return;
  }

  const double kmagic = h.kmagic;
  const unsigned num_x = (unsigned)h.num_x, plane_cells = h.plane_cells, row4 = 4u * (unsigned)h.num_x;
  const bool no_fast = !h.fast_ok;
  long long fill = 0;
  for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
    // slice-major: CTAs running side by side vote the SAME tiles for different groups (the tiles come from L2 once)
    const int group = slices > 1 ? item % n_groups_all : item, slice = slices > 1 ? item / n_groups_all : 0;
    const int t_begin = slices > 1 ? (int)((long long)n_tiles * slice / slices) : 0;
    const int t_end = slices > 1 ? (int)((long long)n_tiles * (slice + 1) / slices) : n_tiles;
    if (tid < D) {
      const int e = group * D + tid;
      const int l = e < count ? (list ? list[e] : e) : -1;
      sh.dir[tid] = l;
      const FastCoef f = make_fast(h.coef[h.dir_begin + (l < 0 ? 0 : l)], h.inv_s);
      sh.coef[tid][0] = f.c0s; sh.coef[tid][1] = f.c1n; sh.coef[tid][2] = f.c2s; sh.coef[tid][3] = f.bxn; sh.coef[tid][4] = f.byn;
    }
    if (TMA) consumer_bar(); else __syncthreads();
    for (int t = t_begin; t < t_end; ++t, ++fill) {
      const int st = (int)(fill % kStreamStages);
      const int i0 = t * kStreamTile;
      double x[kStreamP], y[kStreamP], z[kStreamP];
      unsigned live = 0;
      if (TMA) {
        mbar_wait(full_sa + 8u * st, (unsigned)((fill / kStreamStages) & 1));
        const double *bx = ring + (size_t)st * 3 * kStreamTile, *by = bx + kStreamTile, *bz = by + kStreamTile;
#pragma unroll
        for (int p = 0; p < kStreamP; ++p) {
          const int k = tid + p * kStreamThreads;
          x[p] = bx[k]; y[p] = by[k]; z[p] = bz[k];
          live |= (i0 + k < n) ? (1u << p) : 0u;
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(empty_sa + 8u * st);   // this warp's points are in registers: 16 arrivals free the stage
      } else {
#pragma unroll
        for (int p = 0; p < kStreamP; ++p) {
          const int i = i0 + tid + p * kStreamThreads;
          const bool ok = i < n;
          x[p] = ok ? h.px[i] : 0.0; y[p] = ok ? h.py[i] : 0.0; z[p] = ok ? h.pz[i] : 0.0;
          live |= ok ? (1u << p) : 0u;
        }
      }
      const bool full_tile = i0 + kStreamTile <= n;          // CTA-uniform
#pragma unroll 1
      for (int d = 0; d < D; ++d) {
        const int l = sh.dir[d];
        if (l < 0) break;
        FastCoef f;
        f.c0s = sh.coef[d][0]; f.c1n = sh.coef[d][1]; f.c2s = sh.coef[d][2]; f.bxn = sh.coef[d][3]; f.byn = sh.coef[d][4];
        const unsigned mine_sa = planes_sa + (unsigned)d * stride * 4u;
        const unsigned drop_off = dummy_sa - mine_sa;
        unsigned off[kStreamP];
        unsigned gmin = 0xFFFFFFFFu, hacc = 0;
#pragma unroll
        for (int p = 0; p < kStreamP; ++p) {
          if (BOUNDED) off[p] = vote_off_bounded(f, kmagic, row4, x[p], y[p], z[p], gmin);
          else off[p] = 4u * vote_cell_acc(f, kmagic, num_x, x[p], y[p], z[p], gmin, hacc);
        }
        if (batch_unsure(gmin, hacc) | no_fast) {            // rare: within 8e-6 cell of a boundary
#pragma unroll
          for (int p = 0; p < kStreamP; ++p) {
            const unsigned cell = vote_cell_checked_h(&h, l, x[p], y[p], z[p]);
            off[p] = cell < plane_cells ? 4u * cell : drop_off;
          }
        }
#pragma unroll
        for (int p = 0; p < kStreamP; ++p) {
          bool in = BOUNDED ? true : off[p] < 4u * plane_cells;
          if (!full_tile) in = in && ((live >> p) & 1u);
          red_shared_inc(in ? mine_sa + off[p] : dummy_sa);
        }
      }
    }
    if (TMA) consumer_bar(); else __syncthreads();
    if (slices > 1) {        // partial planes of a slice: into the group's global planes, shared planes back to zero
#pragma unroll 1
      for (int d = 0; d < D; ++d) {
        if (sh.dir[d] < 0) continue;
        unsigned *g = gplanes + (size_t)(group * D + d) * plane_cells;
        unsigned *pl = planes + (size_t)d * stride;
        for (unsigned q = tid; q < stride; q += kStreamThreads) {
          const unsigned v = pl[q];
          if (v) { pl[q] = 0; if (q < plane_cells) atomicAdd(&g[q], v); }
        }
      }
      __syncthreads();
      continue;
    }
#pragma unroll
    for (int d = 0; d < D; ++d) {
      const int l = sh.dir[d];
      if (l < 0) continue;   // uniform across the CTA
      const unsigned long long k = scan_plane<kStreamThreads>(h, planes + (size_t)d * stride, tid, h.dir_begin + l, dense);
      if (lane == 0) sh.warp_best[d][warp] = k;
    }
    if (TMA) consumer_bar(); else __syncthreads();
    if (warp < D && sh.dir[warp] >= 0) {
      unsigned long long k = lane < kStreamWarps ? sh.warp_best[warp][lane] : 0ull;
      k = warp_max_u64(k);
      if (lane == 0) store_record(h, sh.dir[warp], k);
    }
    if (TMA) consumer_bar(); else __syncthreads();   // sh.dir / sh.coef are rewritten by the next item
  }
}

// second half of a point-sliced list re-vote: one CTA per listed plane reduces the global plane to its record and
// zeroes it again (no-op when the vote kernel did not slice: *sliced_flag == 0)
__global__ void __launch_bounds__(256) stream_merge_kernel(const HoughDev h, const int *__restrict__ list,
                                                           unsigned *__restrict__ gplanes, const int *__restrict__ sliced_flag) {
  __shared__ unsigned long long scratch[8];
  const int count = *sliced_flag;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int e = blockIdx.x; e < count; e += gridDim.x) {
    unsigned *g = gplanes + (size_t)e * h.plane_cells;
    unsigned long long k = 0;
    for (unsigned q = threadIdx.x; q < h.plane_cells; q += blockDim.x) {
      const unsigned v = g[q];
      if (v) {
        g[q] = 0;
        const unsigned long long kk = ((unsigned long long)v << 32) | (0xFFFFFFFFu - q);
        k = kk > k ? kk : k;
      }
    }
    k = warp_max_u64(k);
    if (lane == 0) scratch[warp] = k;
    __syncthreads();
    if (warp == 0) {
      k = lane < (int)(blockDim.x >> 5) ? scratch[lane] : 0ull;
      k = warp_max_u64(k);
      if (lane == 0) store_record(h, list ? list[e] : e, k);
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------
// K1 (sliced form): FEW directions, MANY points (granularity 0-3 on 100 k+ points, configs[4]).  A
// direction is one work item in the forms above, so fewer directions than CTAs leave SMs idle.
// Here a work item is (direction l, slice s of S): the CTA votes points [s*n/S, (s+1)*n/S) into a
// shared-memory plane exactly as above, then adds the non-zero cells to the direction's plane in
// global memory (num_x^2 u32 per direction: the accumulator of these granularities is a few MB);
// sliced_scan_kernel then reduces each global plane to its record and zeroes it again.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(kVoteThreads, 2)
vote_sliced_kernel(const HoughDev h, int slices, unsigned *__restrict__ gplanes) {
  extern __shared__ __align__(16) unsigned planes[];
  const int tid = threadIdx.x;
  const unsigned plane_cells = h.plane_cells;
  for (unsigned q = tid; q < plane_cells; q += kVoteThreads) planes[q] = 0;
  __syncthreads();
  const int n = point_count(h);
  const double kmagic = h.kmagic;
  const unsigned num_x = (unsigned)h.num_x;
  const bool no_fast = !h.fast_ok;
  const int items = h.dir_count * slices;
  for (int item = blockIdx.x; item < items; item += gridDim.x) {
    // slice-major: CTAs running side by side vote the SAME points for different directions, so a slice
    // comes from DRAM once and from L2 afterwards (4 M points do not fit L2 with their 28 B each)
    const int sl = item / h.dir_count, l = item - sl * h.dir_count;
    const int i_begin = (int)((long long)n * sl / slices), i_end = (int)((long long)n * (sl + 1) / slices);
    const FastCoef f = make_fast(h.coef[h.dir_begin + l], h.inv_s);
    constexpr int U = 4;   // independent vote chains per thread
    for (int i0 = i_begin + tid; i0 < i_end; i0 += U * kVoteThreads) {
      unsigned cell[U];
      int wi[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int i = i0 + u * kVoteThreads;
        const bool in = i < i_end;
        wi[u] = in ? h.w[i] : 0;
        const double x = in ? h.px[i] : 0.0, y = in ? h.py[i] : 0.0, z = in ? h.pz[i] : 0.0;
        bool sure = true;
        cell[u] = vote_cell_fast(f, kmagic, num_x, x, y, z, &sure);
        if (wi[u] != 0 && (!sure | no_fast))
          cell[u] = vote_cell_exact(h.coef + h.dir_begin + l, h.max_x, h.dx, h.num_x, x, y, z);
      }
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (wi[u] != 0 && cell[u] < plane_cells) atomicAdd(&planes[cell[u]], (unsigned)wi[u]);
    }
    __syncthreads();
    unsigned *g = gplanes + (size_t)l * plane_cells;
    for (unsigned q = tid; q < plane_cells; q += kVoteThreads) {
      const unsigned v = planes[q];
      if (v) { atomicAdd(&g[q], v); planes[q] = 0; }
    }
    __syncthreads();
  }
}

// one CTA per direction: (max, first max cell) of its global plane -> record; the plane is zeroed
__global__ void __launch_bounds__(256) sliced_scan_kernel(const HoughDev h, unsigned *__restrict__ gplanes) {
  __shared__ unsigned long long scratch[8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int l = blockIdx.x; l < h.dir_count; l += gridDim.x) {
    unsigned *g = gplanes + (size_t)l * h.plane_cells;
    unsigned long long k = 0;
    for (unsigned q = threadIdx.x; q < h.plane_cells; q += blockDim.x) {
      const unsigned v = g[q];
      if (v) {
        g[q] = 0;
        const unsigned long long kk = ((unsigned long long)v << 32) | (0xFFFFFFFFu - q);
        k = kk > k ? kk : k;
      }
    }
    k = warp_max_u64(k);
    if (lane == 0) scratch[warp] = k;
    __syncthreads();
    if (warp == 0) {
      k = lane < (int)(blockDim.x >> 5) ? scratch[lane] : 0ull;
      k = warp_max_u64(k);
      if (lane == 0) store_record(h, l, k);
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------
// K1 (resident form, N <= kResThreads * P, weights in {0,1}): one persistent CTA
// per SM keeps ALL live points in registers (P per thread) for its whole list of
// directions, so a vote costs 6 DFMA, ~10 integer ops and one shared-memory RED
// and no load at all.  Two planes are double-buffered: while the CTA votes
// direction k into one, it scans (and re-zeroes) the plane of direction k-1 in
// the other, so there is a single barrier per direction and no scan-only phase.
// The coefficients of direction k+1 are fetched by cp.async into shared memory
// during phase k, so no thread waits on global memory inside the loop.
// ---------------------------------------------------------------------------
constexpr int kResThreads = 512;
constexpr int kResWarps = kResThreads / 32;

// (max, first max cell) of this thread's quads of a plane, zeroing them.  Three quads are loaded
// before any is examined (independent LDS.128 in flight instead of a load -> test -> branch chain
// per quad).  Cells >= plane_cells (slack of the BOUNDED vote form) are zeroed but never counted.
// Quads are visited in ascending order and the update is strict, so bcell is the thread's first
// maximum; the warp / CTA reductions of the caller keep the lowest cell among equal maxima.
template <int THREADS>
__device__ __forceinline__ void scan_plane_local(unsigned *plane, unsigned q_begin, unsigned quads, int t, unsigned plane_cells,
                                                 unsigned &best, unsigned &bcell) {
  uint4 *p4 = reinterpret_cast<uint4 *>(plane);
  best = 0; bcell = 0;
  for (unsigned q0 = q_begin + t; q0 < quads; q0 += 3 * THREADS) {
    uint4 v[3];
#pragma unroll
    for (int u = 0; u < 3; ++u) {
      const unsigned q = q0 + u * THREADS;
      v[u] = q < quads ? p4[q] : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
    for (int u = 0; u < 3; ++u) {
      const unsigned q = q0 + u * THREADS;
      uint4 w = v[u];
      if ((w.x | w.y | w.z | w.w) == 0) continue;
      p4[q] = make_uint4(0, 0, 0, 0);
      if (4 * q + 3 >= plane_cells) {
        if (4 * q >= plane_cells) w.x = 0;
        if (4 * q + 1 >= plane_cells) w.y = 0;
        if (4 * q + 2 >= plane_cells) w.z = 0;
        w.w = 0;
      }
      const unsigned m4 = max(max(w.x, w.y), max(w.z, w.w));
      if (m4 > best) {
        best = m4;
        bcell = 4 * q + (w.x == m4 ? 0u : w.y == m4 ? 1u : w.z == m4 ? 2u : 3u);
      }
    }
  }
}

constexpr int kResG = 4;   // directions voted per barrier phase

struct ResShared {
  unsigned track_best[2][kResG][kResWarps];   // TRACK form: per-warp maxima of (count << 16 | ~cell) of the group voted in phase k (slot k&1)
  unsigned long long warp_best[2][kResG][kResWarps];
  double coef[2][kResG][6];   // raw DirCoef of the directions voted in phase k (slot k&1)
  int dir[3][kResG];          // local direction numbers of phase k (slot k%3; -1: none); the
                              // previous phase's numbers stay readable while the next are staged
  int zero_ctr[3];            // TRACK form: chunk counter of the zero-fill of phase k (slot k%3), reset two phases ahead
};

// TRACK form (needs plane_stride < 65536; N <= 6144 < 65536 always holds here): the increment returns the
// old count and every thread keeps the maximum of (old + 1) << 16 | (0xFFFF - cell) over its votes of a
// direction, so the plane of the previous group is only zero-filled, never scanned; the per-direction fixed
// cost drops from ~1600 to ~300 warp-instructions (ncu, round 1: 33 warp-instructions per warp-vote at
// 2.8 k points, of which 14 were the vote itself).
template <int P, int ABLATE, bool BOUNDED = false, bool TRACK = false>
__device__ __forceinline__ void vote_resident(const HoughDev &h, const int *__restrict__ list, int count,
                                              unsigned *planes, ResShared &sh, int cta, int n_cta) {
  constexpr int G = kResG;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const unsigned stride = h.plane_stride, quads = stride / 4;
  const unsigned planes_sa = (unsigned)__cvta_generic_to_shared(planes);
  const unsigned coef_sa = (unsigned)__cvta_generic_to_shared(&sh.coef[0][0][0]);
  // points are COMPACT here (every weight +1): slot p of thread t is point t + p*512, so all
  // slots below n/512 are full for every thread and only one slot is partially filled
  const int n = point_count(h);
  double px[P], py[P], pz[P];
  unsigned live = 0;
#pragma unroll
  for (int p = 0; p < P; ++p) {
    const int i = tid + p * kResThreads;
    const bool ok = i < n;
    px[p] = ok ? h.px[i] : 0.0; py[p] = ok ? h.py[i] : 0.0; pz[p] = ok ? h.pz[i] : 0.0;
    live |= ok ? (1u << p) : 0u;
  }
  // per WARP: slots below full_w hold 32 live lanes, slots in [full_w, used_w) are partly filled,
  // slots from used_w on are empty for the whole warp and are skipped (warp-uniform branches)
  const int w0 = warp * 32;
  const int used_w = n > w0 ? (n - w0 + kResThreads - 1) / kResThreads : 0;
  const int full_w = n >= w0 + 32 ? (n - w0 - 32) / kResThreads + 1 : 0;
  const unsigned dummy_sa = planes_sa + 2u * G * stride * 4u;   // pad word: votes that must not count
  // thread g < G stages direction g of a group: list entry -> dir[], DirCoef -> coef[] (cp.async)
  auto stage = [&](int group, int slot, int dslot) {
    const int e = group * G + tid;
    int l = -1;
    if (e < count) {
      l = list ? list[e] : e;
      const double *src = reinterpret_cast<const double *>(h.coef + h.dir_begin + l);
#pragma unroll
      for (int c = 0; c < 5; ++c) cp_async8(coef_sa + 48u * (unsigned)(slot * G + tid) + 8u * c, src + c);
    }
    sh.dir[dslot][tid] = l;
  };
  if (tid == 0) { sh.zero_ctr[0] = 0; sh.zero_ctr[1] = 0; sh.zero_ctr[2] = 0; }
  if (tid < G) { stage(cta, 0, 0); cp_async_wait_all(); }
  __syncthreads();   // also orders the kernel's zero-fill of the planes
  const double kmagic = h.kmagic, inv_s = h.inv_s;
  const unsigned num_x = (unsigned)h.num_x, plane_cells = h.plane_cells, row4 = 4u * (unsigned)h.num_x;
  const bool no_fast = !h.fast_ok;   // in a register: h may live in shared memory, and the asm increments clobber memory
  unsigned sink = 0;
  bool prev_any = false;
  int k = 0;
  for (int group = cta;; group += n_cta, ++k) {
    const int slot = k & 1, dslot = k % 3, dprev = (k + 2) % 3;
    const bool have = group * G < count;
    if (tid < G) stage(group + n_cta, slot ^ 1, (k + 1) % 3);   // fetched while this group is voted
    if (have) {
#pragma unroll 1
      for (int g = 0; g < G; ++g) {
        const int l = sh.dir[dslot][g];
        if (l < 0) break;
        DirCoef raw;
        raw.c0 = sh.coef[slot][g][0]; raw.c1 = sh.coef[slot][g][1]; raw.c2 = sh.coef[slot][g][2];
        raw.bx = sh.coef[slot][g][3]; raw.by = sh.coef[slot][g][4];
        const FastCoef f = make_fast(raw, inv_s);
        const unsigned mine_sa = planes_sa + (unsigned)(slot * G + g) * stride * 4u;
        unsigned tbest = 0;     // TRACK: this thread's best (count, ~cell) key of direction g
        const unsigned lim4 = 4u * plane_cells;
        // in_plane: the increment may land where `off` says even when it is not counted (BOUNDED form with every
        // lane live: a dropped vote falls into the slack cells or, on the exact path, onto the pad word)
        auto track = [&](unsigned off, bool counted, bool in_plane) {
          const unsigned old = atom_shared_inc_ret((in_plane || counted) ? mine_sa + off : dummy_sa);
          const unsigned key = ((old + 1u) << 16) | ((off >> 2) ^ 0xFFFFu);   // == 0xFFFF - cell for a counted vote
          tbest = max(tbest, counted ? key : 0u);
        };
        constexpr int CH = 4;   // votes per batch: 8 independent DFMA chains, 4 cell registers
#pragma unroll
        for (int p0 = 0; p0 < P; p0 += CH) {
          if (p0 >= used_w) break;                        // warp-uniform
          const unsigned drop_off = dummy_sa - mine_sa;
          if ((p0 + CH < P ? p0 + CH : P) <= full_w) {    // warp-uniform: every lane of every slot is live
            // off[c]: byte offset of the vote inside the plane (>= 4*plane_cells when the reference drops it)
            unsigned off[CH];
            unsigned gmin = 0xFFFFFFFFu, hacc = 0;
#pragma unroll
            for (int c = 0; c < CH; ++c) {
              const int p = p0 + c;
              if (p >= P) { off[c] = drop_off; continue; }
              if (ABLATE == 2) off[c] = 4u * (((unsigned)(tid + p * kResThreads) * 2654435761u + (unsigned)l * 40503u) % plane_cells);
              else if (BOUNDED) off[c] = vote_off_bounded(f, kmagic, row4, px[p], py[p], pz[p], gmin);
              else off[c] = 4u * vote_cell_acc(f, kmagic, num_x, px[p], py[p], pz[p], gmin, hacc);
            }
            if (ABLATE != 2 && (batch_unsure(gmin, hacc) | no_fast)) {   // rare: within 8e-6 cell of a boundary
#pragma unroll
              for (int c = 0; c < CH; ++c) {   // unrolled: register arrays must not be indexed dynamically
                const int p = p0 + c;
                if (p < P) {
                  const unsigned cell = vote_cell_checked_h(&h, l, px[p], py[p], pz[p]);
                  off[c] = cell < plane_cells ? 4u * cell : drop_off;
                }
              }
            }
            // unconditional increments (ATOMS.POPC.INC merges lanes that hit the same cell and needs no
            // branch this way); a vote the reference drops goes to the pad word
#pragma unroll
            for (int c = 0; c < CH; ++c) {
              if (p0 + c >= P) continue;
              if (ABLATE == 1) sink ^= off[c];
              else if (TRACK) track(off[c], off[c] < lim4, BOUNDED);
              else if (BOUNDED) red_shared_inc(mine_sa + off[c]);
              else red_shared_inc(off[c] < 4u * plane_cells ? mine_sa + off[c] : dummy_sa);
            }
          } else {                                        // the batch that holds this warp's last slots
#pragma unroll
            for (int c = 0; c < CH; ++c) {
              const int p = p0 + c;
              if (p >= P) continue;
              if (p >= used_w) continue;                  // warp-uniform: nothing of this warp in the slot
              unsigned g1 = 0xFFFFFFFFu, h1 = 0, o;
              if (ABLATE == 2) o = 4u * (((unsigned)(tid + p * kResThreads) * 2654435761u + (unsigned)l * 40503u) % plane_cells);
              else if (BOUNDED) o = vote_off_bounded(f, kmagic, row4, px[p], py[p], pz[p], g1);
              else o = 4u * vote_cell_acc(f, kmagic, num_x, px[p], py[p], pz[p], g1, h1);
              if (ABLATE != 2 && (batch_unsure(g1, h1) | no_fast)) {
                const unsigned cell = vote_cell_checked_h(&h, l, px[p], py[p], pz[p]);
                o = cell < plane_cells ? 4u * cell : drop_off;
              }
              // an empty lane of a partly filled slot votes into the pad word
              if (ABLATE == 1) sink ^= o;
              else if (TRACK) track(o, (o < lim4) & ((live >> p) & 1u), false);
              else if (BOUNDED) red_shared_inc(((live >> p) & 1u) ? mine_sa + o : dummy_sa);
              else red_shared_inc(((o < 4u * plane_cells) & ((live >> p) & 1u)) ? mine_sa + o : dummy_sa);
            }
          }
        }
        if (TRACK) {
          tbest = __reduce_max_sync(0xFFFFFFFFu, tbest);
          if (lane == 0) sh.track_best[slot][g][warp] = tbest;
        }
      }
    }
    if (TRACK) {
      if (tid == 0) sh.zero_ctr[(k + 1) % 3] = 0;   // last used in phase k-2, next in phase k+1: no reader now
      if (prev_any) {
        // the previous group's planes back to zero (their maxima were tracked, not scanned).  The warps do not
        // finish their votes together (5 or 6 slots of points each), so the zero-fill is handed out in chunks of
        // kZeroChunk quads to whichever warp is free (ncu, round 2: 14 % of the samples waited at the phase barrier)
        // The G planes of a slot are contiguous; a chunk is 8 STS.128 per lane.  (ncu, round 2: with 128-quad chunks
        // addressed as (plane, offset) the bookkeeping -- a division, 64-bit address arithmetic, a rolled store
        // loop -- made the zero-fill 21 % of the kernel's instructions, 920 per direction for 49 useful stores.)
        constexpr unsigned kZeroChunk = 256;
        const unsigned total_q = (unsigned)G * quads;
        uint4 *p4 = reinterpret_cast<uint4 *>(planes + (size_t)((slot ^ 1) * G) * stride);
        for (;;) {
          unsigned c = 0;
          if (lane == 0) c = (unsigned)atomicAdd(&sh.zero_ctr[dslot], 1);
          c = __shfl_sync(0xFFFFFFFFu, c, 0);
          const unsigned q0 = c * kZeroChunk + lane;
          if (q0 - lane >= total_q) break;
#pragma unroll
          for (unsigned u = 0; u < kZeroChunk / 32; ++u)
            if (q0 + u * 32 < total_q) p4[q0 + u * 32] = make_uint4(0, 0, 0, 0);
        }
      }
      if (tid < G) cp_async_wait_all();
      __syncthreads();   // votes of this group and their per-warp maxima, zero-fill, staging of the next group: done
      if (have && warp < G) {
        const int l = sh.dir[dslot][warp];
        if (l >= 0) {
          unsigned k2 = lane < kResWarps ? sh.track_best[slot][warp][lane] : 0u;
          k2 = __reduce_max_sync(0xFFFFFFFFu, k2);
          if (lane == 0) {
            const unsigned v = k2 >> 16;
            h.dvotes[l] = v; h.dcell[l] = v ? 0xFFFFu - (k2 & 0xFFFFu) : 0u; h.dstale[l] = 0;
          }
        }
      }
      prev_any = have;
      if (!have) break;
      continue;   // track_best is double-buffered by phase parity: the next phase writes the other half
    }
    if (prev_any) {   // scan the previous group's planes: local maxima first, then 2*G independent reductions
      unsigned best[G], bcell[G];
#pragma unroll
      for (int g = 0; g < G; ++g)
        scan_plane_local<kResThreads>(planes + (size_t)((slot ^ 1) * G + g) * stride, 0u, quads, tid, plane_cells, best[g], bcell[g]);
      unsigned m[G];
#pragma unroll
      for (int g = 0; g < G; ++g) m[g] = __reduce_max_sync(0xFFFFFFFFu, best[g]);
#pragma unroll
      for (int g = 0; g < G; ++g) bcell[g] = __reduce_min_sync(0xFFFFFFFFu, best[g] == m[g] ? bcell[g] : 0xFFFFFFFFu);
      if (lane == 0) {
#pragma unroll
        for (int g = 0; g < G; ++g)
          sh.warp_best[slot ^ 1][g][warp] = ((unsigned long long)m[g] << 32) | (0xFFFFFFFFu - bcell[g]);
      }
    }
    // the previous group's direction numbers are read BEFORE the barrier: right after it lanes 0..G-1
    // of warp 0 enter phase k+1 and stage() rewrites this very slot ((k+1)+1 == k+2 mod 3), so a read
    // placed behind the barrier by a late warp 1..G-1 would race with it (a record stored under the
    // wrong direction, or skipped and left holding the previous frame's maximum)
    int prev_l = -1;
    if (prev_any && warp < G) prev_l = sh.dir[dprev][warp];
    if (tid < G) cp_async_wait_all();
    __syncthreads();   // votes of this group, scan of the previous one, staging of the next: done
    if (prev_l >= 0) {
      unsigned long long key = lane < kResWarps ? sh.warp_best[slot ^ 1][warp][lane] : 0ull;
      key = warp_max_u64(key);
      if (lane == 0) store_record(h, prev_l, key);
    }
    prev_any = have;
    if (!have) break;
  }
  if (ABLATE == 1 && sink == 0x9E3779B9u) h.dvotes[0] = sink;
}

template <int P, int ABLATE, bool BOUNDED = false>
__global__ void __launch_bounds__(kResThreads, 1)
vote_resident_kernel(const __grid_constant__ HoughDev h, const int *__restrict__ list,
                     const int *__restrict__ count_ptr, int count_imm) {
  extern __shared__ __align__(16) unsigned planes[];
  __shared__ __align__(16) ResShared sh;
  const int count = count_ptr ? *count_ptr : count_imm;
  uint4 *p4 = reinterpret_cast<uint4 *>(planes);
  for (unsigned q = threadIdx.x; q < 2 * kResG * h.plane_stride / 4 + 1; q += kResThreads) p4[q] = make_uint4(0, 0, 0, 0);
  vote_resident<P, ABLATE, BOUNDED>(h, list, count, planes, sh, blockIdx.x, gridDim.x);
}

// ---------------------------------------------------------------------------
// K2: argmax over the per-direction records.
// fresh_only = 1: maximum over exact (non-stale) records -> *out_key (atomicMax)
// fresh_only = 0: maximum over all records (valid once every stale record whose
//                 bound reaches the exact maximum has been rebuilt)
// ---------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long block_max_u64(unsigned long long k, unsigned long long *scratch) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  k = warp_max_u64(k);
  if (lane == 0) scratch[warp] = k;
  __syncthreads();
  k = lane < nw ? scratch[lane] : 0ull;
  k = warp_max_u64(k);
  __syncthreads();
  return k;
}

__global__ void argmax_kernel(const HoughDev h, int fresh_only, unsigned long long *out_key) {
  __shared__ unsigned long long scratch[32];
  unsigned long long k = 0;
  for (int l = blockIdx.x * blockDim.x + threadIdx.x; l < h.dir_count; l += gridDim.x * blockDim.x) {
    if (fresh_only && h.dstale[l]) continue;
    const unsigned long long kk = line_key(h.dvotes[l], h.dcell[l], h.dir_begin + l, h.num_b);
    k = kk > k ? kk : k;
  }
  k = block_max_u64(k, scratch);
  if (threadIdx.x == 0) atomicMax(out_key, k);
}

// stale records whose upper bound reaches the best exact maximum must be rebuilt
__global__ void worklist_kernel(const HoughDev h, const unsigned long long *fresh_key, int *list, int *count) {
  const unsigned fvotes = (unsigned)(*fresh_key >> 32);
  for (int l = blockIdx.x * blockDim.x + threadIdx.x; l < h.dir_count; l += gridDim.x * blockDim.x) {
    if (h.dstale[l] && h.dvotes[l] >= fvotes) list[atomicAdd(count, 1)] = l;
  }
}

// ---------------------------------------------------------------------------
// K4: Hough::subtract for points that are being removed.  Their votes leave
// every plane; a direction's record stays exact unless one of them voted for its
// first-maximum cell (then the record becomes an upper bound).  One thread per
// (direction, chunk of removed points); no read-modify-write at all.
// idx: indices into the point arrays of the removed points.
// ---------------------------------------------------------------------------
constexpr int kUnvoteChunk = 32;
__global__ void unvote_kernel(const HoughDev h, const int *__restrict__ idx, int n_idx) {
  const int n_chunks = (n_idx + kUnvoteChunk - 1) / kUnvoteChunk;
  const long long total = (long long)h.dir_count * n_chunks;
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total;
       t += (long long)gridDim.x * blockDim.x) {
    const int l = (int)(t % h.dir_count), chunk = (int)(t / h.dir_count);
    if (h.dstale[l]) continue;
    const unsigned target = h.dcell[l];
    if (h.dvotes[l] == 0) continue;
    const FastCoef c = make_fast(h.coef[h.dir_begin + l], h.inv_s);
    const int e0 = chunk * kUnvoteChunk, e1 = min(e0 + kUnvoteChunk, n_idx);
    bool hit = false;
    for (int e = e0; e < e1; ++e) {
      const int i = idx[e];
      if (i < 0) continue;   // a subtracted point that was never added
      hit |= vote_cell(h, c, h.dir_begin + l, h.px[i], h.py[i], h.pz[i]) == target;
    }
    if (hit) h.dstale[l] = 1;
  }
}

}  // namespace h3d
#endif  // H3D_KERNELS_CUH_
