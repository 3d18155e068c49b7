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
    const uint4 v = p4[q];
    if ((v.x | v.y | v.z | v.w) == 0) continue;
    p4[q] = make_uint4(0, 0, 0, 0);
    const unsigned c0 = 4 * q;
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
    for (int i0 = tid; i0 < h.n; i0 += U * kVoteThreads) {
      double x[U], y[U], z[U];
      int wi[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int i = i0 + u * kVoteThreads;
        const bool in = i < h.n;
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
// K1 (resident form, N <= kResThreads * P, weights in {0,1}): one persistent CTA
// per SM keeps ALL live points in registers (P per thread) for its whole list of
// directions, so a vote costs 6 DFMA, ~12 integer ops and one shared atomic and
// no load at all.  Two planes are double-buffered: while the CTA votes direction
// k into one, it scans (and re-zeroes) the plane of direction k-1 in the other,
// so there is a single barrier per direction and no scan-only phase.
// ---------------------------------------------------------------------------
constexpr int kResThreads = 512;
constexpr int kResWarps = kResThreads / 32;

template <int P, int ABLATE>
__device__ __forceinline__ void vote_resident(const HoughDev &h, const int *__restrict__ list, int count,
                                              unsigned *planes, unsigned long long (*warp_best)[32], int cta,
                                              int n_cta) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const unsigned stride = h.plane_stride;
  double px[P], py[P], pz[P];
  unsigned live = 0;
#pragma unroll
  for (int p = 0; p < P; ++p) {
    const int i = tid + p * kResThreads;
    const bool ok = i < h.n && h.w[i] != 0;
    px[p] = ok ? h.px[i] : 0.0; py[p] = ok ? h.py[i] : 0.0; pz[p] = ok ? h.pz[i] : 0.0;
    live |= ok ? (1u << p) : 0u;
  }
  unsigned sink = 0;
  int prev = -1;     // local direction whose plane waits to be scanned
  int k = 0;
  for (int e = cta;; e += n_cta, ++k) {
    const bool have = e < count;
    unsigned *mine = planes + (k & 1) * stride;
    unsigned *other = planes + ((k & 1) ^ 1) * stride;
    int l = -1;
    if (have) {
      l = list ? list[e] : e;
      const int j = h.dir_begin + l;
      const FastCoef f = make_fast(h.coef[j], h.inv_s);
      const double kmagic = h.kmagic;
      const unsigned num_x = (unsigned)h.num_x, plane_cells = h.plane_cells;
      unsigned cell[P];
      unsigned unsure = 0;
#pragma unroll
      for (int p = 0; p < P; ++p) {   // branch-free: P independent chains of 3 DFMA per axis
        bool sure = true;
        if (ABLATE == 2) cell[p] = ((unsigned)(tid + p * kResThreads) * 2654435761u + (unsigned)l * 40503u) % plane_cells;
        else cell[p] = vote_cell_fast(f, kmagic, num_x, px[p], py[p], pz[p], &sure);
        unsure |= sure ? 0u : (1u << p);
      }
      unsure = h.fast_ok ? (unsure & live) : live;
      if (ABLATE != 2 && unsure) {   // rare: a projection within 8e-6 cell of a boundary
#pragma unroll
        for (int p = 0; p < P; ++p)
          if (unsure & (1u << p)) cell[p] = vote_cell_exact(h.coef + j, h.max_x, h.dx, h.num_x, px[p], py[p], pz[p]);
      }
#pragma unroll
      for (int p = 0; p < P; ++p) {
        if (ABLATE == 1) sink ^= cell[p];
        else if ((live & (1u << p)) && cell[p] < plane_cells) atomicAdd(&mine[cell[p]], 1u);
      }
    }
    if (prev >= 0) {
      const unsigned long long key = scan_plane<kResThreads>(h, other, tid, h.dir_begin + prev, nullptr);
      if (lane == 0) warp_best[(k & 1) ^ 1][warp] = key;
    }
    __syncthreads();
    if (prev >= 0 && warp == 0) {
      unsigned long long key = lane < kResWarps ? warp_best[(k & 1) ^ 1][lane] : 0ull;
      key = warp_max_u64(key);
      if (lane == 0) store_record(h, prev, key);
    }
    prev = l;
    if (!have) break;
  }
  if (ABLATE == 1 && sink == 0x9E3779B9u) h.dvotes[0] = sink;
}

template <int P, int ABLATE>
__global__ void __launch_bounds__(kResThreads, 1)
vote_resident_kernel(const HoughDev h, const int *__restrict__ list, const int *__restrict__ count_ptr,
                     int count_imm) {
  extern __shared__ __align__(16) unsigned planes[];
  __shared__ unsigned long long warp_best[2][32];
  const int count = count_ptr ? *count_ptr : count_imm;
  uint4 *p4 = reinterpret_cast<uint4 *>(planes);
  for (unsigned q = threadIdx.x; q < 2 * h.plane_stride / 4; q += kResThreads) p4[q] = make_uint4(0, 0, 0, 0);
  __syncthreads();
  vote_resident<P, ABLATE>(h, list, count, planes, warp_best, blockIdx.x, gridDim.x);
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
