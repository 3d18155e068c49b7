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

// ---------------------------------------------------------------------------
// K1: build the planes of a list of directions in shared memory.
//
// A work item is a group of D consecutive entries of `list` (local direction
// numbers; identity when list == nullptr).  The CTA zeroes D planes once, then
// per item: every thread projects its points on the D directions (coefficients
// in registers, points streamed coalesced from the SoA arrays, which stay in
// L1/L2), increments shared-memory cells, and after one barrier the planes are
// scanned with uint4 loads for (max, first max cell) and zeroed in the same
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
    DirCoef c[D];
#pragma unroll
    for (int d = 0; d < D; ++d) {
      const int e = item * D + d;
      dl[d] = e < count ? (list ? list[e] : e) : -1;
      c[d] = h.coef[h.dir_begin + (dl[d] < 0 ? 0 : dl[d])];
    }

    for (int i = tid; i < h.n; i += kVoteThreads) {
      const int wi = h.w[i];
      if (wi == 0) continue;
      const double x = h.px[i], y = h.py[i], z = h.pz[i];
#pragma unroll
      for (int d = 0; d < D; ++d) {
        if (dl[d] < 0) continue;
        unsigned cell;
        if (ABLATE == 2) {
          cell = ((unsigned)i * 2654435761u + (unsigned)dl[d] * 40503u) % h.plane_cells;
        } else {
          cell = vote_cell(h, c[d], x, y, z);
        }
        if (ABLATE == 1) {
          sink ^= cell;
        } else if (cell != 0xFFFFFFFFu) {
          atomicAdd(&planes[d * stride + cell], (unsigned)wi);
        }
      }
    }
    __syncthreads();

    // scan + zero; thread-local first maximum over an ascending cell sequence
#pragma unroll
    for (int d = 0; d < D; ++d) {
      if (dl[d] < 0) continue;   // uniform across the CTA
      uint4 *p4 = reinterpret_cast<uint4 *>(planes + d * stride);
      unsigned best = 0, bcell = 0;
      const int j = h.dir_begin + dl[d];
      for (unsigned q = tid; q < stride / 4; q += kVoteThreads) {
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
      if (lane == 0) warp_best[d][warp] = ((unsigned long long)m << 32) | (0xFFFFFFFFu - cm);
    }
    __syncthreads();
    if (warp < D && dl[warp] >= 0) {
      unsigned long long k = lane < kVoteWarps ? warp_best[warp][lane] : 0ull;
      k = warp_max_u64(k);
      if (lane == 0) {
        const unsigned votes = (unsigned)(k >> 32);
        h.dvotes[dl[warp]] = votes;
        h.dcell[dl[warp]] = votes ? 0xFFFFFFFFu - (unsigned)k : 0u;
        h.dstale[dl[warp]] = 0;
      }
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
    const DirCoef c = h.coef[h.dir_begin + l];
    const int e0 = chunk * kUnvoteChunk, e1 = min(e0 + kUnvoteChunk, n_idx);
    bool hit = false;
    for (int e = e0; e < e1; ++e) {
      const int i = idx[e];
      if (i < 0) continue;   // a subtracted point that was never added
      hit |= vote_cell(h, c, h.px[i], h.py[i], h.pz[i]) == target;
    }
    if (hit) h.dstale[l] = 1;
  }
}

}  // namespace h3d
#endif  // H3D_KERNELS_CUH_
