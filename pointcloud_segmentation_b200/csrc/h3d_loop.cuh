// h3d_loop.cuh -- the iterative loop on the device (general, multi-kernel form).
//   hough3dlines()  wrap.h:172-259, 339-347  /  lib/hough3dlines.cpp:200-317
// One stream, no host round trip inside the loop: every kernel reads the loop state from
// device memory and returns at once when `done` is set, so the host can enqueue opt_nlines
// iterations blindly.  Removed points are compacted away in order, like
// PointCloud::removePoints (lib/pointcloud.cpp:109-128); the live count lives in LoopState.
#ifndef H3D_LOOP_CUH_
#define H3D_LOOP_CUH_

#include "../../include/h3d.h"
#include "h3d_fit.cuh"
#include "h3d_kernels.cuh"

namespace h3d {

struct LoopState {
  // configuration
  double opt_dx;          // inlier radius (wrap.h:245,250)
  double shift[3];        // X.shift (wrap.h:257)
  int minvotes, nlines_max, max_iters;
  unsigned pool_cap;
  // running state
  int done;
  int n_alive;            // |X|
  int n_lines, n_iters;
  int ny;                 // |Y| of the current pass
  int ny_prev;            // |Y| removed by the previous iteration (to un-vote)
  unsigned pool_used;
  int work_count;
  unsigned long long fresh_key, best_key;
  double a[3], b[3];      // current line
  unsigned long long votes_cast;
  double ua[3], ub[3];    // the line whose inlier test produced the set being removed (pass-1 fit): un-vote strip test
};

struct LoopBuffers {
  LoopState *st;
  h3d_iter *iters;        // max_iters records
  unsigned *pool;         // pool_cap original indices
  unsigned *orig;         // original index of live point i (NaN/Inf strip, removals)
  unsigned char *flags;   // inlier flags of the current pass
  int *yidx;              // ordered inlier indices of the current pass
  double *ypx, *ypy, *ypz;  // coordinates removed by the previous iteration (to un-vote)
  double *px, *py, *pz;   // mutable aliases of the point arrays (compaction)
  int *list;              // directions to rebuild
  const double *dirs;     // 3*num_b direction vectors
  // large clouds (kBigCloud points and more): chunked multi-CTA compaction / removal
  int *ccount, *coff;     // inliers per 1024-point chunk of the current pass, their exclusive prefix
  double *sx, *sy, *sz;   // scratch copy of the kept points (removal goes out of place and back)
  unsigned *so;
};

constexpr int kBigCloud = 32768;   // from here on the loop uses the chunked kernels below
constexpr int kChunk = 1024;

// ---- start of an iteration: reset the reduction targets -------------------------------
__global__ void loop_begin_kernel(LoopBuffers lb) {
  LoopState &s = *lb.st;
  if (s.done) return;
  s.fresh_key = 0; s.best_key = 0; s.work_count = 0;
}

// K4 with the removed set of the previous iteration
__global__ void loop_unvote_kernel(const HoughDev h, LoopBuffers lb) {
  const LoopState &s = *lb.st;
  if (s.done || s.ny_prev == 0) return;
  const int n_idx = s.ny_prev;
  const int n_chunks = (n_idx + kUnvoteChunk - 1) / kUnvoteChunk;
  const long long total = (long long)h.dir_count * n_chunks;
  const double ua[3] = {s.ua[0], s.ua[1], s.ua[2]}, ub[3] = {s.ub[0], s.ub[1], s.ub[2]};
  const double cell_off = h.max_x / h.dx + 0.5;
  const double reach = 1.4142136 * (0.5 + s.opt_dx / h.dx + 0.02), reach2 = reach * reach;
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total;
       t += (long long)gridDim.x * blockDim.x) {
    const int l = (int)(t % h.dir_count), chunk = (int)(t / h.dir_count);
    if (h.dstale[l] || h.dvotes[l] == 0) continue;
    const unsigned target = h.dcell[l];
    const FastCoef c = make_fast(h.coef[h.dir_begin + l], h.inv_s);
    {
      // O(1) reject (as in the fused loop): the removed set lies within opt_dx of the line (ua, ub); a record whose
      // max cell is farther than `reach` cells from that line's image in the plane cannot have lost a vote.  Edge
      // cells, which the reference's flat-index aliasing can reach, are always tested.
      const unsigned num_x = (unsigned)h.num_x, cx = target / num_x, cy = target - cx * num_x;
      if (cy != 0 && cy != num_x - 1) {
        double ilo, ihi;
        if (!strip_interval(c, ua, ub, cell_off, reach2, cx, cy, &ilo, &ihi)) continue;
      }
    }
    const int e0 = chunk * kUnvoteChunk, e1 = min(e0 + kUnvoteChunk, n_idx);
    bool hit = false;
    for (int e = e0; e < e1; ++e) {
      hit |= vote_cell(h, c, h.dir_begin + l, lb.ypx[e], lb.ypy[e], lb.ypz[e]) == target;
    }
    if (hit) h.dstale[l] = 1;
  }
}

__global__ void loop_argmax_kernel(const HoughDev h, LoopBuffers lb, int fresh_only) {
  __shared__ unsigned long long scratch[32];
  LoopState &s = *lb.st;
  if (s.done) return;
  unsigned long long k = 0;
  for (int l = blockIdx.x * blockDim.x + threadIdx.x; l < h.dir_count; l += gridDim.x * blockDim.x) {
    if (fresh_only && h.dstale[l]) continue;
    const unsigned long long kk = line_key(h.dvotes[l], h.dcell[l], h.dir_begin + l, h.num_b);
    k = kk > k ? kk : k;
  }
  k = block_max_u64(k, scratch);
  if (threadIdx.x == 0) atomicMax(fresh_only ? &s.fresh_key : &s.best_key, k);
}

__global__ void loop_worklist_kernel(const HoughDev h, LoopBuffers lb) {
  LoopState &s = *lb.st;
  if (s.done) return;
  const unsigned fvotes = (unsigned)(s.fresh_key >> 32);
  for (int l = blockIdx.x * blockDim.x + threadIdx.x; l < h.dir_count; l += gridDim.x * blockDim.x)
    if (h.dstale[l] && h.dvotes[l] >= fvotes) lb.list[atomicAdd(&s.work_count, 1)] = l;
}

// inverse map of Hough::getLine, lib/hough.cpp:114-132 (same conversions, no FMA)
__device__ inline void line_from_index(const HoughDev &h, const double *dirs, unsigned index, double a[3],
                                       double b[3]) {
  const size_t num_x = (size_t)h.num_x, num_b = (size_t)h.num_b;
  double x = (int)(index / (num_x * num_b));
  index -= (int)x * num_x * num_b;
  x = __dsub_rn(__dmul_rn(x, h.dx), h.max_x);
  double y = (int)index / num_b;
  index -= (int)y * num_b;
  y = __dsub_rn(__dmul_rn(y, h.dx), h.max_x);
  const double bx = dirs[3 * (size_t)index], by = dirs[3 * (size_t)index + 1], bz = dirs[3 * (size_t)index + 2];
  b[0] = bx; b[1] = by; b[2] = bz;
  const double opz = __dadd_rn(1.0, bz);
  const double kxx = __ddiv_rn(__dmul_rn(bx, bx), opz), kxy = __ddiv_rn(__dmul_rn(bx, by), opz);
  const double kyy = __ddiv_rn(__dmul_rn(by, by), opz);
  a[0] = __dsub_rn(__dmul_rn(x, __dsub_rn(1.0, kxx)), __dmul_rn(y, kxy));
  a[1] = __dadd_rn(__dmul_rn(x, -kxy), __dmul_rn(y, __dsub_rn(1.0, kyy)));
  a[2] = __dsub_rn(__dmul_rn(-x, bx), __dmul_rn(y, by));
}

// Hough::getLine result of this iteration -> current line + record
__global__ void loop_line_kernel(const HoughDev h, LoopBuffers lb) {
  LoopState &s = *lb.st;
  if (s.done || threadIdx.x != 0) return;
  const unsigned votes = (unsigned)(s.best_key >> 32);
  const unsigned flat = votes ? 0xFFFFFFFFu - (unsigned)s.best_key : 0u;
  line_from_index(h, lb.dirs, flat, s.a, s.b);
  s.votes_cast += (unsigned long long)s.work_count * (unsigned long long)s.n_alive;
  if (s.n_iters < s.max_iters) {
    h3d_iter &r = lb.iters[s.n_iters];
    r.votes = votes; r.status = 0; r.argmax = flat;
    for (int c = 0; c < 3; ++c) { r.a_raw[c] = s.a[c]; r.b_raw[c] = s.b[c]; }
    r.rc1 = 0; r.rc2 = 0; r.n1 = 0; r.n2 = 0; r.off1 = s.pool_used; r.off2 = s.pool_used;
    for (int c = 0; c < 3; ++c) { r.a1[c] = 0; r.b1[c] = 0; r.a2[c] = 0; r.b2[c] = 0; }
  }
}

// K3a: inlier flags of the live points against the current line
__global__ void loop_mask_kernel(const HoughDev h, LoopBuffers lb) {
  const LoopState &s = *lb.st;
  if (s.done) return;
  const double a[3] = {s.a[0], s.a[1], s.a[2]}, b[3] = {s.b[0], s.b[1], s.b[2]};
  const double dx = s.opt_dx;
  const int n = s.n_alive;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    lb.flags[i] = point_near_line(h.px[i], h.py[i], h.pz[i], a, b, dx) ? 1 : 0;
}

// K3b: ordered compaction of the flags into yidx (one CTA walks the array in chunks), and the
// inlier list of this pass into the pool (original numbering)
__global__ void __launch_bounds__(1024, 1) loop_compact_kernel(const HoughDev h, LoopBuffers lb, int pass) {
  __shared__ int warp_sum[32];
  __shared__ int base_s;
  LoopState &s = *lb.st;
  if (s.done) return;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) base_s = 0;
  __syncthreads();
  const unsigned pool0 = s.pool_used;
  const int n = s.n_alive;
  for (int c0 = 0; c0 < n; c0 += 1024) {
    const int i = c0 + tid;
    const bool keep = i < n && lb.flags[i];
    const unsigned bal = __ballot_sync(0xFFFFFFFFu, keep);
    if (lane == 0) warp_sum[warp] = __popc(bal);
    __syncthreads();
    int off = base_s;
    for (int q = 0; q < warp; ++q) off += warp_sum[q];
    if (keep) {
      const int pos = off + __popc(bal & ((1u << lane) - 1u));
      lb.yidx[pos] = i;
      lb.ypx[pos] = lb.px[i]; lb.ypy[pos] = lb.py[i]; lb.ypz[pos] = lb.pz[i];   // see big_scatter_kernel
      if (pool0 + (unsigned)pos < s.pool_cap) lb.pool[pool0 + pos] = lb.orig[i];
    }
    __syncthreads();
    if (tid == 0) { int t = 0; for (int q = 0; q < 32; ++q) t += warp_sum[q]; base_s += t; }
    __syncthreads();
  }
  if (tid == 0) {
    const int ny = base_s;
    s.ny = ny;
    const unsigned used = pool0 + (unsigned)ny;
    s.pool_used = used < s.pool_cap ? used : s.pool_cap;
    if (s.n_iters < s.max_iters) {
      h3d_iter &r = lb.iters[s.n_iters];
      if (pass == 1) { r.n1 = (unsigned)ny; r.off1 = pool0; r.off2 = s.pool_used; }
      else { r.n2 = (unsigned)ny; r.off2 = pool0; }
    }
  }
}

// pass 1: first LSQ (break on rc == 0); pass 2: minvotes test, second LSQ, accept the line
__global__ void __launch_bounds__(192, 1) loop_lsq_kernel(const HoughDev h, LoopBuffers lb, int pass) {
  __shared__ LsqShared sh;
  __shared__ LsqOut out;
  LoopState &s = *lb.st;
  if (s.done) return;
  const bool rec_ok = s.n_iters < s.max_iters;
  h3d_iter *r = rec_ok ? &lb.iters[s.n_iters] : nullptr;
  if (pass == 2 && (unsigned)s.ny < (unsigned)s.minvotes) {   // wrap.h:251-252
    __syncthreads();
    if (threadIdx.x == 0) { if (r) r->status = 2; s.n_iters++; s.done = 1; }
    return;
  }
  extern __shared__ __align__(16) float lsq_stage[];   // 3 * kLsqStageCap floats
  lsq_cta(lb.ypx, lb.ypy, lb.ypz, s.ny, sh, &out, nullptr, lsq_stage, kLsqStageCap);   // compact Y
  if (threadIdx.x != 0) return;
  if (pass == 2) for (int c = 0; c < 3; ++c) { s.ua[c] = s.a[c]; s.ub[c] = s.b[c]; }   // Y lies within opt_dx of THIS line
  for (int c = 0; c < 3; ++c) { s.a[c] = out.a[c]; s.b[c] = out.b[c]; }
  if (pass == 1) {
    if (r) { for (int c = 0; c < 3; ++c) { r->a1[c] = out.a[c]; r->b1[c] = out.b[c]; } r->rc1 = out.rc; }
    if (out.rc == 0.0) { if (r) r->status = 1; s.n_iters++; s.done = 1; }   // wrap.h:248
  } else {
    if (r) { for (int c = 0; c < 3; ++c) { r->a2[c] = out.a[c]; r->b2[c] = out.b[c]; } r->rc2 = out.rc; }
    if (out.rc == 0.0) { if (r) r->status = 3; s.n_iters++; s.done = 1; return; }   // wrap.h:255
    if (r) for (int c = 0; c < 3; ++c) r->a2[c] = __dadd_rn(out.a[c], s.shift[c]);   // wrap.h:257
    s.n_lines++;
  }
}

// end of an accepted iteration: X.removePoints(Y) (wrap.h:339) as a stable in-place compaction
// of the point arrays (the flags of pass 2 mark Y), keeping the removed coordinates for the
// next iteration's un-vote; then the loop condition (wrap.h:341-342).  One CTA, chunked.
__global__ void __launch_bounds__(1024, 1) loop_remove_kernel(LoopBuffers lb) {
  __shared__ int warp_keep[32], warp_drop[32];
  __shared__ int base_keep, base_drop;
  LoopState &s = *lb.st;
  if (s.done) return;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) { base_keep = 0; base_drop = 0; }
  __syncthreads();
  const int n = s.n_alive;
  for (int c0 = 0; c0 < n; c0 += 1024) {
    const int i = c0 + tid;
    const bool in = i < n;
    const bool drop = in && lb.flags[i];
    const bool keep = in && !drop;
    double x = 0, y = 0, z = 0;
    unsigned o = 0;
    if (in) { x = lb.px[i]; y = lb.py[i]; z = lb.pz[i]; o = lb.orig[i]; }
    const unsigned bk = __ballot_sync(0xFFFFFFFFu, keep), bd = __ballot_sync(0xFFFFFFFFu, drop);
    if (lane == 0) { warp_keep[warp] = __popc(bk); warp_drop[warp] = __popc(bd); }
    __syncthreads();   // every thread of the chunk holds its point in registers from here on
    int ok = base_keep, od = base_drop;
    for (int q = 0; q < warp; ++q) { ok += warp_keep[q]; od += warp_drop[q]; }
    const unsigned below = (1u << lane) - 1u;
    if (keep) { const int p = ok + __popc(bk & below); lb.px[p] = x; lb.py[p] = y; lb.pz[p] = z; lb.orig[p] = o; }
    if (drop) { const int p = od + __popc(bd & below); lb.ypx[p] = x; lb.ypy[p] = y; lb.ypz[p] = z; }
    __syncthreads();
    if (tid == 0) {
      int tk = 0, td = 0;
      for (int q = 0; q < 32; ++q) { tk += warp_keep[q]; td += warp_drop[q]; }
      base_keep += tk; base_drop += td;
    }
    __syncthreads();
  }
  if (tid == 0) {
    s.n_alive = base_keep;
    s.ny_prev = base_drop;
    s.n_iters++;
    if (!((s.n_alive > 1) && (s.nlines_max == 0 || s.nlines_max > s.n_lines))) s.done = 1;
  }
}

// ---------------------------------------------------------------------------------------------
// The same K3 / removal steps for LARGE clouds (configs[3]: 1 M points): the single-CTA kernels above
// walk the cloud chunk by chunk, which is fine for the few thousand points of a ToF frame and far too
// serial for a million.  Here every 1024-point chunk is one CTA's work item: flags + per-chunk counts,
// a one-CTA scan of the (n/1024) counts, an ordered scatter.  Same order, same results.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kChunk) big_mask_kernel(const HoughDev h, LoopBuffers lb) {
  const LoopState &s = *lb.st;
  if (s.done) return;
  const double a[3] = {s.a[0], s.a[1], s.a[2]}, b[3] = {s.b[0], s.b[1], s.b[2]};
  const double dx = s.opt_dx;
  const int n = s.n_alive, nchunks = (n + kChunk - 1) / kChunk;
  for (int c = blockIdx.x; c < nchunks; c += gridDim.x) {
    const int i = c * kChunk + threadIdx.x;
    const bool keep = i < n && point_near_line(h.px[i], h.py[i], h.pz[i], a, b, dx);
    if (i < n) lb.flags[i] = keep ? 1 : 0;
    const int cnt = __syncthreads_count(keep);
    if (threadIdx.x == 0) lb.ccount[c] = cnt;
  }
}

// exclusive prefix of the chunk counts + the bookkeeping of loop_compact_kernel's tail
__global__ void __launch_bounds__(1024, 1) big_scan_kernel(LoopBuffers lb, int pass) {
  __shared__ int warp_sum[32];
  __shared__ int base_s;
  LoopState &s = *lb.st;
  if (s.done) return;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) base_s = 0;
  __syncthreads();
  const int n = s.n_alive, nchunks = (n + kChunk - 1) / kChunk;
  for (int c0 = 0; c0 < nchunks; c0 += 1024) {
    const int c = c0 + tid;
    const int v = c < nchunks ? lb.ccount[c] : 0;
    int incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(0xFFFFFFFFu, incl, d); if (lane >= d) incl += t; }
    if (lane == 31) warp_sum[warp] = incl;
    __syncthreads();
    int off = base_s;
    for (int q = 0; q < warp; ++q) off += warp_sum[q];
    if (c < nchunks) lb.coff[c] = off + incl - v;
    __syncthreads();
    if (tid == 0) { int t = 0; for (int q = 0; q < 32; ++q) t += warp_sum[q]; base_s += t; }
    __syncthreads();
  }
  if (tid == 0 && pass != 0) {
    const unsigned pool0 = s.pool_used;
    const int ny = base_s;
    s.ny = ny;
    lb.coff[nchunks] = (int)pool0;   // where this pass's list starts in the pool (read by the scatter)
    const unsigned used = pool0 + (unsigned)ny;
    s.pool_used = used < s.pool_cap ? used : s.pool_cap;
    if (s.n_iters < s.max_iters) {
      h3d_iter &r = lb.iters[s.n_iters];
      if (pass == 1) { r.n1 = (unsigned)ny; r.off1 = pool0; r.off2 = s.pool_used; }
      else { r.n2 = (unsigned)ny; r.off2 = pool0; }
    }
  }
}

// position of a flagged thread among the flagged threads of its 1024-thread chunk, in thread order
__device__ __forceinline__ int chunk_rank(bool flag, int *warp_sum) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned bal = __ballot_sync(0xFFFFFFFFu, flag);
  __syncthreads();   // warp_sum of the previous use has been read
  if (lane == 0) warp_sum[warp] = __popc(bal);
  __syncthreads();
  int off = 0;
  for (int q = 0; q < warp; ++q) off += warp_sum[q];
  return off + __popc(bal & ((1u << lane) - 1u));
}

__global__ void __launch_bounds__(kChunk) big_scatter_kernel(LoopBuffers lb) {
  __shared__ int warp_sum[32];
  const LoopState &s = *lb.st;
  if (s.done) return;
  const int n = s.n_alive, nchunks = (n + kChunk - 1) / kChunk;
  const unsigned pool0 = (unsigned)lb.coff[nchunks];
  for (int c = blockIdx.x; c < nchunks; c += gridDim.x) {
    if (lb.ccount[c] == 0) continue;   // uniform
    const int i = c * kChunk + threadIdx.x;
    const bool keep = i < n && lb.flags[i];
    const int r = chunk_rank(keep, warp_sum);
    if (keep) {
      const int pos = lb.coff[c] + r;
      lb.yidx[pos] = i;
      // Y's coordinates, compact and in order: orthogonal_LSQ reads them coalesced (gathering px[yidx[k]] inside its
      // reduction chains cost two dependent L2 round trips per element: 0.42 ms per fit of an 11 k-point line).
      // The un-vote of this iteration has already consumed the previous Y from these arrays.
      lb.ypx[pos] = lb.px[i]; lb.ypy[pos] = lb.py[i]; lb.ypz[pos] = lb.pz[i];
      if (pool0 + (unsigned)pos < s.pool_cap) lb.pool[pool0 + pos] = lb.orig[i];
    }
  }
}

// X.removePoints(Y): kept points to the scratch arrays (ordered), removed coordinates to ypx/ypy/ypz;
// ccount / coff still describe the flags of pass 2
__global__ void __launch_bounds__(kChunk) big_remove_kernel(LoopBuffers lb) {
  __shared__ int warp_sum[32];
  const LoopState &s = *lb.st;
  if (s.done) return;
  const int n = s.n_alive, nchunks = (n + kChunk - 1) / kChunk;
  for (int c = blockIdx.x; c < nchunks; c += gridDim.x) {
    const int i = c * kChunk + threadIdx.x;
    const bool in = i < n;
    const bool drop = in && lb.flags[i];
    const bool keep = in && !drop;
    const int doff = lb.coff[c], koff = c * kChunk - doff;
    double x = 0, y = 0, z = 0;
    unsigned o = 0;
    if (in) { x = lb.px[i]; y = lb.py[i]; z = lb.pz[i]; o = lb.orig[i]; }
    if (lb.ccount[c] == 0) {           // uniform: nothing of this chunk is removed
      if (in) { const int p = koff + threadIdx.x; lb.sx[p] = x; lb.sy[p] = y; lb.sz[p] = z; lb.so[p] = o; }
      continue;
    }
    const int rk = chunk_rank(keep, warp_sum);
    const int rd = chunk_rank(drop, warp_sum);
    if (keep) { const int p = koff + rk; lb.sx[p] = x; lb.sy[p] = y; lb.sz[p] = z; lb.so[p] = o; }
    if (drop) { const int p = doff + rd; lb.ypx[p] = x; lb.ypy[p] = y; lb.ypz[p] = z; }
  }
}

__global__ void big_copyback_kernel(LoopBuffers lb) {
  const LoopState &s = *lb.st;
  if (s.done) return;
  const int nk = s.n_alive - s.ny;   // ny = |Y| of pass 2
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nk; i += gridDim.x * blockDim.x) {
    lb.px[i] = lb.sx[i]; lb.py[i] = lb.sy[i]; lb.pz[i] = lb.sz[i]; lb.orig[i] = lb.so[i];
  }
}

// the loop condition (wrap.h:341-342), after the copy-back has read n_alive
__global__ void big_finish_kernel(LoopBuffers lb) {
  LoopState &s = *lb.st;
  if (s.done) return;
  s.ny_prev = s.ny;
  s.n_alive = s.n_alive - s.ny;
  s.n_iters++;
  if (!((s.n_alive > 1) && (s.nlines_max == 0 || s.nlines_max > s.n_lines))) s.done = 1;
}

}  // namespace h3d
#endif  // H3D_LOOP_CUH_
