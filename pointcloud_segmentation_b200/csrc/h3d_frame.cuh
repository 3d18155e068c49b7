// h3d_frame.cuh -- the per-frame fast path of the pipeline (sm_100a): everything that happens before
// Hough::add in ONE launch, with no host round trip in the middle of a frame.
//
//   cloudFiltering()  node.cpp:386-421       PassThrough x,y,z + VoxelGrid (K0)
//   hough3dlines()    wrap.h:175-226         NaN/Inf strip + widen, getMinMax3D, shiftToOrigin,
//                                            the d / opt_dx guards, new Hough(minP, maxP, dx)
//
// Round 1 ran this as eleven launches and read the bounding box back to run the Hough constructor on
// the host.  Here one thread-block CLUSTER (8 CTAs x 1024 threads, hardware cluster barriers instead of
// kernel boundaries) does the counting sort by voxel key, the centroids, the widening, the bounding box,
// the shift and the constructor's arithmetic, and leaves a FrameDev descriptor in device memory that the
// vote and loop kernels read instead of host-built launch parameters.
//
// Voxel order.  VoxelGrid emits one centroid per voxel in ascending voxel index
// i0 + i1*div0 + i2*div0*div1 with ijk = floor(p*inv_leaf) - min_b taken from the cloud's bounding box.
// That order is the lexicographic order of (i2, i1, i0), which does not depend on min_b / div_b: the grid
// of the WINDOW (known when the pipeline is created) gives the same order with no bounding-box pass.
// floor(p*inv_leaf) - (float)min_b is exact in float for both origins (small integers).
//
// The voxel histogram count[] is zero when the kernel starts and zero again when it ends (the scatter
// takes the counts back down), so no memset launch is needed between frames.
#ifndef H3D_FRAME_CUH_
#define H3D_FRAME_CUH_

#include <cooperative_groups.h>

#include "h3d_loop.cuh"
#include "h3d_prefilter.cuh"

namespace h3d {

// ---- Hough::Hough discretisation (lib/hough.cpp:26-43), one implementation for host and device ----------
struct HoughGeom {
  int rc;                 // 0 ok; 1 num_x out of range; 2 accumulator would have >= 2^32 cells
  int num_x;
  unsigned plane_cells, plane_stride;
  int fast_ok, bounded;
  double max_x, dx, inv_s, kmagic;
};

__host__ __device__ inline double geom_round_to_nearest(double v) {   // lib/hough.cpp:22-24
  return (v > 0.0) ? floor(v + 0.5) : ceil(v - 0.5);
}

// minP / maxP: corners of the SHIFTED bounding box; var_dx == 0 means range/64 (lib/hough.cpp:36-38).
// max_x, dx and num_x follow the reference expression by expression (no FMA: this file is compiled with
// -fmad=false / -ffp-contract=off); inv_s and kmagic only parametrise the guarded fast cell path.
__host__ __device__ inline HoughGeom hough_geometry(const double minP[3], const double maxP[3], double var_dx,
                                                    unsigned long long num_b) {
  HoughGeom g;
  g.rc = 0; g.num_x = 0; g.plane_cells = 0; g.plane_stride = 4; g.fast_ok = 0; g.bounded = 0;
  g.inv_s = 0.0; g.kmagic = 0.0;
  const double nmax = sqrt((maxP[0] * maxP[0]) + (maxP[1] * maxP[1]) + (maxP[2] * maxP[2]));
  const double nmin = sqrt((minP[0] * minP[0]) + (minP[1] * minP[1]) + (minP[2] * minP[2]));
  const double max_x = nmax < nmin ? nmin : nmax;
  const double range_x = 2 * max_x;
  double dx = var_dx;
  if (dx == 0.0) dx = range_x / 64.0;
  g.max_x = max_x; g.dx = dx;
  const double nxd = geom_round_to_nearest(range_x / dx);
  if (!(nxd >= 0.0) || nxd > 65535.0) { g.rc = 1; return g; }
  const unsigned long long num_x = (unsigned long long)nxd;
  if (num_x * num_x * num_b >= (1ull << 32)) { g.rc = 2; return g; }
  g.num_x = (int)num_x;
  g.plane_cells = (unsigned)(num_x * num_x);
  g.plane_stride = (g.plane_cells + 2u * (unsigned)num_x + 4u + 3u) & ~3u;   // + slack cells (BOUNDED vote form)
  g.fast_ok = (num_x >= 1 && num_x <= 4000 && dx > 0.0 && max_x == max_x && max_x <= 1.7e308) ? 1 : 0;
  if (g.fast_ok) {
    g.inv_s = 1048576.0 / dx;                                  // 2^20 / dx
    const double bias = rint(fma(max_x, g.inv_s, 524288.0));   // round(max_x * 2^20/dx + 2^19)
    g.kmagic = 6755399441055744.0 + bias;                      // 1.5 * 2^52 + bias (exact: bias < 2^32)
    if (!(bias >= 0.0 && bias < 4294967296.0)) g.fast_ok = 0;
  }
  // box symmetric about the origin (PointCloud::shiftToOrigin): with every point inside it,
  // |x'| <= max_x (1 + 1e-9), so 0 <= x_i, y_i <= num_x + 1 and the fast path needs no range test
  g.bounded = (g.fast_ok && num_x < 2000) ? 1 : 0;
  for (int c = 0; c < 3; ++c)
    if (!(fabs(minP[c] + maxP[c]) <= 1e-9 * max_x)) g.bounded = 0;
  return g;
}

// ---- what the frame's first kernel leaves for the others (device memory; exported to the host at the end) --
struct FrameDev {
  int status;             // 0: voted and looped on the fast path; 1: hough3dlines() returns 1 (wrap.h:202-212);
                          // 2: the frame does not fit the fast path (host re-runs it on the general path)
  int n;                  // points after the pre-filter (== after the NaN/Inf strip: centroids are finite)
  int n_in_window, pad0;
  HoughGeom g;
  double d;               // (maxP - minP).norm(), wrap.h:201
  double ns[3];           // what shiftToOrigin subtracted = X.shift
  double mn[3], mx[3];    // bounding box before the shift
  double cell_off;        // max_x/dx + 0.5 (strip test of the un-vote)
  int nb_planes, n_cap;   // shared-memory layout of the loop kernel for this frame
  int bb[6];              // scratch: ordered-int float bounding box (min x,y,z, max x,y,z)
  int pad1[2];
  unsigned long long t_ns[10];   // measurement: globaltimer at the start and after each phase of the K0 cluster kernel (thread 0)
};

// ---- the voxel grid of the window (host, once per pipeline) ------------------------------------------------
struct K0Grid {
  float inv_leaf;
  int min_b[3], dim[3], mul[3];
  int cells;
};

inline bool make_window_grid(const K0Window &w, float leaf, size_t cell_cap, K0Grid *g) {
  const float inv = 1.0f / leaf;                       // VoxelGrid::setLeafSize: inverse_leaf_size_ = 1 / leaf
  g->inv_leaf = inv;
  double cells = 1.0;
  for (int c = 0; c < 3; ++c) {
    if (!(w.lo[c] > -3.0e38f && w.hi[c] < 3.0e38f && w.lo[c] <= w.hi[c])) return false;   // finite window only
    const float lo = floorf(w.lo[c] * inv), hi = floorf(w.hi[c] * inv);
    if (!(lo > -1.0e9f && hi < 1.0e9f)) return false;
    g->min_b[c] = (int)lo;
    g->dim[c] = (int)hi - (int)lo + 1;
    cells *= (double)g->dim[c];
  }
  if (cells > (double)cell_cap || cells > 268435456.0) return false;
  g->mul[0] = 1; g->mul[1] = g->dim[0]; g->mul[2] = g->dim[0] * g->dim[1];
  g->cells = (int)cells;
  return true;
}

__device__ __forceinline__ int window_voxel_key(const float3 p, const K0Grid &g) {
  const int i0 = (int)(floorf(p.x * g.inv_leaf) - (float)g.min_b[0]);
  const int i1 = (int)(floorf(p.y * g.inv_leaf) - (float)g.min_b[1]);
  const int i2 = (int)(floorf(p.z * g.inv_leaf) - (float)g.min_b[2]);
  return i0 * g.mul[0] + i1 * g.mul[1] + i2 * g.mul[2];
}

constexpr int kK0Cluster = 8, kK0Threads = 1024, kK0Warps = kK0Threads / 32;
constexpr int kFrameMaxPoints = kResThreads * 12;   // the resident vote / fused loop hold <= 6144 points

struct FrameArgs {
  const float *raw;          // frame: n points, `stride` floats apart
  int n, stride;
  K0Window win;
  K0Grid grid;
  int *count;                // grid.cells ints: zero on entry, zero on exit
  int *start, *ncell;        // grid.cells ints each (written for occupied voxels only): list start, list length
  int *members;              // >= n raw indices, grouped by voxel (scatter order)
  int *sorted;               // >= n: the same lists in ascending raw index
  int2 *occ;                 // >= n: (start, count) of the occupied voxels in ascending voxel order
  int *tile_sums;            // 2 * kK0Cluster
  float *filt;               // 3 floats per centroid (the filtered cloud, kept for h3d_pipeline_filtered / post-tests)
  int out_cap;
  double *px, *py, *pz;      // centred SoA for the vote and the loop
  int *w;
  unsigned *orig;
  FrameDev *fd;
  // Hough ctor / loop configuration
  double opt_dx;
  unsigned long long num_b;
  LoopState *st;
  int minvotes, nlines_max, max_iters;
  unsigned pool_cap;
  unsigned long long *xchg;  // fused loop exchange words, zeroed here
  int xchg_words;
  unsigned *bar;             // 4 words, zeroed here
  // shared-memory budget of the vote / loop kernels (fast-path eligibility, decided on the device)
  unsigned smem_loop, tables_bytes, smem_vote;
};

// ---------------------------------------------------------------------------------------------------------
// K0 + prep + Hough ctor: one cluster.
// ---------------------------------------------------------------------------------------------------------
// Launched as ONE cluster of 1, 2, 4 or 8 CTAs (cudaLaunchAttributeClusterDimension): 8 for a frame that owns
// the GPU (lowest latency), 2 when frames are in flight beside each other -- a cluster needs all its SMs free in
// one GPC at the same time, and every CTA of the vote / loop kernels owns its SM's whole register file, so a
// wide cluster would wait for other frames' kernels to end while a narrow one costs 50 us of two SMs.
__global__ void __launch_bounds__(kK0Threads, 1)
frame_prefilter_kernel(const __grid_constant__ FrameArgs A) {
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  __shared__ int2 wsum[kK0Warps];
  __shared__ int2 wpre[kK0Warps];
  __shared__ int bb_s[6];
  constexpr int kBigList = 1024, kBigVoxel = 32;     // voxels with more members than kBigVoxel are summed by whole warps
  __shared__ int big_list[kBigList];
  __shared__ int big_n;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) big_n = 0;
  const int rank = (int)cluster.block_rank(), n_blk = (int)cluster.num_blocks();   // n_blk <= kK0Cluster
  const int gtid = rank * kK0Threads + tid, T = n_blk * kK0Threads;
  const K0Grid &G = A.grid;
  int stamp_k = 0;
  auto stamp = [&]() { if (gtid == 0 && stamp_k < 10) { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); A.fd->t_ns[stamp_k] = t; } ++stamp_k; };
  stamp();

  if (gtid < 6) A.fd->bb[gtid] = gtid < 3 ? INT_MAX : INT_MIN;
  if (tid < 6) bb_s[tid] = tid < 3 ? INT_MAX : INT_MIN;
  for (int k = gtid; k < A.xchg_words; k += T) A.xchg[k] = 0ull;
  if (gtid < 4) A.bar[gtid] = 0u;

  // Every phase is a handful of dependent L2 round trips per thread, so the loops are written in batches:
  // all loads of a batch are issued before the first result is used (a rolled loop costs one ~0.7 us round
  // trip per iteration: 65 us for the first version of this kernel).
  constexpr int U = 5;   // points per thread per batch: 5 * 8192 threads cover a 224x172 frame in one batch

  // ---- phase A: PassThrough window + voxel histogram ---------------------------------------------------
  for (int i0 = gtid; i0 < A.n; i0 += U * T) {
    float3 p[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int i = i0 + u * T;
      p[u] = i < A.n ? load_xyz(A.raw, i, A.stride) : make_float3(NAN, NAN, NAN);
    }
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (in_window(p[u], A.win)) atomicAdd(&A.count[window_voxel_key(p[u], G)], 1);
  }
  cluster.sync();
  stamp();

  // ---- phase B: exclusive scans over the grid (point offsets, output slots) ----------------------------
  // CTA `rank` owns cells [t0, t1); warp w of it the chunk [c0, c1); a warp reads 128 consecutive cells per
  // step, a lane four of them (one 16-byte load)
  const int tile = (((G.cells + n_blk - 1) / n_blk) + 4095) & ~4095;
  const int t0 = min(rank * tile, G.cells), t1 = min(t0 + tile, G.cells);
  const int chunk = tile / kK0Warps;                     // multiple of 128
  const int c0 = min(t0 + warp * chunk, t1), c1 = min(c0 + chunk, t1);
  auto load4 = [&](int c) -> int4 {                      // cells c..c+3 of this warp's chunk (0 beyond its end)
    if (c + 3 < c1) return *reinterpret_cast<const int4 *>(A.count + c);
    int4 v = make_int4(0, 0, 0, 0);
    if (c < c1) v.x = A.count[c];
    if (c + 1 < c1) v.y = A.count[c + 1];
    if (c + 2 < c1) v.z = A.count[c + 2];
    return v;
  };
  constexpr int SB = 4;                                  // steps per batch
  {
    int sc = 0, sn = 0;
    for (int cb = c0; cb < c1; cb += SB * 128) {
      int4 v[SB];
#pragma unroll
      for (int q = 0; q < SB; ++q) v[q] = load4(cb + q * 128 + 4 * lane);
#pragma unroll
      for (int q = 0; q < SB; ++q) {
        sc += v[q].x + v[q].y + v[q].z + v[q].w;
        sn += (v[q].x > 0) + (v[q].y > 0) + (v[q].z > 0) + (v[q].w > 0);
      }
    }
    sc = __reduce_add_sync(0xFFFFFFFFu, sc); sn = __reduce_add_sync(0xFFFFFFFFu, sn);
    if (lane == 0) wsum[warp] = make_int2(sc, sn);
  }
  __syncthreads();
  if (warp == 0) {
    const int2 v = wsum[lane];
    int2 inc = v;
    for (int o = 1; o < 32; o <<= 1) {
      const int x = __shfl_up_sync(0xFFFFFFFFu, inc.x, o), y = __shfl_up_sync(0xFFFFFFFFu, inc.y, o);
      if (lane >= o) { inc.x += x; inc.y += y; }
    }
    wpre[lane] = make_int2(inc.x - v.x, inc.y - v.y);
    if (lane == 31) { A.tile_sums[2 * rank] = inc.x; A.tile_sums[2 * rank + 1] = inc.y; }
  }
  cluster.sync();
  stamp();
  int n_window = 0, n_out = 0;
  {
    int bc = 0, bn = 0;
    for (int q = 0; q < n_blk; ++q) {
      const int sc = A.tile_sums[2 * q], sn = A.tile_sums[2 * q + 1];
      if (q < rank) { bc += sc; bn += sn; }
      n_window += sc; n_out += sn;
    }
    int run_c = bc + wpre[warp].x, run_n = bn + wpre[warp].y;
    for (int cb = c0; cb < c1; cb += SB * 128) {
      int4 v[SB];
#pragma unroll
      for (int q = 0; q < SB; ++q) v[q] = load4(cb + q * 128 + 4 * lane);
#pragma unroll
      for (int q = 0; q < SB; ++q) {
        const int s4 = v[q].x + v[q].y + v[q].z + v[q].w;
        if (__ballot_sync(0xFFFFFFFFu, s4 > 0) == 0) continue;               // most of the grid is empty
        const int n4 = (v[q].x > 0) + (v[q].y > 0) + (v[q].z > 0) + (v[q].w > 0);
        int ic = s4, in = n4;
        for (int o = 1; o < 32; o <<= 1) {
          const int x = __shfl_up_sync(0xFFFFFFFFu, ic, o), y = __shfl_up_sync(0xFFFFFFFFu, in, o);
          if (lane >= o) { ic += x; in += y; }
        }
        int sc = run_c + ic - s4, sn = run_n + in - n4;                      // exclusive prefixes of this lane's 4 cells
        const int c = cb + q * 128 + 4 * lane;
        const int vv[4] = {v[q].x, v[q].y, v[q].z, v[q].w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          if (vv[e] > 0) {
            A.start[c + e] = sc; A.ncell[c + e] = vv[e];
            A.occ[sn] = make_int2(sc, vv[e]);
            sc += vv[e]; ++sn;
          }
        }
        run_c += __shfl_sync(0xFFFFFFFFu, ic, 31);
        run_n += __shfl_sync(0xFFFFFFFFu, in, 31);
      }
    }
  }
  cluster.sync();
  stamp();

  // ---- phase C: scatter the raw indices into their voxel's list; the histogram goes back to zero -------
  // (a frame of up to U * T points is one batch: the keys / list starts / list lengths of a thread's points stay in
  // registers for phase C'; larger frames recompute them there)
  const bool one_batch = A.n <= U * T;
  int key[U], st[U], nc[U];
  for (int i0 = gtid; i0 < A.n; i0 += U * T) {
    int old[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int i = i0 + u * T;
      const float3 p = i < A.n ? load_xyz(A.raw, i, A.stride) : make_float3(NAN, NAN, NAN);
      key[u] = in_window(p, A.win) ? window_voxel_key(p, G) : -1;
    }
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (key[u] >= 0) { st[u] = A.start[key[u]]; nc[u] = A.ncell[key[u]]; old[u] = atomicSub(&A.count[key[u]], 1); }
#pragma unroll
    for (int u = 0; u < U; ++u) if (key[u] >= 0) A.members[st[u] + old[u] - 1] = i0 + u * T;
  }
  cluster.sync();
  stamp();

  // ---- phase C': the lists in input order (the canonical summation order of the centroids): the rank of a
  // point is the number of smaller raw indices in its voxel's list; one thread per POINT, a list is a few to a
  // few tens of entries long and its loads are independent ------------------------------------------------
  for (int i0 = gtid; i0 < A.n; i0 += U * T) {
    if (!one_batch) {
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int i = i0 + u * T;
        const float3 p = i < A.n ? load_xyz(A.raw, i, A.stride) : make_float3(NAN, NAN, NAN);
        key[u] = in_window(p, A.win) ? window_voxel_key(p, G) : -1;
        if (key[u] >= 0) { st[u] = A.start[key[u]]; nc[u] = A.ncell[key[u]]; }
      }
    }
    int rk[U], cmax = 0;
#pragma unroll
    for (int u = 0; u < U; ++u) { rk[u] = 0; if (key[u] < 0) nc[u] = 0; cmax = max(cmax, nc[u]); }
    for (int a0 = 0; a0 < cmax; a0 += 4) {
      int v[U][4];
#pragma unroll
      for (int u = 0; u < U; ++u)
#pragma unroll
        for (int q = 0; q < 4; ++q) v[u][q] = a0 + q < nc[u] ? A.members[st[u] + a0 + q] : INT_MAX;
#pragma unroll
      for (int u = 0; u < U; ++u)
#pragma unroll
        for (int q = 0; q < 4; ++q) rk[u] += v[u][q] < i0 + u * T;
    }
#pragma unroll
    for (int u = 0; u < U; ++u) if (key[u] >= 0) A.sorted[st[u] + rk[u]] = i0 + u * T;
  }
  cluster.sync();
  stamp();

  // ---- phase D: centroids: sequential float sums over the ordered lists; float bounding box -------------
  // one thread per occupied voxel (8 list entries and 8 points in flight per trip); a crowded voxel (a near
  // surface: a hundred pixels) is a warp's job: 32 loads per trip, the sum runs on lane 0's values by shuffle
  const int n_keep = min(n_out, A.out_cap);
  {
    int mn[3] = {INT_MAX, INT_MAX, INT_MAX}, mx[3] = {INT_MIN, INT_MIN, INT_MIN};
    auto emit = [&](int s, float cx, float cy, float cz, int cnt) {
      const float fc = (float)cnt;
      cx = __fdiv_rn(cx, fc); cy = __fdiv_rn(cy, fc); cz = __fdiv_rn(cz, fc);
      A.filt[3 * (size_t)s] = cx; A.filt[3 * (size_t)s + 1] = cy; A.filt[3 * (size_t)s + 2] = cz;
      const int o[3] = {float_to_ordered(cx), float_to_ordered(cy), float_to_ordered(cz)};
#pragma unroll
      for (int c = 0; c < 3; ++c) { mn[c] = min(mn[c], o[c]); mx[c] = max(mx[c], o[c]); }
    };
    for (int s = gtid; s < n_keep; s += T) {
      const int2 oc = A.occ[s];
      if (oc.y > kBigVoxel) {
        const int slot = atomicAdd(&big_n, 1);
        if (slot < kBigList) { big_list[slot] = s; continue; }
      }
      const int *m = A.sorted + oc.x;
      float cx = 0.0f, cy = 0.0f, cz = 0.0f;
#pragma unroll 8
      for (int a = 0; a < oc.y; ++a) {
        const float3 p = load_xyz(A.raw, m[a], A.stride);
        cx = __fadd_rn(cx, p.x); cy = __fadd_rn(cy, p.y); cz = __fadd_rn(cz, p.z);
      }
      emit(s, cx, cy, cz, oc.y);
    }
    __syncthreads();
    for (int e = warp; e < min(big_n, kBigList); e += kK0Warps) {
      const int s = big_list[e];
      const int2 oc = A.occ[s];
      const int *m = A.sorted + oc.x;
      float cx = 0.0f, cy = 0.0f, cz = 0.0f;
      int idx = lane < oc.y ? m[lane] : -1;
      for (int a0 = 0; a0 < oc.y; a0 += 32) {
        const int nxt = a0 + 32 + lane < oc.y ? m[a0 + 32 + lane] : -1;     // next trip's entries, in flight
        float3 p = make_float3(0.0f, 0.0f, 0.0f);
        if (idx >= 0) p = load_xyz(A.raw, idx, A.stride);
        const int cnt = min(32, oc.y - a0);
        for (int q = 0; q < cnt; ++q) {
          cx = __fadd_rn(cx, __shfl_sync(0xFFFFFFFFu, p.x, q));
          cy = __fadd_rn(cy, __shfl_sync(0xFFFFFFFFu, p.y, q));
          cz = __fadd_rn(cz, __shfl_sync(0xFFFFFFFFu, p.z, q));
        }
        idx = nxt;
      }
      if (lane == 0) emit(s, cx, cy, cz, oc.y);
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      mn[c] = __reduce_min_sync(0xFFFFFFFFu, mn[c]);
      mx[c] = __reduce_max_sync(0xFFFFFFFFu, mx[c]);
    }
    if (lane == 0) {
#pragma unroll
      for (int c = 0; c < 3; ++c) { atomicMin(&bb_s[c], mn[c]); atomicMax(&bb_s[3 + c], mx[c]); }
    }
    __syncthreads();
    if (tid < 3) atomicMin(&A.fd->bb[tid], bb_s[tid]);
    else if (tid < 6) atomicMax(&A.fd->bb[tid], bb_s[tid]);
  }
  cluster.sync();
  stamp();

  // ---- phase E: widen (wrap.h:175-189; the centroids are finite, nothing to strip), getMinMax3D,
  // shiftToOrigin (lib/pointcloud.cpp:18-26), the guards of wrap.h:202-212 and the Hough ctor ------------
  double lo[3], hi[3], ns[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    lo[c] = n_keep ? (double)ordered_to_float(A.fd->bb[c]) : 0.0;          // float -> double is exact and monotonic
    hi[c] = n_keep ? (double)ordered_to_float(A.fd->bb[3 + c]) : 0.0;
    ns[c] = (lo[c] + hi[c]) / 2.0;                                         // (p1 + p2) / 2.0
  }
  for (int s = gtid; s < n_keep; s += T) {
    A.px[s] = __dsub_rn((double)A.filt[3 * (size_t)s], ns[0]);
    A.py[s] = __dsub_rn((double)A.filt[3 * (size_t)s + 1], ns[1]);
    A.pz[s] = __dsub_rn((double)A.filt[3 * (size_t)s + 2], ns[2]);
    A.w[s] = 1;
    A.orig[s] = (unsigned)s;
  }
  if (gtid == 0) {
    FrameDev &F = *A.fd;
    F.n = n_keep; F.n_in_window = n_window; F.status = 0;
    const double ex = hi[0] - lo[0], ey = hi[1] - lo[1], ez = hi[2] - lo[2];
    const double d = sqrt((ex * ex) + (ey * ey) + (ez * ez));             // (maxP - minP).norm()
    F.d = d;
    double minPs[3], maxPs[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      F.ns[c] = ns[c]; F.mn[c] = lo[c]; F.mx[c] = hi[c];
      minPs[c] = lo[c] - ns[c];                                           // min(p - s) == min(p) - s: rounding is monotonic
      maxPs[c] = hi[c] - ns[c];
    }
    HoughGeom g = hough_geometry(minPs, maxPs, A.opt_dx, A.num_b);
    if (n_out > A.out_cap) F.status = 2;                                  // more voxels than the pipeline was sized for
    else if (d == 0.0 || A.opt_dx >= d) F.status = 1;                     // wrap.h:202-205, 209-212
    else if (g.rc != 0) F.status = 2;                                     // the general path reports the error
    F.g = g;
    F.cell_off = g.max_x / g.dx + 0.5;
    F.nb_planes = 0; F.n_cap = 0;
    if (F.status == 0) {
      // does the frame fit the register-resident vote and the shared-memory loop?  (same tests as the host's)
      const unsigned long long plane_bytes = (unsigned long long)g.plane_stride * 4ull;
      const unsigned long long n_cap = ((unsigned long long)n_keep + 1ull) & ~1ull;
      bool ok = g.bounded && n_keep >= 1 && n_keep <= kFrameMaxPoints && g.plane_stride < 65536u;
      ok = ok && (2ull * kResG * plane_bytes + 16ull <= A.smem_vote);
      ok = ok && (A.tables_bytes + 2ull * kResG * plane_bytes + 16ull <= A.smem_loop);
      ok = ok && (A.tables_bytes + 16ull + 32ull * n_cap + 2ull * plane_bytes <= A.smem_loop);
      if (!ok) F.status = 2;
      else {
        unsigned long long nb = (A.smem_loop - A.tables_bytes - 16ull - 32ull * n_cap) / plane_bytes;
        if (nb > 2ull * kResG) nb = 2ull * kResG;
        nb &= ~1ull;
        F.nb_planes = (int)nb; F.n_cap = (int)n_cap;
      }
    }
    LoopState st;
    st.opt_dx = A.opt_dx;
    for (int c = 0; c < 3; ++c) st.shift[c] = 0.0 + ns[c];                // X.shift = 0 + newshift
    st.minvotes = A.minvotes; st.nlines_max = A.nlines_max; st.max_iters = A.max_iters; st.pool_cap = A.pool_cap;
    st.done = F.status != 0; st.n_alive = n_keep; st.n_lines = 0; st.n_iters = 0; st.ny = 0; st.ny_prev = 0;
    st.pool_used = 0; st.work_count = 0; st.fresh_key = 0; st.best_key = 0;
    for (int c = 0; c < 3; ++c) { st.a[c] = 0.0; st.b[c] = 0.0; }
    st.votes_cast = 0;
    *A.st = st;
  }
  stamp();
}

// HoughDev of a frame on the fast path: the launch-time part (tables, buffers) + the geometry the frame's
// first kernel computed.  Called by one thread into SHARED memory; the others read it after a barrier.
__device__ __forceinline__ void patch_hough(HoughDev &h, const HoughDev &fixed, const FrameDev &F) {
  h = fixed;
  h.num_x = F.g.num_x; h.plane_cells = F.g.plane_cells; h.plane_stride = F.g.plane_stride;
  h.max_x = F.g.max_x; h.dx = F.g.dx; h.inv_s = F.g.inv_s; h.kmagic = F.g.kmagic;
  h.fast_ok = F.g.fast_ok; h.bounded = F.g.bounded;
  h.n = F.n; h.n_dev = nullptr;
}

// Hough::add of a frame on the fast path: the BOUNDED resident vote, geometry from the descriptor.
// The slot count P = ceil(n / 512) is picked ON THE DEVICE: the kernel holds the twelve instantiations of the
// vote loop and branches once (only the taken one occupies the instruction cache; the register allocation is
// that of the widest, and only its path spills).
// MODE 0: RED increments + plane scan; 1: TRACK (atomics return the old count, no scan); 2 (default): TRACK up
// to 6 slots (3072 points), scan above -- measured on the 16 bench frames (profiles/r02_track_ab.txt): TRACK
// wins 88 vs 103 us at 1.3 k points and loses 195 vs 179 us at 4.9 k (ATOMS with a return value costs ~13 cycles
// per warp-vote against ~9 for the merging RED form; the scan it saves is a fixed ~0.2 us per direction).
// Shared-memory layout: [HoughDev | ResShared | planes ...], carved out of the one extern array.
constexpr unsigned kVoteFrameHdr = 256;                                              // HoughDev
constexpr unsigned kVoteFrameRes = (unsigned)((sizeof(ResShared) + 255) / 256 * 256);
constexpr unsigned kVoteFramePlanesOff = kVoteFrameHdr + kVoteFrameRes;
static_assert(sizeof(HoughDev) <= kVoteFrameHdr, "HoughDev must fit its slot");

template <int MODE>
__global__ void __launch_bounds__(kResThreads, 1)
vote_frame_kernel(const __grid_constant__ HoughDev fixed, const FrameDev *__restrict__ fd) {
  extern __shared__ __align__(16) unsigned char vf_smem[];
  HoughDev &hs = *reinterpret_cast<HoughDev *>(vf_smem);
  unsigned *planes = reinterpret_cast<unsigned *>(vf_smem + kVoteFramePlanesOff);
  ResShared &sh = *reinterpret_cast<ResShared *>(vf_smem + kVoteFrameHdr);
  if (fd->status != 0) return;
  if (threadIdx.x == 0) patch_hough(hs, fixed, *fd);
  __syncthreads();
  uint4 *p4 = reinterpret_cast<uint4 *>(planes);
  for (unsigned q = threadIdx.x; q < 2 * kResG * hs.plane_stride / 4 + 1; q += kResThreads) p4[q] = make_uint4(0, 0, 0, 0);
  const int cta = blockIdx.x, n_cta = gridDim.x;
  // (One out-of-line function per slot count was tried for the sake of per-variant register allocation: ptxas then
  // rebuilt the shared-window address -- S2UR SR_CgaCtaId / ULEA / IMAD -- in front of every shared access of the
  // callee, +24 % instructions, 154 us against 132 us for the host-dispatched kernel; ncu, round 2.  Inlined, the
  // variants share the register allocation of the widest one and only its path spills.)
#define H3D_FRAME_VOTE(PP) vote_resident<PP, 0, true, (MODE == 1 || (MODE == 2 && PP <= 6))>(hs, nullptr, hs.dir_count, planes, sh, cta, n_cta); break;
  switch ((hs.n + kResThreads - 1) / kResThreads) {
    case 0: case 1: H3D_FRAME_VOTE(1)
    case 2: H3D_FRAME_VOTE(2)
    case 3: H3D_FRAME_VOTE(3)
    case 4: H3D_FRAME_VOTE(4)
    case 5: H3D_FRAME_VOTE(5)
    case 6: H3D_FRAME_VOTE(6)
    case 7: H3D_FRAME_VOTE(7)
    case 8: H3D_FRAME_VOTE(8)
    case 9: H3D_FRAME_VOTE(9)
    case 10: H3D_FRAME_VOTE(10)
    case 11: H3D_FRAME_VOTE(11)
    default: H3D_FRAME_VOTE(12)
  }
#undef H3D_FRAME_VOTE
}

// ---- results to the host in one piece: written straight into mapped pinned host memory ----------------
struct ExportArgs {
  const FrameDev *fd;
  const LoopState *st;
  const h3d_iter *iters;
  const unsigned *pool;
  const h3d_segment *segs;   // nullable
  int max_iters, seg_cap;
  unsigned pool_host_cap;    // entries the host staging buffer holds
  // host staging (device-visible addresses of pinned memory)
  FrameDev *fd_out;
  LoopState *st_out;
  h3d_iter *iters_out;
  unsigned *pool_out;
  h3d_segment *segs_out;
};

__global__ void __launch_bounds__(256) frame_export_kernel(const ExportArgs A) {
  const int gt = blockIdx.x * blockDim.x + threadIdx.x, T = gridDim.x * blockDim.x;
  const LoopState st = *A.st;
  if (gt == 0) { *A.fd_out = *A.fd; *A.st_out = st; }
  const int n_rec = min(st.n_iters, A.max_iters);
  {
    const uint2 *src = reinterpret_cast<const uint2 *>(A.iters);
    uint2 *dst = reinterpret_cast<uint2 *>(A.iters_out);
    const int words = n_rec * (int)(sizeof(h3d_iter) / sizeof(uint2));
    for (int k = gt; k < words; k += T) dst[k] = src[k];
  }
  const unsigned n_pool = min(st.pool_used, A.pool_host_cap);
  for (unsigned k = gt; k < n_pool; k += T) A.pool_out[k] = A.pool[k];
  if (A.segs) {
    const uint2 *src = reinterpret_cast<const uint2 *>(A.segs);
    uint2 *dst = reinterpret_cast<uint2 *>(A.segs_out);
    const int words = min(n_rec, A.seg_cap) * (int)(sizeof(h3d_segment) / sizeof(uint2));
    for (int k = gt; k < words; k += T) dst[k] = src[k];
  }
}

}  // namespace h3d
#endif  // H3D_FRAME_CUH_
