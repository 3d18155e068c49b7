// h3d_fused.cuh -- the whole frame loop as ONE persistent cooperative kernel (N <= 6144 points).
//   hough->add(X) + do { subtract, getLine, pointsCloseToLine, orthogonal_LSQ x2, removePoints }
//   wrap.h:228-342 / lib/hough3dlines.cpp:244-317
//
// One CTA per SM.  Division of labour:
//   * each CTA OWNS a fixed, strided set of directions (the ones it voted in the first pass); their
//     records (max votes, first max cell, stale flag) live in ITS shared memory (16 B each; the
//     coefficients are re-read from L2 when a record is tested or re-voted).  Un-voting removed points and re-voting stale planes are CTA-local.
//   * the line fit (inliers -> ordered compaction -> orthogonal_LSQ -> inliers -> LSQ -> remove) is
//     computed REDUNDANTLY and deterministically by every CTA from its own copy of the points, so no
//     CTA waits for another one's fit.
//   * two 8-byte exchanges per iteration, each an atomicMax + grid barrier: (A) the packed winner
//     key (Hough::getLine), (B) the best EXACT record after the un-vote, which is the bound that
//     decides which stale planes must be re-voted before the next getLine.
// Un-vote (Hough::subtract) never touches a plane: a record stays exact unless a removed point
// voted for its max cell.  The removed points Y lie within dx of the fitted line, so in the
// plane of a direction they project into a strip around a segment; a record whose max cell is
// farther than the strip's half-width from that segment cannot be hit (O(1) test per record),
// the few others are checked point by point with the exact cell arithmetic.
//
// Code size is a first-order concern: the SM's instruction cache holds 32 KB (2048 instructions)
// and one loop iteration walks through every phase once, so a loop body unrolled over register-
// resident points (an earlier version: 13 k instructions) is bound by instruction fetch from L2
// (ncu: icc hit rate 68 %, the first execution of a phase 4x slower than the second).  Hence:
// the FIRST vote keeps the points in registers (vote_resident, a tight hot loop), but the loop
// phases read a compact copy of the live points in SHARED memory with rolled loops, removed
// points are compacted away in place (order kept), Y is an index list, both fit passes share
// one copy of the code, and the eigen-solver's divisions and square roots are out of line.
#ifndef H3D_FUSED_CUH_
#define H3D_FUSED_CUH_

#include "h3d_frame.cuh"

namespace h3d {

struct FusedArgs {
  HoughDev h;                 // n = points after the NaN strip, n_dev = nullptr
  LoopState *st;              // configuration in, final state out
  h3d_iter *iters;
  unsigned *pool;
  unsigned *orig;             // original index of live point i (compacted in place by CTA 0)
  const double *dirs;
  unsigned long long *xchg;   // zeroed: [2*it] winner key, [2*it+1] best exact votes of iteration it
  unsigned *bar;              // zeroed grid-barrier counter
  int n_own_max;              // records per CTA (multiple of kResG)
  int nb_planes;              // planes available to a re-vote round (2..8)
  int n_cap;                  // capacity of the shared-memory point arrays (>= h.n, even)
  double near_t;              // largest nn with sqrt_rn(nn) <= opt_dx (pointsCloseToLine without the sqrt)
  double cell_off;            // max_x/dx + 0.5: cell coordinate of x' = 0
  unsigned long long *prof;   // measurement hook (nullable): [it][cta][kProfSlots] globaltimer stamps
  int prof_iters;
  const FrameDev *fd;         // frame fast path (nullable): geometry, point count and shared-memory layout come from
                              // the descriptor the frame's first kernel wrote; h.n / nb_planes / n_cap / cell_off above are unused
};
constexpr int kProfSlots = 16;
constexpr int kFusedMaxSlots = 12;

// A timer read placed right after __syncthreads() is NOT held back by the barrier (BAR.SYNC.DEFER_BLOCKING lets the
// warp run on until it touches memory): the volatile shared-memory load in front of it is, and instructions of a
// warp issue in order, so the stamp is taken once the barrier has completed.
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  unsigned probe;
  asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(probe) : "r"(0u) : "memory");
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t) : "r"(probe) : "memory");
  return t;
}

struct FusedShared {
  ResShared res;              // first vote (vote_resident) and re-vote rounds (warp_best)
  LsqOut lsq;
  LsqShared lsq_s;
  long long ipart[kResWarps][3];
  unsigned long long warp_key[kResWarps];
  int warp_cnt[kFusedMaxSlots][kResWarps];
  int red_lo[kResWarps], red_hi[kResWarps];
  int tmin_i, tmax_i;         // ordered-int encodings of float t bounds of Y
  int ns, nlist;
  unsigned rv_best[2 * kResG + 2][kResWarps];   // re-vote rounds: per-warp plane maxima
};

// live points (compact, cloud order) and the current inlier list, all in shared memory
struct FusedPoints {
  double *x, *y, *z;
  int *yidx;
  float *yt;     // t = b * (p - a) of the inliers of the current pass, in list order (un-vote filter)
};

__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned *p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ void grid_sync(unsigned *bar, unsigned &target, unsigned n_cta) {
  __syncthreads();
  if (threadIdx.x == 0) {
    target += n_cta;
    __threadfence();
    atomicAdd(bar, 1u);
    while (ld_acquire_u32(bar) < target) { }
  }
  __syncthreads();
}

__device__ __forceinline__ int float_ordered(float f) {
  const int i = __float_as_int(f);
  return i >= 0 ? i : i ^ 0x7FFFFFFF;
}
__device__ __forceinline__ float ordered_float(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7FFFFFFF); }

// pointsCloseToLine without the square root: nn <= near_t  <=>  sqrt_rn(nn) <= dx (sqrt_rn is
// monotonic; near_t is found on the host).  Same operation order as point_near_line.  *t_out: b*(p-a).
__device__ __forceinline__ bool point_near_line_t(double px, double py, double pz, const double (&a)[3],
                                                  const double (&b)[3], double near_t, double *t_out) {
  const double ex = __dsub_rn(px, a[0]), ey = __dsub_rn(py, a[1]), ez = __dsub_rn(pz, a[2]);
  const double t = __dadd_rn(__dadd_rn(__dmul_rn(b[0], ex), __dmul_rn(b[1], ey)), __dmul_rn(b[2], ez));
  const double qx = __dadd_rn(a[0], __dmul_rn(t, b[0]));
  const double qy = __dadd_rn(a[1], __dmul_rn(t, b[1]));
  const double qz = __dadd_rn(a[2], __dmul_rn(t, b[2]));
  const double dxv = __dsub_rn(px, qx), dyv = __dsub_rn(py, qy), dzv = __dsub_rn(pz, qz);
  const double nn = __dadd_rn(__dadd_rn(__dmul_rn(dxv, dxv), __dmul_rn(dyv, dyv)), __dmul_rn(dzv, dzv));
  *t_out = t;
  return nn <= near_t;
}

// frame-level fixed-point facts for the exact parallel anchor (see h3d_fit.cuh): every
// coordinate of the frame is a multiple of 2^lo and below 2^(hi+1); ok = both are sane
struct FixedPoint {
  int lo, hi, ok;
};
__device__ __forceinline__ void fixed_point_scan(double v, int &lo, int &hi) {
  const long long bits = __double_as_longlong(v);
  const int e = (int)((bits >> 52) & 0x7FF);
  const long long mant = bits & 0xFFFFFFFFFFFFFLL;
  if (e == 0x7FF || (e == 0 && mant != 0)) hi = 100000;           // not finite / subnormal: no fast path
  else if (e != 0) {
    lo = min(lo, e - 1075 + (__ffsll(mant | 0x10000000000000LL) - 1));
    hi = max(hi, e - 1023);
  }
}

// TREE32 chain of one warp over v(k), k = lane, lane+32, ... (4 loads in flight)
template <class F>
__device__ __forceinline__ float tree32_chain(F v, int n, int lane) {
  float acc = 0.0f;
  int k = lane;
#pragma unroll 1
  for (; k + 96 < n; k += 128) {
    const float v0 = v(k), v1 = v(k + 32), v2 = v(k + 64), v3 = v(k + 96);
    acc = __fadd_rn(__fadd_rn(__fadd_rn(__fadd_rn(acc, v0), v1), v2), v3);
  }
#pragma unroll 1
  for (; k < n; k += 32) acc = __fadd_rn(acc, v(k));
  return tree32_combine(acc);
}

// orthogonal_LSQ sums over Y = points[yidx[0..n)] (wrap.h:126-138) by the whole CTA: anchor,
// float column means, centred scatter in the canonical TREE32 order.  Leaves sh.lsq.a and
// sh.lsq_s.sc ready; the eigen-solve (lsq_eig) is a separate single-thread step so that the
// caller can overlap it.  Ends with a barrier.
__device__ __forceinline__ void lsq_sums_fused(const FusedPoints &Q, int n, const FixedPoint &fx, FusedShared &sh) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const bool fast = fx.ok && (n == 0 || fx.hi + 1 - fx.lo + (32 - __clz(n)) <= 53);
  if (warp < 3) {          // colwise().mean() of the float copy (wrap.h:129-137), TREE32
    const double *src = warp == 0 ? Q.x : warp == 1 ? Q.y : Q.z;
    const float acc = tree32_chain([&](int k) { return (float)src[Q.yidx[k]]; }, n, lane);
    if (lane == 0) sh.lsq_s.mean[warp] = __fdiv_rn(acc, (float)n);
  } else if (fast) {       // exact int64 sums, any order (provably the sequential double sum)
    const double scale = __longlong_as_double((long long)(1023 - fx.lo) << 52);   // 2^-lo
    long long acc[3] = {0, 0, 0};
#pragma unroll 1
    for (int k = tid - 96; k < n; k += kResThreads - 96) {
      const int i = Q.yidx[k];
      acc[0] += __double2ll_rn(__dmul_rn(Q.x[i], scale));
      acc[1] += __double2ll_rn(__dmul_rn(Q.y[i], scale));
      acc[2] += __double2ll_rn(__dmul_rn(Q.z[i], scale));
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
#pragma unroll
      for (int o = 16; o; o >>= 1) acc[c] += __shfl_xor_sync(0xFFFFFFFFu, acc[c], o);
      if (lane == 0) sh.ipart[warp][c] = acc[c];
    }
  } else if (warp < 6) {   // the reference's sequential double sum (rare)
    const double *src = warp == 3 ? Q.x : warp == 4 ? Q.y : Q.z;
    const double m = mean_axis_warp([&](int k) { return src[Q.yidx[k]]; }, n, lane);
    if (lane == 0) sh.lsq.a[warp - 3] = m;
  }
  __syncthreads();
  if (warp < 6) {          // centred^T * centred (wrap.h:137-138), TREE32 per entry
    const int c0 = warp < 3 ? 0 : warp < 5 ? 1 : 2, c1 = warp < 3 ? warp : warp < 5 ? warp - 2 : 2;
    const double *s0 = c0 == 0 ? Q.x : c0 == 1 ? Q.y : Q.z, *s1 = c1 == 0 ? Q.x : c1 == 1 ? Q.y : Q.z;
    const float m0 = sh.lsq_s.mean[c0], m1 = sh.lsq_s.mean[c1];
    const float acc = tree32_chain([&](int k) {
      const int i = Q.yidx[k];
      return __fmul_rn(__fsub_rn((float)s0[i], m0), __fsub_rn((float)s1[i], m1));
    }, n, lane);
    if (lane == 0) sh.lsq_s.sc[warp] = acc;
  } else if (warp == 6 && fast && lane < 3) {
    long long s = 0;
    for (int w = 3; w < kResWarps; ++w) s += sh.ipart[w][lane];
    const double sum = __dmul_rn((double)s, __longlong_as_double((long long)(1023 + fx.lo) << 52));
    sh.lsq.a[lane] = n > 0 ? __ddiv_rn(sum, (double)n) : sum;
  }
  __syncthreads();
}

// single thread: SelfAdjointEigenSolver on the scatter matrix (wrap.h:141-147)
__device__ __forceinline__ void lsq_eig(FusedShared &sh) {
  const float *sc = sh.lsq_s.sc;
  const float m[9] = {sc[0], sc[1], sc[2], sc[1], sc[3], sc[4], sc[2], sc[4], sc[5]};
  const Eig3f e = eig3f(m);
  sh.lsq.b[0] = e.evecs[6]; sh.lsq.b[1] = e.evecs[7]; sh.lsq.b[2] = e.evecs[8];   // last column
  sh.lsq.rc = (double)e.evals[2];
}

__device__ __forceinline__ void named_bar_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" :: "r"(id), "r"(threads) : "memory");
}

// X.pointsCloseToLine(a, b, dx, &Y): inlier mask of this thread's points (bit p = point
// tid + p*512), Y as an ordered index list in Q.yidx, |Y| in *ny_out (uniform).  CTA 0 also appends
// the original indices to the pool.  track_t: also reduce the range of t = b*(p - a) over Y into
// sh.tmin_i / sh.tmax_i.  Leaves the per-(slot, warp) inlier counts in sh.warp_cnt.  Ends with a
// barrier.
__device__ __forceinline__ unsigned inlier_pass(const FusedPoints &Q, int n_alive, const double (&a)[3],
                                                const double (&b)[3], double near_t, FusedShared &sh, bool track_t,
                                                bool to_pool, unsigned *pool, unsigned pool_base, unsigned pool_cap,
                                                const unsigned *orig, int *ny_out, unsigned long long *dbg) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int slots = (n_alive + kResThreads - 1) / kResThreads;
  unsigned m = 0;
  float tlo = 3.0e38f, thi = -3.0e38f;
  // two slots per trip: two independent ~40-deep FP64 dependency chains in flight per thread
#pragma unroll 1
  for (int p = 0; p < slots; p += 2) {
    const int i0 = tid + p * kResThreads, i1 = i0 + kResThreads;
    const bool v0 = i0 < n_alive, v1 = i1 < n_alive;
    double t0, t1;
    bool in0 = point_near_line_t(v0 ? Q.x[i0] : 0.0, v0 ? Q.y[i0] : 0.0, v0 ? Q.z[i0] : 0.0, a, b, near_t, &t0);
    bool in1 = point_near_line_t(v1 ? Q.x[i1] : 0.0, v1 ? Q.y[i1] : 0.0, v1 ? Q.z[i1] : 0.0, a, b, near_t, &t1);
    in0 = in0 && v0; in1 = in1 && v1;
    if (in0) { const float tf = (float)t0; tlo = fminf(tlo, tf); thi = fmaxf(thi, tf); }
    if (in1) { const float tf = (float)t1; tlo = fminf(tlo, tf); thi = fmaxf(thi, tf); }
    m |= (in0 ? (1u << p) : 0u) | (in1 ? (2u << p) : 0u);
    const unsigned bal0 = __ballot_sync(0xFFFFFFFFu, in0), bal1 = __ballot_sync(0xFFFFFFFFu, in1);
    if (lane == 0) { sh.warp_cnt[p][warp] = __popc(bal0); if (p + 1 < slots) sh.warp_cnt[p + 1][warp] = __popc(bal1); }
  }
  if (dbg && tid == 0) *dbg = globaltimer_ns();
  if (track_t) {
    const int lo_i = __reduce_min_sync(0xFFFFFFFFu, float_ordered(tlo));
    const int hi_i = __reduce_max_sync(0xFFFFFFFFu, float_ordered(thi));
    if (lane == 0) { atomicMin(&sh.tmin_i, lo_i); atomicMax(&sh.tmax_i, hi_i); }
  }
  __syncthreads();
  const unsigned below = (1u << lane) - 1u;
  int run = 0;
#pragma unroll 1
  for (int p = 0; p < slots; ++p) {
    const int c = lane < kResWarps ? sh.warp_cnt[p][lane] : 0;
    const int tot = __reduce_add_sync(0xFFFFFFFFu, c);
    const int pre = __reduce_add_sync(0xFFFFFFFFu, lane < warp ? c : 0);
    const bool in = (m >> p) & 1u;
    const unsigned bal = __ballot_sync(0xFFFFFFFFu, in);
    if (in) {
      const int pos = run + pre + __popc(bal & below);
      const int i = tid + p * kResThreads;
      Q.yidx[pos] = i;
      if (track_t) {   // t = b * (p - a) again (same operations as the test above), kept for the un-vote's filter
        const double ex = __dsub_rn(Q.x[i], a[0]), ey = __dsub_rn(Q.y[i], a[1]), ez = __dsub_rn(Q.z[i], a[2]);
        Q.yt[pos] = (float)__dadd_rn(__dadd_rn(__dmul_rn(b[0], ex), __dmul_rn(b[1], ey)), __dmul_rn(b[2], ez));
      }
      if (to_pool && pool_base + (unsigned)pos < pool_cap) pool[pool_base + pos] = orig[i];
    }
    run += tot;
  }
  __syncthreads();
  *ny_out = run;
  return m;
}

// Re-vote of two directions (j1 < 0: only one) over the compact live points, BOUNDED vote form
// (the host launches this kernel only when HoughDev::bounded holds).  The planes of late iterations
// are sparse, so instead of scanning them the increments RETURN the old count: the last voter of a
// cell sees its final count - 1, hence max over all votes of (old + 1, lowest cell) is the plane's
// (max votes, first max cell).  key = (old + 1) << 16 | (0xFFFF - cell); the caller zero-fills the
// planes afterwards.  Votes the reference drops (slack cells, exact path) do not enter the key.
__device__ __forceinline__ void revote_pair(const HoughDev &h, const FusedPoints &Q, int n_alive, const FastCoef &f0,
                                            const FastCoef &f1, int j0, int j1, unsigned plane0_sa, unsigned plane1_sa,
                                            unsigned dummy_sa, unsigned &best0, unsigned &best1) {
  const double kmagic = h.kmagic;
  const unsigned row4 = 4u * (unsigned)h.num_x, plane_cells = h.plane_cells, lim = 4u * h.plane_cells;
  const bool two = j1 >= 0, no_fast = !h.fast_ok;
  unsigned b0 = 0, b1 = 0;
#pragma unroll 2
  for (int i = threadIdx.x; i < n_alive; i += kResThreads) {
    const double x = Q.x[i], y = Q.y[i], z = Q.z[i];
    unsigned gmin = 0xFFFFFFFFu;
    unsigned o0 = vote_off_bounded(f0, kmagic, row4, x, y, z, gmin);
    unsigned o1 = vote_off_bounded(f1, kmagic, row4, x, y, z, gmin);
    if ((gmin < 0x10000u) | no_fast) {   // rare: within 8e-6 cell of a boundary
      const unsigned c0 = vote_cell_checked(h.coef + j0, h.inv_s, kmagic, h.max_x, h.dx, h.num_x, h.fast_ok, x, y, z);
      o0 = c0 < plane_cells ? 4u * c0 : lim;
      if (two) {
        const unsigned c1 = vote_cell_checked(h.coef + j1, h.inv_s, kmagic, h.max_x, h.dx, h.num_x, h.fast_ok, x, y, z);
        o1 = c1 < plane_cells ? 4u * c1 : lim;
      }
    }
    const bool in0 = o0 < lim, in1 = two & (o1 < lim);
    // BOUNDED form: a vote that is not counted still lands inside its plane's stride (slack cells; `lim` on the exact
    // path), so plane 0 needs no redirect; plane 1 does not exist when the pair is a single record
    const unsigned old0 = atom_shared_inc_ret(plane0_sa + o0);
    const unsigned old1 = atom_shared_inc_ret(two ? plane1_sa + o1 : dummy_sa);
    const unsigned k0 = ((old0 + 1u) << 16) | ((o0 >> 2) ^ 0xFFFFu), k1 = ((old1 + 1u) << 16) | ((o1 >> 2) ^ 0xFFFFu);
    b0 = max(b0, in0 ? k0 : 0u);
    b1 = max(b1, in1 ? k1 : 0u);
  }
  best0 = b0; best1 = b1;
}

// The same for FOUR directions per pass over the points (records j[k] < 0: none): the three coordinate loads, the
// loop overhead and the fixed cost of a pass (coefficient fetch from L2, key reductions) are shared by four
// planes instead of two.  Used when the frame's layout has at least four planes.
__device__ __forceinline__ void revote_quad(const HoughDev &h, const FusedPoints &Q, int n_alive, const FastCoef (&f)[4],
                                            const int (&j)[4], unsigned plane0_sa, unsigned plane_bytes, unsigned dummy_sa,
                                            unsigned (&best)[4]) {
  const double kmagic = h.kmagic;
  const unsigned row4 = 4u * (unsigned)h.num_x, plane_cells = h.plane_cells, lim = 4u * h.plane_cells;
  const bool no_fast = !h.fast_ok;
  unsigned b[4] = {0u, 0u, 0u, 0u};
#pragma unroll 1
  for (int i = threadIdx.x; i < n_alive; i += kResThreads) {
    const double x = Q.x[i], y = Q.y[i], z = Q.z[i];
    unsigned gmin = 0xFFFFFFFFu;
    unsigned o[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) o[k] = vote_off_bounded(f[k], kmagic, row4, x, y, z, gmin);
    if ((gmin < 0x10000u) | no_fast) {   // rare: within 8e-6 cell of a boundary
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (j[k] < 0) continue;
        const unsigned c = vote_cell_checked(h.coef + j[k], h.inv_s, kmagic, h.max_x, h.dx, h.num_x, h.fast_ok, x, y, z);
        o[k] = c < plane_cells ? 4u * c : lim;
      }
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const bool have = j[k] >= 0, in = have & (o[k] < lim);
      const unsigned old = atom_shared_inc_ret(have ? plane0_sa + (unsigned)k * plane_bytes + o[k] : dummy_sa);
      const unsigned key = ((old + 1u) << 16) | ((o[k] >> 2) ^ 0xFFFFu);
      b[k] = max(b[k], in ? key : 0u);
    }
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) best[k] = b[k];
}

// ---- shared-memory layout of the loop kernel ---------------------------------------------------------------
// [HoughDev | FusedShared | record tables | planes (+ pad) | live points, inlier list, its t values], all carved
// out of the one extern array so that the out-of-line phase functions below address it as SHARED memory (LDS /
// STS / ATOMS) and get a register allocation of their own: inlined into one body, every hot loop of the kernel
// shared 128 registers with everything else that was live and spilled (224 bytes with the re-vote prefetch).
constexpr unsigned kFusedHoughBytes = 256;
constexpr unsigned kFusedHdr = kFusedHoughBytes + (unsigned)((sizeof(FusedShared) + 255) / 256 * 256);
static_assert(sizeof(HoughDev) <= kFusedHoughBytes, "HoughDev must fit its slot");

struct FusedView {
  const HoughDev *h;
  FusedShared *sh;
  unsigned *rv, *rcell, *rstale;   // per own record: max votes (bound when stale), first max cell, stale flag
  int *rlist;
  unsigned *planes;
  unsigned planes_sa, dummy_sa, stride, quads;
  FusedPoints Q;
};
__device__ __forceinline__ FusedView fused_view(int n_own, int nb, int n_cap) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  FusedView V;
  V.h = reinterpret_cast<const HoughDev *>(smem_raw);
  V.sh = reinterpret_cast<FusedShared *>(smem_raw + kFusedHoughBytes);
  unsigned char *body = smem_raw + kFusedHdr;
  V.rv = reinterpret_cast<unsigned *>(body);
  V.rcell = V.rv + n_own;
  V.rstale = V.rcell + n_own;
  V.rlist = reinterpret_cast<int *>(V.rstale + n_own);
  V.planes = reinterpret_cast<unsigned *>(body + (size_t)n_own * 16u);   // n_own % 4 == 0 -> multiple of 16
  V.stride = V.h->plane_stride; V.quads = V.stride / 4;
  V.planes_sa = (unsigned)__cvta_generic_to_shared(V.planes);
  V.dummy_sa = V.planes_sa + (unsigned)nb * V.stride * 4u;              // pad word of the loop layout
  V.Q.x = reinterpret_cast<double *>(V.planes + (size_t)nb * V.stride + 4);
  V.Q.y = V.Q.x + n_cap; V.Q.z = V.Q.y + n_cap;
  V.Q.yidx = reinterpret_cast<int *>(V.Q.z + n_cap);
  V.Q.yt = reinterpret_cast<float *>(V.Q.yidx + n_cap);
  return V;
}

// Which direction (index inside the instance's slice) is record r of CTA `cta`.  With few CTAs (frames in flight, 9-18
// SMs each) groups of G consecutive directions stay together (their coefficients share cache lines).  With the whole
// GPU on one frame the map is a pure interleave: directions next to each other on the sphere turn stale together, and
// with 140 records per CTA a clump of them on one CTA made every other CTA wait at the next barrier for its re-votes.
__device__ __forceinline__ int own_dir(int r, int cta, int n_cta) {
  constexpr int G = kResG;
  return n_cta >= 64 ? r * n_cta + cta : (cta + (r / G) * n_cta) * G + (r % G);
}

// K4 inside the loop: Hough::subtract(Y) for the CTA's own directions, run by warps 1..15 while warp 0 solves the
// second eigen-problem.  Y (pass 2, ny2 points) lies within opt_dx of the line (a, b) of pass 2.
//
// A record turns stale iff some point e of Y votes for its max cell.  Only points whose parameter t_e along the
// line falls into the record's strip interval can (strip_interval), so Y is first bucketed by t (counting sort
// into kUvBuckets buckets; the scratch is the head of the plane area, which is all zero between re-vote rounds
// and is zeroed again here) and every record then looks at the few points of its bucket range only:
//   stage 1  one THREAD per record: interval, bucket range; up to kUvInline candidates are tested on the spot
//            (exact cell arithmetic, first hit ends the record), longer ranges go to a list;
//   stage 2  one WARP per listed record, 32 candidates per step.
// Before (every surviving record took a warp that walked all of Y with a float-t filter) the phase cost 20-90 us
// per iteration with 2276 records per CTA (9 SMs per frame), most of it per-record overhead of stage 2.
constexpr int kUvBuckets = 256, kUvInline = 8;

__device__ __noinline__ void fused_unvote(const double (&a)[3], const double (&b)[3], int ny2, int n_own, int nb, int n_cap,
                                          int cta, int n_cta, double cell_off, double reach2) {
  constexpr int G = kResG;
  const FusedView V = fused_view(n_own, nb, n_cap);
  const HoughDev &h = *V.h;
  FusedShared &sh = *V.sh;
  unsigned *rv = V.rv, *rcell = V.rcell, *rstale = V.rstale;
  int *rlist = V.rlist;
  const FusedPoints &Q = V.Q;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int ut = tid - 32, un = kResThreads - 32;
  const float tl = ordered_float(sh.tmin_i), th = ordered_float(sh.tmax_i);
  const double t0 = (double)tl - (fabs((double)tl) * 1e-6 + 1e-9), t1 = (double)th + (fabs((double)th) * 1e-6 + 1e-9);
  const unsigned num_x = (unsigned)h.num_x;

  // ---- Y bucketed by t: start[0..kUvBuckets] | fill[kUvBuckets] | sidx[ny2] | st[ny2] in the (zero) plane area ----
  int *start = reinterpret_cast<int *>(V.planes);
  int *fill = start + kUvBuckets + 1;
  int *sidx = fill + kUvBuckets;
  float *st = reinterpret_cast<float *>(sidx + ny2);
  const unsigned scratch_words = 2u * kUvBuckets + 1u + 2u * (unsigned)ny2;
  // uniform.  Not with few records per CTA AND a short Y (a frame on all SMs: the three barriers of the bucketing
  // would cost more than they save, and the phase hides behind the eigen-solve anyway; with a long Y the warp stage
  // walked 400 inliers for each of ~140 records, 10 us on the critical path) or tiny planes with a huge Y.
  const bool bucketed = (n_own >= 512 || ny2 >= 128) && scratch_words <= (unsigned)nb * V.stride;
  const float bscale = th > tl ? (float)kUvBuckets / (th - tl) : 0.0f;
  // monotonic in t (float subtraction, multiplication by a positive constant, truncation, clamp): a point with
  // flo <= t_e <= fhi lies in a bucket of [bucket_of(flo), bucket_of(fhi)]
  auto bucket_of = [&](float t) {
    const float v = (t - tl) * bscale;
    return v >= (float)(kUvBuckets - 1) ? kUvBuckets - 1 : (v > 0.0f ? (int)v : 0);
  };
  if (bucketed) {
    for (int e = ut; e < ny2; e += un) atomicAdd(&start[bucket_of(Q.yt[e]) + 1], 1);
    named_bar_sync(1, un);
    if (warp == 1) {   // inclusive scan of the counts: start[k] = points in buckets below k
      constexpr int PER = kUvBuckets / 32;
      int c[PER], tot = 0;
#pragma unroll
      for (int q = 0; q < PER; ++q) { c[q] = start[1 + lane * PER + q]; tot += c[q]; }
      int incl = tot;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xFFFFFFFFu, incl, o); if (lane >= o) incl += v; }
      int run = incl - tot;
#pragma unroll
      for (int q = 0; q < PER; ++q) { run += c[q]; start[1 + lane * PER + q] = run; }
    }
    named_bar_sync(1, un);
    for (int e = ut; e < ny2; e += un) {
      const float te = Q.yt[e];
      const int bk = bucket_of(te);
      const int pos = start[bk] + atomicAdd(&fill[bk], 1);
      sidx[pos] = Q.yidx[e]; st[pos] = te;
    }
    named_bar_sync(1, un);
  }

  // ---- stage 1: one thread per record ----
  constexpr int SU = 3;   // records per thread per batch: their coefficient loads (L2) are issued together
  for (int r0 = ut; r0 < n_own; r0 += SU * un) {
    DirCoef cj[SU];
    bool live[SU];
#pragma unroll
    for (int u = 0; u < SU; ++u) {
      const int r = r0 + u * un;
      live[u] = r < n_own && !rstale[r] && rv[r] != 0;
      if (live[u]) cj[u] = h.coef[h.dir_begin + own_dir(r, cta, n_cta)];
    }
#pragma unroll
    for (int u = 0; u < SU; ++u) {
      if (!live[u]) continue;
      const int r = r0 + u * un;
      const unsigned target = rcell[r];
      const unsigned cx = target / num_x, cy = target - cx * num_x;
      // y_i == num_x or -1 aliases into a neighbouring row (flat bound only, lib/hough.cpp:90-92): edge cells take
      // every point of Y
      const bool edge = cy == 0 || cy == num_x - 1;
      const FastCoef fc = make_fast(cj[u], h.inv_s);
      double ilo = -INFINITY, ihi = INFINITY;
      if (!edge && !(strip_interval(fc, a, b, cell_off, reach2, cx, cy, &ilo, &ihi) && ilo <= t1 && ihi >= t0)) continue;
      if (!bucketed) { rlist[atomicAdd(&sh.ns, 1)] = r; continue; }
      const float flo = (float)(ilo - (fabs(ilo) * 1e-6 + 1e-6)), fhi = (float)(ihi + (fabs(ihi) * 1e-6 + 1e-6));
      const int e0 = start[bucket_of(flo)], e1 = start[bucket_of(fhi) + 1];
      const int j = h.dir_begin + own_dir(r, cta, n_cta);
      // middle-out over the sorted range: a vote for the cell needs P(t) within half a cell of its centre, i.e.
      // t in the central third of the interval, so a record whose max cell lies ON the removed line (almost
      // every record when a dominant line goes) is hit within the first few tests
      const int len = e1 - e0, tries = min(len, kUvInline);
      int hi_i = e0 + (len >> 1);
      if (ihi - ilo < 1e30) {   // start at the bucket of the interval's centre (the range may be clipped at an end of Y)
        const int bc = bucket_of((float)(0.5 * (ilo + ihi)));
        hi_i = min(max((start[bc] + start[bc + 1]) >> 1, e0), max(e1 - 1, e0));
      }
      int lo_i = hi_i - 1;
      bool hit = false;
#pragma unroll 1
      for (int q = 0; q < tries; ++q) {   // hi, lo, hi, lo, ... while both sides last, then the side that is left
        const bool up = (q & 1) == 0 ? hi_i < e1 : lo_i < e0;
        const int e = up ? hi_i++ : lo_i--;
        const float te = st[e];
        if (!(te >= flo && te <= fhi)) continue;
        const int i = sidx[e];
        if (vote_cell(h, fc, j, Q.x[i], Q.y[i], Q.z[i]) == target) { hit = true; break; }
      }
      if (hit) rstale[r] = 1;
      else if (len > kUvInline) rlist[atomicAdd(&sh.ns, 1)] = r;   // the rest of a long range: a warp takes it
    }
  }
  named_bar_sync(1, un);

  // ---- stage 2: one WARP per listed record, 32 candidates per step, done at the first hit ----
  const int ns = sh.ns;
  // the coefficients of a record come from L2 (~0.7 us): lane q of the warp fetches those of the q-th record
  // the warp will test, 32 records per round trip, and hands them over by shuffle
  DirCoef pre = {0.0, 0.0, 0.0, 0.0, 0.0};
  int pre_base = -1;
#pragma unroll 1
  for (int s = warp - 1, q = 0; s < ns; s += kResWarps - 1, ++q) {
    if ((q & 31) == 0) {
      const int sq = s + lane * (kResWarps - 1);
      if (sq < ns) {
        const int rq = rlist[sq];
        pre = h.coef[h.dir_begin + own_dir(rq, cta, n_cta)];
      }
      pre_base = q;
    }
    const int r = rlist[s];
    const int j = h.dir_begin + own_dir(r, cta, n_cta);
    DirCoef cj;
    {
      const int src = q - pre_base;
      cj.c0 = __shfl_sync(0xFFFFFFFFu, pre.c0, src); cj.c1 = __shfl_sync(0xFFFFFFFFu, pre.c1, src);
      cj.c2 = __shfl_sync(0xFFFFFFFFu, pre.c2, src); cj.bx = __shfl_sync(0xFFFFFFFFu, pre.bx, src);
      cj.by = __shfl_sync(0xFFFFFFFFu, pre.by, src);
    }
    const FastCoef fc = make_fast(cj, h.inv_s);
    const unsigned target = rcell[r];
    const unsigned cx = target / num_x, cy = target - cx * num_x;
    double ilo = -INFINITY, ihi = INFINITY;
    if (cy != 0 && cy != num_x - 1) strip_interval(fc, a, b, cell_off, reach2, cx, cy, &ilo, &ihi);
    const float flo = (float)(ilo - (fabs(ilo) * 1e-6 + 1e-6)), fhi = (float)(ihi + (fabs(ihi) * 1e-6 + 1e-6));
    const int e_begin = bucketed ? start[bucket_of(flo)] : 0, e_end = bucketed ? start[bucket_of(fhi) + 1] : ny2;
#pragma unroll 1
    for (int e0 = e_begin; e0 < e_end; e0 += 32) {
      const int e = e0 + lane;
      bool cand = false;
      if (e < e_end) { const float te = bucketed ? st[e] : Q.yt[e]; cand = te >= flo && te <= fhi; }
      if (!__any_sync(0xFFFFFFFFu, cand)) continue;
      bool hit = false;
      if (cand) { const int i = bucketed ? sidx[e] : Q.yidx[e]; hit = vote_cell(h, fc, j, Q.x[i], Q.y[i], Q.z[i]) == target; }
      if (__any_sync(0xFFFFFFFFu, hit)) { if (lane == 0) rstale[r] = 1; break; }
    }
  }
  if (bucketed) {   // the scratch back to zero: the re-vote rounds count on an all-zero plane area
    named_bar_sync(1, un);
    for (unsigned w = (unsigned)ut; w < scratch_words; w += (unsigned)un) V.planes[w] = 0u;
  }
}

// Re-vote of the stale records listed in rlist[0..nlist) (their bound reaches the best exact record).
__device__ __noinline__ void fused_revote(int nlist, int n_alive, int n_own, int nb, int n_cap, int cta, int n_cta) {
  constexpr int G = kResG;
  const FusedView V = fused_view(n_own, nb, n_cap);
  const HoughDev &h = *V.h;
  FusedShared &sh = *V.sh;
  unsigned *rv = V.rv, *rcell = V.rcell, *rstale = V.rstale;
  int *rlist = V.rlist;
  unsigned *planes = V.planes;
  const unsigned planes_sa = V.planes_sa, dummy_sa = V.dummy_sa, stride = V.stride, quads = V.quads;
  const FusedPoints &Q = V.Q;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // (Fetching the coefficients a round ahead through shared memory and zero-filling only the reachable rows was
    // tried and measured SLOWER -- 425 vs 365 us for a 786-plane rebuild at 18 SMs: the 16 warps already overlap
    // the L2 latency of the two coefficient loads per pair, and the per-round bookkeeping cost more than it hid.)
#pragma unroll 1
    for (int b0 = 0; b0 < nlist; b0 += nb) {
      const int cnt = min(nb, nlist - b0);
      if (nb >= 4) {
#pragma unroll 1
        for (int bi = 0; bi < cnt; bi += 4) {
          int j[4];
          FastCoef f[4];
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const int r = bi + k < cnt ? rlist[b0 + bi + k] : -1;
            j[k] = r >= 0 ? h.dir_begin + own_dir(r, cta, n_cta) : -1;
          }
#pragma unroll
          for (int k = 0; k < 4; ++k) f[k] = make_fast(h.coef[j[k] >= 0 ? j[k] : j[0]], h.inv_s);
          unsigned key[4];
          revote_quad(h, Q, n_alive, f, j, planes_sa + (unsigned)bi * stride * 4u, stride * 4u, dummy_sa, key);
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const unsigned kk = __reduce_max_sync(0xFFFFFFFFu, key[k]);
            if (lane == 0) sh.rv_best[bi + k][warp] = kk;
          }
        }
      } else {
#pragma unroll 1
      for (int bi = 0; bi < cnt; bi += 2) {
        const int r0 = rlist[b0 + bi], r1 = bi + 1 < cnt ? rlist[b0 + bi + 1] : -1;
        const int j0 = h.dir_begin + own_dir(r0, cta, n_cta);
        const int j1 = r1 >= 0 ? h.dir_begin + own_dir(r1, cta, n_cta) : -1;
        unsigned k0, k1;
        revote_pair(h, Q, n_alive, make_fast(h.coef[j0], h.inv_s), make_fast(h.coef[j1 >= 0 ? j1 : j0], h.inv_s), j0, j1,
                    planes_sa + (unsigned)bi * stride * 4u, planes_sa + (unsigned)(bi + 1) * stride * 4u, dummy_sa, k0, k1);
        k0 = __reduce_max_sync(0xFFFFFFFFu, k0);
        k1 = __reduce_max_sync(0xFFFFFFFFu, k1);
        if (lane == 0) { sh.rv_best[bi][warp] = k0; sh.rv_best[bi + 1][warp] = k1; }
      }
      }
      __syncthreads();
      {   // the planes of this round back to zero (no scan: the keys already hold the maxima)
        uint4 *p4 = reinterpret_cast<uint4 *>(planes);
        for (unsigned q = tid; q < (unsigned)cnt * quads; q += kResThreads) p4[q] = make_uint4(0, 0, 0, 0);
      }
      if (warp < cnt) {
        unsigned k = lane < kResWarps ? sh.rv_best[warp][lane] : 0u;
        k = __reduce_max_sync(0xFFFFFFFFu, k);
        if (lane == 0) {
          const int r = rlist[b0 + warp];
          const unsigned v = k >> 16;
          rv[r] = v; rcell[r] = v ? 0xFFFFu - (k & 0xFFFFu) : 0u; rstale[r] = 0;
        }
      }
      __syncthreads();
    }
}

__global__ void __launch_bounds__(kResThreads, 1) fused_loop_kernel(const FusedArgs A) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  HoughDev &hs = *reinterpret_cast<HoughDev *>(smem_raw);
  constexpr int G = kResG;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int cta = blockIdx.x, n_cta = gridDim.x;
  if (A.fd && A.fd->status != 0) return;   // the frame ended before the vote (wrap.h:202-212) or takes the general path
  if (tid == 0) { if (A.fd) patch_hough(hs, A.h, *A.fd); else hs = A.h; }
  __syncthreads();
  const int n_cap = A.fd ? A.fd->n_cap : A.n_cap;
  const double cell_off = A.fd ? A.fd->cell_off : A.cell_off;
  // per-CTA record tables first, then the planes (+ pad word); behind the planes of a re-vote
  // round (the first vote uses 2*G planes over the same area) the live points and the inlier list
  const int n_own = A.n_own_max;
  const int nb = A.fd ? A.fd->nb_planes : A.nb_planes;
  const FusedView V = fused_view(n_own, nb, n_cap);
  const HoughDev &h = *V.h;
  FusedShared &sh = *V.sh;
  unsigned *rv = V.rv, *rcell = V.rcell, *rstale = V.rstale;
  int *rlist = V.rlist;
  unsigned *planes = V.planes;
  const unsigned quads = V.quads;
  const FusedPoints Q = V.Q;

  if (A.prof && tid == 0 && A.prof_iters > 0) A.prof[(size_t)cta * kProfSlots + 15] = globaltimer_ns();   // kernel start
  // hough->add(X) ran just before this kernel (vote_resident_kernel, BOUNDED form): the records of
  // every direction are in global memory.  Planes of the re-vote rounds start at zero.
  {
    uint4 *p4 = reinterpret_cast<uint4 *>(planes);
    for (unsigned q = tid; q < (unsigned)nb * quads + 1; q += kResThreads) p4[q] = make_uint4(0, 0, 0, 0);
  }
  // own records: entry r = k*G + g is direction (cta + k*n_cta)*G + g
  for (int r = tid; r < n_own; r += kResThreads) {
    const int l = own_dir(r, cta, n_cta);
    if (l < h.dir_count) {
      rv[r] = h.dvotes[l]; rcell[r] = h.dcell[l];
    } else { rv[r] = 0; rcell[r] = 0; }
    rstale[r] = 0;
  }
  // the points, for the loop, into shared memory; and the frame's fixed-point facts
  FixedPoint fx;
  {
    int lo = 0x7FFFFFFF, hi = -0x7FFFFFFF;
#pragma unroll 1
    for (int i = tid; i < h.n; i += kResThreads) {
      const double x = h.px[i], y = h.py[i], z = h.pz[i];
      Q.x[i] = x; Q.y[i] = y; Q.z[i] = z;
      fixed_point_scan(x, lo, hi); fixed_point_scan(y, lo, hi); fixed_point_scan(z, lo, hi);
    }
    lo = __reduce_min_sync(0xFFFFFFFFu, lo);
    hi = __reduce_max_sync(0xFFFFFFFFu, hi);
    if (lane == 0) { sh.red_lo[warp] = lo; sh.red_hi[warp] = hi; }
    __syncthreads();
    lo = sh.red_lo[0]; hi = sh.red_hi[0];
    for (int w = 1; w < kResWarps; ++w) { lo = min(lo, sh.red_lo[w]); hi = max(hi, sh.red_hi[w]); }
    fx.lo = lo; fx.hi = hi;
    fx.ok = (lo != 0x7FFFFFFF && lo > -900 && hi < 900) ? 1 : 0;
  }
  const LoopState cfg = *A.st;
  int n_alive = h.n, n_lines = 0, it = 0;
  unsigned pool_used = 0, bar_target = 0;
  unsigned long long revotes = 0;
  const double reach = 1.4142136 * (0.5 + cfg.opt_dx / h.dx + 0.02), reach2 = reach * reach;
  const unsigned below = (1u << lane) - 1u;
  __syncthreads();

  auto stamp = [&](int slot, unsigned long long extra) {
    if (A.prof && tid == 0 && it < A.prof_iters)
      A.prof[((size_t)it * n_cta + cta) * kProfSlots + slot] = extra ? extra : globaltimer_ns();
  };

  for (;;) {
    stamp(0, 0);
    // ---- (A) publish this CTA's best record, meet the others, read the winner (K2) -------------------
    unsigned long long key;
    {
      unsigned long long k = 0;
      for (int r = tid; r < n_own; r += kResThreads) {
        const int l = own_dir(r, cta, n_cta);
        if (l < h.dir_count) { const unsigned long long kk = line_key(rv[r], rcell[r], h.dir_begin + l, h.num_b); k = kk > k ? kk : k; }
      }
      k = block_max_u64(k, sh.warp_key);
      if (tid == 0) { atomicMax(&A.xchg[2 * it], k); sh.tmin_i = 0x7FFFFFFF; sh.tmax_i = -0x7FFFFFFF; sh.ns = 0; sh.nlist = 0; }
      stamp(1, 0);
      grid_sync(A.bar, bar_target, (unsigned)n_cta);
      stamp(2, 0);
      key = *reinterpret_cast<volatile unsigned long long *>(&A.xchg[2 * it]);
    }
    const unsigned votes = (unsigned)(key >> 32);
    const unsigned flat = votes ? 0xFFFFFFFFu - (unsigned)key : 0u;
    double a[3], b[3];
    line_from_index(h, A.dirs, flat, a, b);                    // Hough::getLine, lib/hough.cpp:114-132
    const bool rec = cta == 0 && tid == 0 && it < cfg.max_iters;
    h3d_iter *R = rec ? &A.iters[it] : nullptr;
    if (rec) {
      R->votes = votes; R->status = 0; R->argmax = flat;
      for (int c = 0; c < 3; ++c) { R->a_raw[c] = a[c]; R->b_raw[c] = b[c]; R->a1[c] = 0; R->b1[c] = 0; R->a2[c] = 0; R->b2[c] = 0; }
      R->rc1 = 0; R->rc2 = 0; R->n1 = 0; R->n2 = 0; R->off1 = pool_used; R->off2 = pool_used;
    }
    int status = 0;
    // ---- X.pointsCloseToLine + orthogonal_LSQ, twice (K3), wrap.h:245-255; one copy of the code -------
    unsigned m2 = 0;
    int ny2 = 0;
    bool last = false;
#pragma unroll 1
    for (int pass = 0; pass < 2; ++pass) {
      int ny = 0;
      const unsigned m = inlier_pass(Q, n_alive, a, b, A.near_t, sh, pass == 1, cta == 0, A.pool, pool_used,
                                     cfg.pool_cap, A.orig, &ny,
                                     (A.prof && it < A.prof_iters) ? A.prof + ((size_t)it * n_cta + cta) * kProfSlots + 12 + pass : nullptr);
      if (rec) {
        if (pass == 0) { R->n1 = (unsigned)ny; R->off1 = pool_used; }
        else { R->n2 = (unsigned)ny; R->off2 = pool_used; }
      }
      pool_used = min(pool_used + (unsigned)ny, cfg.pool_cap);
      if (rec && pass == 0) R->off2 = pool_used;
      if (pass == 1) {
        m2 = m; ny2 = ny;
        if ((unsigned)ny2 < (unsigned)cfg.minvotes) { status = 2; break; }                 // wrap.h:251-252
        last = !((n_alive - ny2 > 1) && (cfg.nlines_max == 0 || cfg.nlines_max > n_lines + 1));   // wrap.h:341-342
      }
      stamp(3 + 2 * pass, 0);
      lsq_sums_fused(Q, ny, fx, sh);
      stamp(4 + 2 * pass, 0);
      if (warp == 0) {
        if (tid == 0) lsq_eig(sh);
      } else if (pass == 1 && !last) {
        // ---- hough->subtract(Y) for the own directions (K4), overlapped with the eigen-solve ----------
        fused_unvote(a, b, ny2, n_own, nb, n_cap, cta, n_cta, cell_off, reach2);
      }
      __syncthreads();
      if (pass == 1) stamp(14, (unsigned long long)sh.ns + 1);   // measurement: records the un-vote handed to its warp stage
      const double rc = sh.lsq.rc;
      if (rec) {
        if (pass == 0) { for (int c = 0; c < 3; ++c) { R->a1[c] = sh.lsq.a[c]; R->b1[c] = sh.lsq.b[c]; } R->rc1 = rc; }
        else { for (int c = 0; c < 3; ++c) { R->a2[c] = sh.lsq.a[c]; R->b2[c] = sh.lsq.b[c]; } R->rc2 = rc; }
      }
      if (rc == 0.0) { status = pass == 0 ? 1 : 3; break; }                                // wrap.h:248, 255
#pragma unroll
      for (int c = 0; c < 3; ++c) { a[c] = sh.lsq.a[c]; b[c] = sh.lsq.b[c]; }
    }
    stamp(7, 0);
    if (status) { if (rec) R->status = (unsigned)status; ++it; break; }
    if (rec) for (int c = 0; c < 3; ++c) R->a2[c] = __dadd_rn(sh.lsq.a[c], cfg.shift[c]);   // wrap.h:257
    ++n_lines;
    if (last) { n_alive -= ny2; ++it; break; }

    // ---- X.removePoints(Y), wrap.h:339: stable in-place compaction, slot by slot ----------------------
    // (position of a kept point = its index - inliers before it; sh.warp_cnt still holds the per-
    // (slot, warp) inlier counts of pass 2; a slot is read by everyone before anyone overwrites it)
    {
      const int slots = (n_alive + kResThreads - 1) / kResThreads;
      int run = 0;
#pragma unroll 1
      for (int p = 0; p < slots; ++p) {
        const int i = tid + p * kResThreads;
        const bool valid = i < n_alive, in = (m2 >> p) & 1u;
        double x = 0.0, y = 0.0, z = 0.0;
        unsigned o = 0;
        if (valid) { x = Q.x[i]; y = Q.y[i]; z = Q.z[i]; if (cta == 0) o = A.orig[i]; }
        const int c = lane < kResWarps ? sh.warp_cnt[p][lane] : 0;
        const int tot = __reduce_add_sync(0xFFFFFFFFu, c);
        const int pre = __reduce_add_sync(0xFFFFFFFFu, lane < warp ? c : 0);
        const unsigned bal = __ballot_sync(0xFFFFFFFFu, in);
        const int pos = i - (run + pre + __popc(bal & below));
        __syncthreads();
        if (valid && !in) { Q.x[pos] = x; Q.y[pos] = y; Q.z[pos] = z; if (cta == 0) A.orig[pos] = o; }
        run += tot;
      }
      n_alive -= ny2;
    }
    stamp(10, 0);
    // ---- (B) best EXACT record of all CTAs: stale records whose bound reaches it must be rebuilt -----
    unsigned lb;
    {
      unsigned long long fv = 0;
      for (int r = tid; r < n_own; r += kResThreads) if (!rstale[r]) fv = max(fv, (unsigned long long)rv[r]);
      fv = block_max_u64(fv, sh.warp_key);
      if (tid == 0) atomicMax(&A.xchg[2 * it + 1], fv);
      grid_sync(A.bar, bar_target, (unsigned)n_cta);
      lb = (unsigned)*reinterpret_cast<volatile unsigned long long *>(&A.xchg[2 * it + 1]);
    }
    stamp(11, 0);
    for (int r = tid; r < n_own; r += kResThreads)
      if (rstale[r] && rv[r] >= lb) rlist[atomicAdd(&sh.nlist, 1)] = r;
    __syncthreads();
    const int nlist = sh.nlist;
    stamp(8, (unsigned long long)nlist + 1);
    revotes += (unsigned long long)nlist * (unsigned long long)n_alive;
    fused_revote(nlist, n_alive, n_own, nb, n_cap, cta, n_cta);
    stamp(9, 0);
    ++it;
  }
  // ---- results --------------------------------------------------------------------------------------
  if (tid == 0) {
    if (revotes) atomicAdd(&A.st->votes_cast, revotes);
    if (cta == 0) {
      A.st->n_iters = it; A.st->n_lines = n_lines; A.st->n_alive = n_alive; A.st->pool_used = pool_used; A.st->done = 1;
    }
  }
}

}  // namespace h3d
#endif  // H3D_FUSED_CUH_
