// h3d_fused.cuh -- the whole frame loop as ONE persistent cooperative kernel (N <= 6144 points).
//   hough->add(X) + do { subtract, getLine, pointsCloseToLine, orthogonal_LSQ x2, removePoints }
//   wrap.h:228-342 / lib/hough3dlines.cpp:244-317
//
// One CTA per SM, every CTA keeps ALL live points in registers.  Division of labour:
//   * each CTA OWNS a fixed, strided set of directions (the ones it voted in the first pass); their
//     records (max votes, first max cell, stale flag) and scaled coefficients live in ITS shared
//     memory.  Un-voting removed points and re-voting stale planes are CTA-local.
//   * the line fit (inliers -> ordered compaction -> orthogonal_LSQ -> inliers -> LSQ -> remove) is
//     computed REDUNDANTLY and deterministically by every CTA from its registers, so no CTA waits
//     for another one's fit.
//   * two 8-byte exchanges per iteration, each an atomicMax + grid barrier: (A) the packed winner
//     key (Hough::getLine), (B) the best EXACT record after the un-vote, which is the bound that
//     decides which stale planes must be re-voted before the next getLine.
// Un-vote (Hough::subtract) never touches a plane: a record stays exact unless a removed point
// voted for its max cell.  The removed points Y lie within dx of the fitted line, so in the
// plane of a direction they project into a strip around a segment; a record whose max cell is
// farther than the strip's half-width from that segment cannot be hit (O(1) test per record),
// the few others are checked point by point with the exact cell arithmetic.
// While no plane is being voted the (all-zero) plane area doubles as scratch for the inlier
// coordinates and the LSQ products; it is re-zeroed before the next vote.  When enough points have
// been removed to free a register slot the live points are re-packed (order kept), so late
// iterations vote, test and scan fewer slots.
#ifndef H3D_FUSED_CUH_
#define H3D_FUSED_CUH_

#include "h3d_loop.cuh"

namespace h3d {

struct FusedArgs {
  HoughDev h;                 // n = points after the NaN strip, n_dev = nullptr
  LoopState *st;              // configuration in, final state out
  h3d_iter *iters;
  unsigned *pool;
  const unsigned *orig;       // original index of point i at kernel start
  unsigned *orig_tmp[2];      // ping-pong buffers for the re-packed numbering (CTA 0 only)
  const double *dirs;
  unsigned long long *xchg;   // zeroed: [2*it] winner key, [2*it+1] best exact votes of iteration it
  unsigned *bar;              // zeroed grid-barrier counter
  int *overflow;              // set to 1 when an inlier set does not fit the scratch area
  int n_own_max;              // records per CTA (multiple of kResG)
  unsigned smem_bytes;        // dynamic shared memory of the launch
  double near_t;              // largest nn with sqrt_rn(nn) <= opt_dx (pointsCloseToLine without the sqrt)
  double cell_off;            // max_x/dx + 0.5: cell coordinate of x' = 0
  unsigned long long *prof;   // measurement hook (nullable): [it][cta][kProfSlots] globaltimer stamps
  int prof_iters;
};
constexpr int kProfSlots = 16;

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

struct FusedShared {
  ResShared res;              // first vote (vote_resident)
  LsqOut lsq;
  LsqShared lsq_s;
  long long ipart[kResWarps][3];
  unsigned long long warp_key[kResWarps];
  int warp_cnt[12][kResWarps];
  int red_lo[kResWarps], red_hi[kResWarps];
  int tmin_i, tmax_i;         // ordered-int encodings of float t bounds of Y
  int ns, nlist;
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

// Ordered positions of the set bits of the per-thread slot masks, point order = slot-major,
// thread-minor (index i = tid + p*512).  Phase 1 publishes the warp counts (one barrier), phase 2
// lets every warp derive its own offsets with REDUX sums.  body(p, pos) runs for every set bit
// (p is a compile-time constant after unrolling).  Returns the total.  Ends WITHOUT a barrier.
template <int P, class Body>
__device__ __forceinline__ int for_each_ordered(unsigned m, int (*warp_cnt)[kResWarps], Body body) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int p = 0; p < P; ++p) {
    const unsigned bal = __ballot_sync(0xFFFFFFFFu, (m >> p) & 1u);
    if (lane == 0) warp_cnt[p][warp] = __popc(bal);
  }
  __syncthreads();
  const unsigned below = (1u << lane) - 1u;
  int run = 0;
#pragma unroll
  for (int p = 0; p < P; ++p) {
    const int c = lane < kResWarps ? warp_cnt[p][lane] : 0;
    const int tot = __reduce_add_sync(0xFFFFFFFFu, c);
    const int pre = __reduce_add_sync(0xFFFFFFFFu, lane < warp ? c : 0);
    const unsigned bal = __ballot_sync(0xFFFFFFFFu, (m >> p) & 1u);
    if ((m >> p) & 1u) body(p, run + pre + __popc(bal & below));
    run += tot;
  }
  return run;
}

// Scratch view of the idle plane area (+ whatever dynamic shared memory lies behind it):
// the current inlier set Y in cloud order, as doubles and as float casts.
struct FusedScratch {
  double *base;   // shared memory
  int cap;
  __device__ __forceinline__ double *yd(int c) const { return base + (size_t)c * cap; }
  __device__ __forceinline__ float *yf(int c) const { return reinterpret_cast<float *>(base + (size_t)3 * cap) + (size_t)c * cap; }
};

// frame-level fixed-point facts for the exact parallel anchor (see h3d_fit.cuh): every
// coordinate of the frame is a multiple of 2^lo and below 2^(hi+1); ok = both are sane
struct FixedPoint {
  int lo, hi, ok;
};

// orthogonal_LSQ sums over Y (wrap.h:126-138) by the whole CTA: anchor, float column means,
// centred scatter in the canonical TREE32 order.  Leaves sh.lsq.a and sh.lsq_s.sc ready; the
// eigen-solve (lsq_eig) is a separate single-thread step so that the caller can overlap it.
// Ends with a barrier.
__device__ __forceinline__ void lsq_sums_fused(const FusedScratch &S, int n, const FixedPoint &fx, FusedShared &sh) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const bool fast = fx.ok && (n == 0 || fx.hi + 1 - fx.lo + (32 - __clz(n)) <= 53);
  if (fast) {             // exact int64 sums, any order (provably the sequential double sum)
    const double scale = __longlong_as_double((long long)(1023 - fx.lo) << 52);   // 2^-lo
    long long acc[3] = {0, 0, 0};
#pragma unroll 1
    for (int k = tid; k < n; k += kResThreads) {
#pragma unroll
      for (int c = 0; c < 3; ++c) acc[c] += __double2ll_rn(__dmul_rn(S.yd(c)[k], scale));
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
#pragma unroll
      for (int o = 16; o; o >>= 1) acc[c] += __shfl_xor_sync(0xFFFFFFFFu, acc[c], o);
      if (lane == 0) sh.ipart[warp][c] = acc[c];
    }
  } else if (warp >= 3 && warp < 6) {
    const int c = warp - 3;
    const double *src = S.yd(c);
    const double m = mean_axis_warp([&](int k) { return src[k]; }, n, lane);
    if (lane == 0) sh.lsq.a[c] = m;
  }
  if (warp < 3) {          // colwise().mean() of the float copy (wrap.h:129-137), TREE32
    const float *src = S.yf(warp);
    float acc = 0.0f;
    int k = lane;
#pragma unroll 1
    for (; k + 96 < n; k += 128) {
      const float v0 = src[k], v1 = src[k + 32], v2 = src[k + 64], v3 = src[k + 96];
      acc = __fadd_rn(__fadd_rn(__fadd_rn(__fadd_rn(acc, v0), v1), v2), v3);
    }
#pragma unroll 1
    for (; k < n; k += 32) acc = __fadd_rn(acc, src[k]);
    acc = tree32_combine(acc);
    if (lane == 0) sh.lsq_s.mean[warp] = __fdiv_rn(acc, (float)n);
  }
  __syncthreads();
  if (warp < 6) {          // centred^T * centred (wrap.h:137-138), TREE32 per entry, products on the fly
    const int c0 = warp < 3 ? 0 : warp < 5 ? 1 : 2, c1 = warp < 3 ? warp : warp < 5 ? warp - 2 : 2;
    const float *s0 = S.yf(c0), *s1 = S.yf(c1);
    const float m0 = sh.lsq_s.mean[c0], m1 = sh.lsq_s.mean[c1];
    float acc = 0.0f;
    int k = lane;
#pragma unroll 1
    for (; k + 96 < n; k += 128) {
      const float u0 = s0[k], u1 = s0[k + 32], u2 = s0[k + 64], u3 = s0[k + 96];
      const float w0 = s1[k], w1 = s1[k + 32], w2 = s1[k + 64], w3 = s1[k + 96];
      acc = __fadd_rn(acc, __fmul_rn(__fsub_rn(u0, m0), __fsub_rn(w0, m1)));
      acc = __fadd_rn(acc, __fmul_rn(__fsub_rn(u1, m0), __fsub_rn(w1, m1)));
      acc = __fadd_rn(acc, __fmul_rn(__fsub_rn(u2, m0), __fsub_rn(w2, m1)));
      acc = __fadd_rn(acc, __fmul_rn(__fsub_rn(u3, m0), __fsub_rn(w3, m1)));
    }
#pragma unroll 1
    for (; k < n; k += 32) acc = __fadd_rn(acc, __fmul_rn(__fsub_rn(s0[k], m0), __fsub_rn(s1[k], m1)));
    acc = tree32_combine(acc);
    if (lane == 0) sh.lsq_s.sc[warp] = acc;
  } else if (warp == 6 && fast && lane < 3) {
    long long s = 0;
    for (int w = 0; w < kResWarps; ++w) s += sh.ipart[w][lane];
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

// votes of all (live) register points for one direction into the plane at shared address plane_sa
template <int P>
__device__ __forceinline__ void vote_one_direction(const HoughDev &h, const FastCoef &f, int j, const double (&px)[P],
                                                   const double (&py)[P], const double (&pz)[P], unsigned live,
                                                   unsigned plane_sa, unsigned dummy_sa) {
  const double kmagic = h.kmagic;
  const unsigned num_x = (unsigned)h.num_x, plane_cells = h.plane_cells;
  constexpr int CH = 4;
#pragma unroll
  for (int p0 = 0; p0 < P; p0 += CH) {
    if (__all_sync(0xFFFFFFFFu, (live >> p0) == 0)) break;
    unsigned cell[CH];
    unsigned gmin = 0xFFFFFFFFu, hacc = 0;
#pragma unroll
    for (int c = 0; c < CH; ++c) {
      const int p = p0 + c;
      if (p >= P) { cell[c] = 0xFFFFFFFFu; continue; }
      cell[c] = vote_cell_acc(f, kmagic, num_x, px[p], py[p], pz[p], gmin, hacc);
    }
    if (batch_unsure(gmin, hacc) | !h.fast_ok) {
#pragma unroll
      for (int c = 0; c < CH; ++c) {
        const int p = p0 + c;
        if (p < P)
          cell[c] = vote_cell_checked(h.coef + j, h.inv_s, kmagic, h.max_x, h.dx, h.num_x, h.fast_ok, px[p], py[p], pz[p]);
      }
    }
#pragma unroll
    for (int c = 0; c < CH; ++c) {
      const int p = p0 + c;
      if (p >= P) continue;
      red_shared_inc(((cell[c] < plane_cells) & ((live >> p) & 1u)) ? plane_sa + 4u * cell[c] : dummy_sa);
    }
  }
}

// X.pointsCloseToLine(a, b, dx, &Y) over the register points + ordered compaction of Y into the
// scratch area (doubles and float casts).  Returns this thread's inlier mask; *ny_out = |Y|
// (uniform).  CTA 0 also appends the original indices to the pool.  track_t: also reduce the
// range of t = b*(p - a) over Y into sh.tmin_i / sh.tmax_i.  Ends with a barrier.
template <int P>
__device__ __forceinline__ unsigned inlier_pass(const double (&px)[P], const double (&py)[P], const double (&pz)[P],
                                                unsigned live, int used_slots, const double (&a)[3], const double (&b)[3],
                                                double near_t, FusedShared &sh, const FusedScratch &S, bool track_t,
                                                bool to_pool, unsigned *pool, unsigned pool_base, unsigned pool_cap,
                                                const unsigned *orig, int *ny_out, unsigned long long *dbg) {
  const int tid = threadIdx.x;
  unsigned m = 0;
  float tlo = 3.0e38f, thi = -3.0e38f;
#pragma unroll
  for (int p = 0; p < P; ++p) {
    if (p < used_slots) {   // uniform
      double t;
      if (point_near_line_t(px[p], py[p], pz[p], a, b, near_t, &t) && ((live >> p) & 1u)) {
        m |= 1u << p;
        if (track_t) { const float tf = (float)t; tlo = fminf(tlo, tf); thi = fmaxf(thi, tf); }
      }
    }
  }
  if (dbg && tid == 0) *dbg = globaltimer_ns();
  if (track_t) {
    const int lo_i = __reduce_min_sync(0xFFFFFFFFu, float_ordered(tlo));
    const int hi_i = __reduce_max_sync(0xFFFFFFFFu, float_ordered(thi));
    if ((tid & 31) == 0) { atomicMin(&sh.tmin_i, lo_i); atomicMax(&sh.tmax_i, hi_i); }
  }
  const int cap = S.cap;
  const int ny = for_each_ordered<P>(m, sh.warp_cnt, [&](int p, int pos) {
    if (pos < cap) {
      S.yd(0)[pos] = px[p]; S.yd(1)[pos] = py[p]; S.yd(2)[pos] = pz[p];
      S.yf(0)[pos] = (float)px[p]; S.yf(1)[pos] = (float)py[p]; S.yf(2)[pos] = (float)pz[p];
    }
    if (to_pool && pool_base + (unsigned)pos < pool_cap) pool[pool_base + pos] = orig[tid + p * kResThreads];
  });
  __syncthreads();
  *ny_out = ny;
  return m;
}

// Can a point of Y vote for cell (cx, cy) of the direction with scaled coefficients f?  Y lies in
// the slab { a + t b + e : t in [t0, t1], |e| <= radius }, whose image in the (x', y') plane, in
// cell units, is within `reach`/sqrt(2)-0.5 cells of the segment; the cell is reachable only if
// its centre is within `reach` (Euclidean) of that segment.  Conservative, never exact.
__device__ __forceinline__ bool strip_may_hit(const FastCoef &f, const double (&a)[3], const double (&b)[3], double t0,
                                              double t1, double cell_off, double reach2, unsigned cx, unsigned cy) {
  const double s = 1.0 / 1048576.0;   // the coefficients carry 2^20/dx
  const double xa = (f.c0s * a[0] + f.c1n * a[1] + f.bxn * a[2]) * s + cell_off - ((double)cx + 0.5);
  const double ya = (f.c1n * a[0] + f.c2s * a[1] + f.byn * a[2]) * s + cell_off - ((double)cy + 0.5);
  const double gx = (f.c0s * b[0] + f.c1n * b[1] + f.bxn * b[2]) * s;
  const double gy = (f.c1n * b[0] + f.c2s * b[1] + f.byn * b[2]) * s;
  const double gg = gx * gx + gy * gy;
  double t = gg > 1e-30 ? -(xa * gx + ya * gy) / gg : t0;
  t = fmin(fmax(t, t0), t1);
  const double ex = xa + t * gx, ey = ya + t * gy;
  return ex * ex + ey * ey <= reach2;
}

__device__ __forceinline__ FastCoef load_fast(const double *own_coef, int r) {
  FastCoef f;
  f.c0s = own_coef[5 * r]; f.c1n = own_coef[5 * r + 1]; f.c2s = own_coef[5 * r + 2];
  f.bxn = own_coef[5 * r + 3]; f.byn = own_coef[5 * r + 4];
  return f;
}

__device__ __forceinline__ void named_bar_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" :: "r"(id), "r"(threads) : "memory");
}

template <int P>
__global__ void __launch_bounds__(kResThreads, 1) fused_loop_kernel(const FusedArgs A) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ __align__(16) FusedShared sh;
  constexpr int G = kResG;
  const HoughDev &h = A.h;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int cta = blockIdx.x, n_cta = gridDim.x;
  // per-CTA record tables first, then the planes (+ pad word), then free scratch
  const int n_own = A.n_own_max;
  double *own_coef = reinterpret_cast<double *>(smem_raw);               // 5 scaled coefficients per record
  unsigned *rv = reinterpret_cast<unsigned *>(own_coef + 5 * n_own);     // max votes (bound when stale)
  unsigned *rcell = rv + n_own;                                          // first max cell
  unsigned *rstale = rcell + n_own;
  int *rlist = reinterpret_cast<int *>(rstale + n_own);
  const unsigned tables_bytes = (unsigned)n_own * 56u;                   // n_own % 4 == 0 -> multiple of 16
  unsigned *planes = reinterpret_cast<unsigned *>(smem_raw + tables_bytes);
  const unsigned stride = h.plane_stride, quads = stride / 4;
  const unsigned plane_words = 2u * G * stride;
  const unsigned planes_sa = (unsigned)__cvta_generic_to_shared(planes);
  const unsigned dummy_sa = planes_sa + plane_words * 4u;
  FusedScratch S;
  S.cap = (int)(((A.smem_bytes - tables_bytes) / 36u) & ~1u);
  S.base = reinterpret_cast<double *>(planes);

  // ---- hough->add(X): every plane once (K1) ------------------------------------------------------
  uint4 *p4 = reinterpret_cast<uint4 *>(planes);
  for (unsigned q = tid; q < plane_words / 4 + 1; q += kResThreads) p4[q] = make_uint4(0, 0, 0, 0);
  vote_resident<P, 0>(h, nullptr, h.dir_count, planes, sh.res, cta, n_cta);
  __syncthreads();
  // own records: entry r = k*G + g is direction (cta + k*n_cta)*G + g
  for (int r = tid; r < n_own; r += kResThreads) {
    const int l = (cta + (r / G) * n_cta) * G + (r % G);
    if (l < h.dir_count) {
      rv[r] = h.dvotes[l]; rcell[r] = h.dcell[l];
      const FastCoef f = make_fast(h.coef[h.dir_begin + l], h.inv_s);
      own_coef[5 * r] = f.c0s; own_coef[5 * r + 1] = f.c1n; own_coef[5 * r + 2] = f.c2s;
      own_coef[5 * r + 3] = f.bxn; own_coef[5 * r + 4] = f.byn;
    } else { rv[r] = 0; rcell[r] = 0; }
    rstale[r] = 0;
  }
  // the points again, for the loop; and the frame's fixed-point facts (exact parallel anchor)
  double px[P], py[P], pz[P];
  unsigned live = 0;
  FixedPoint fx;
  {
    int lo = 0x7FFFFFFF, hi = -0x7FFFFFFF;
#pragma unroll
    for (int p = 0; p < P; ++p) {
      const int i = tid + p * kResThreads;
      const bool ok = i < h.n;
      px[p] = ok ? h.px[i] : 0.0; py[p] = ok ? h.py[i] : 0.0; pz[p] = ok ? h.pz[i] : 0.0;
      live |= ok ? (1u << p) : 0u;
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const long long bits = __double_as_longlong(c == 0 ? px[p] : c == 1 ? py[p] : pz[p]);
        const int e = (int)((bits >> 52) & 0x7FF);
        const long long mant = bits & 0xFFFFFFFFFFFFFLL;
        if (e == 0x7FF || (e == 0 && mant != 0)) hi = 100000;           // not finite / subnormal: no fast path
        else if (e != 0) {
          lo = min(lo, e - 1075 + (__ffsll(mant | 0x10000000000000LL) - 1));
          hi = max(hi, e - 1023);
        }
      }
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
  int n_alive = h.n, n_lines = 0, it = 0, dirty_n = 0;
  int used_slots = (h.n + kResThreads - 1) / kResThreads;
  unsigned pool_used = 0, bar_target = 0;
  unsigned long long revotes = 0;
  const unsigned *cur_orig = A.orig;
  int orig_flip = 0;
  const double reach = 1.4142136 * (0.5 + cfg.opt_dx / h.dx + 0.02), reach2 = reach * reach;
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
        const int l = (cta + (r / G) * n_cta) * G + (r % G);
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
    stamp(14, 0);
    const bool rec = cta == 0 && tid == 0 && it < cfg.max_iters;
    h3d_iter *R = rec ? &A.iters[it] : nullptr;
    if (rec) {
      R->votes = votes; R->status = 0; R->argmax = flat;
      for (int c = 0; c < 3; ++c) { R->a_raw[c] = a[c]; R->b_raw[c] = b[c]; R->a1[c] = 0; R->b1[c] = 0; R->a2[c] = 0; R->b2[c] = 0; }
      R->rc1 = 0; R->rc2 = 0; R->n1 = 0; R->n2 = 0; R->off1 = pool_used; R->off2 = pool_used;
    }
    int status = 0;
    // ---- X.pointsCloseToLine + orthogonal_LSQ, twice (K3), wrap.h:245-255 ----------------------------
    // one copy of the code for both passes (instruction-cache footprint)
    unsigned m2 = 0;
    int ny2 = 0;
    bool last = false;
#pragma unroll 1
    for (int pass = 0; pass < 2; ++pass) {
      int ny = 0;
      const unsigned m = inlier_pass<P>(px, py, pz, live, used_slots, a, b, A.near_t, sh, S, pass == 1, cta == 0, A.pool,
                                        pool_used, cfg.pool_cap, cur_orig, &ny,
                                        (A.prof && it < A.prof_iters) ? A.prof + ((size_t)it * n_cta + cta) * kProfSlots + 12 + pass : nullptr);
      dirty_n = max(dirty_n, min(ny, S.cap));
      if (ny > S.cap) { status = 9; break; }
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
      lsq_sums_fused(S, ny, fx, sh);
      stamp(4 + 2 * pass, 0);
      if (warp == 0) {
        if (tid == 0) lsq_eig(sh);
      } else if (pass == 1 && !last) {
        // ---- hough->subtract(Y) for the own directions (K4), overlapped with the eigen-solve ----------
        // Y (pass 2) is in the scratch area; it lies within opt_dx of the line (a, b) of pass 2
        const int ut = tid - 32, un = kResThreads - 32;
        const float tl = ordered_float(sh.tmin_i), th = ordered_float(sh.tmax_i);
        const double t0 = (double)tl - (fabs((double)tl) * 1e-6 + 1e-9), t1 = (double)th + (fabs((double)th) * 1e-6 + 1e-9);
        const unsigned num_x = (unsigned)h.num_x;
        for (int r = ut; r < n_own; r += un) {
          if (rstale[r] || rv[r] == 0) continue;
          const unsigned cx = rcell[r] / num_x, cy = rcell[r] - cx * num_x;
          // y_i == num_x or -1 aliases into a neighbouring row (flat bound only, lib/hough.cpp:90-92)
          if (cy == 0 || cy == num_x - 1 ||
              strip_may_hit(load_fast(own_coef, r), a, b, t0, t1, A.cell_off, reach2, cx, cy))
            rlist[atomicAdd(&sh.ns, 1)] = r;
        }
        named_bar_sync(1, un);
        const int ns = sh.ns;
        const int total = ns * ny2;
        for (int idx = ut; idx < total; idx += un) {
          const int s = idx / ny2, e = idx - s * ny2;
          const int r = rlist[s];
          const int j = h.dir_begin + (cta + (r / G) * n_cta) * G + (r % G);
          if (vote_cell(h, load_fast(own_coef, r), j, S.yd(0)[e], S.yd(1)[e], S.yd(2)[e]) == rcell[r]) rstale[r] = 1;
        }
      }
      __syncthreads();
      const double rc = sh.lsq.rc;
      if (rec) {
        if (pass == 0) { for (int c = 0; c < 3; ++c) { R->a1[c] = sh.lsq.a[c]; R->b1[c] = sh.lsq.b[c]; } R->rc1 = rc; }
        else { for (int c = 0; c < 3; ++c) { R->a2[c] = sh.lsq.a[c]; R->b2[c] = sh.lsq.b[c]; } R->rc2 = rc; }
      }
      if (rc == 0.0) { status = pass == 0 ? 1 : 3; break; }                                // wrap.h:248, 255
#pragma unroll
      for (int c = 0; c < 3; ++c) { a[c] = sh.lsq.a[c]; b[c] = sh.lsq.b[c]; }
    }
    if (status == 9) { if (tid == 0) *A.overflow = 1; break; }
    stamp(7, 0);
    if (status) { if (rec) R->status = (unsigned)status; ++it; break; }
    if (rec) for (int c = 0; c < 3; ++c) R->a2[c] = __dadd_rn(sh.lsq.a[c], cfg.shift[c]);   // wrap.h:257
    ++n_lines;
    live &= ~m2;                                                                           // X.removePoints(Y)
    n_alive -= ny2;
    if (last) { ++it; break; }

    // ---- give the plane area back: zero what the scratch used ----------------------------------------
    for (int q = tid; q < dirty_n; q += kResThreads) {
      S.yd(0)[q] = 0.0; S.yd(1)[q] = 0.0; S.yd(2)[q] = 0.0;
      S.yf(0)[q] = 0.0f; S.yf(1)[q] = 0.0f; S.yf(2)[q] = 0.0f;
    }
    dirty_n = 0;
    // ---- re-pack the live points when a register slot can be freed (order kept) ----------------------
    const int new_slots = (n_alive + kResThreads - 1) / kResThreads;
    if (new_slots < used_slots) {
      // one ordered pass stages x, y, z (and, on CTA 0, the numbering) by new position
      double *stage = reinterpret_cast<double *>(planes);
      const int n_old = used_slots * kResThreads;   // >= the old live count: stride of the staging arrays
      unsigned *odst = A.orig_tmp[orig_flip];
      const unsigned *osrc = cur_orig;
      __syncthreads();   // zero-fill above is complete before the stage is written
      (void)for_each_ordered<P>(live, sh.warp_cnt, [&](int p, int pos) {
        stage[pos] = px[p]; stage[n_old + pos] = py[p]; stage[2 * n_old + pos] = pz[p];
        if (cta == 0) odst[pos] = osrc[tid + p * kResThreads];
      });
      __syncthreads();
      live = 0;
#pragma unroll
      for (int p = 0; p < P; ++p) {
        const int i = tid + p * kResThreads;
        const bool ok = i < n_alive;
        px[p] = ok ? stage[i] : 0.0; py[p] = ok ? stage[n_old + i] : 0.0; pz[p] = ok ? stage[2 * n_old + i] : 0.0;
        live |= ok ? (1u << p) : 0u;
      }
      cur_orig = odst;
      orig_flip ^= 1;
      __syncthreads();
      for (int q = tid; q < n_alive; q += kResThreads) { stage[q] = 0.0; stage[n_old + q] = 0.0; stage[2 * n_old + q] = 0.0; }
      used_slots = new_slots;
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
    for (int b0 = 0; b0 < nlist; b0 += 2 * G) {
      const int nb = min(2 * G, nlist - b0);
      for (int bi = 0; bi < nb; ++bi) {
        const int r = rlist[b0 + bi];
        const int j = h.dir_begin + (cta + (r / G) * n_cta) * G + (r % G);
        vote_one_direction<P>(h, load_fast(own_coef, r), j, px, py, pz, live, planes_sa + (unsigned)bi * stride * 4u, dummy_sa);
      }
      __syncthreads();
      for (int bi = 0; bi < nb; ++bi) {
        const unsigned long long k = scan_plane_fast<kResThreads>(planes + (size_t)bi * stride, quads, tid);
        if (lane == 0) sh.res.warp_best[bi / G][bi % G][warp] = k;
      }
      __syncthreads();
      if (warp < nb) {
        unsigned long long k = lane < kResWarps ? sh.res.warp_best[warp / G][warp % G][lane] : 0ull;
        k = warp_max_u64(k);
        if (lane == 0) {
          const int r = rlist[b0 + warp];
          const unsigned v = (unsigned)(k >> 32);
          rv[r] = v; rcell[r] = v ? 0xFFFFFFFFu - (unsigned)k : 0u; rstale[r] = 0;
        }
      }
      __syncthreads();
    }
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
