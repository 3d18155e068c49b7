// h3d_fused.cuh -- the whole frame loop as ONE persistent cooperative kernel (N <= 6144 points).
//   hough->add(X) + do { subtract, getLine, pointsCloseToLine, orthogonal_LSQ x2, removePoints }
//   wrap.h:228-342 / lib/hough3dlines.cpp:244-317
//
// One CTA per SM, every CTA keeps ALL points in registers.  Division of labour:
//   * each CTA OWNS a fixed, strided set of directions (the ones it voted in the first pass); their
//     records (max votes, first max cell, stale flag) and scaled coefficients live in ITS shared
//     memory.  Un-voting removed points and re-voting stale planes are CTA-local.
//   * the line fit (inliers -> ordered compaction -> sequential-order LSQ -> inliers -> LSQ ->
//     remove) is computed REDUNDANTLY and deterministically by every CTA from its registers, so
//     no CTA waits for another one's fit.
//   * the only exchange per iteration is one 8-byte atomicMax of the packed winner key plus one
//     grid barrier.
// A stale plane is re-voted when its bound reaches the CTA's own best exact record (a weaker but
// barrier-free bound than the global best exact record; everything left stale is still below
// the global winner, so the result is unchanged).
// While no plane is being voted the (all-zero) plane area doubles as scratch for the inlier
// coordinates and the LSQ products; it is re-zeroed before the next vote.
#ifndef H3D_FUSED_CUH_
#define H3D_FUSED_CUH_

#include "h3d_loop.cuh"

namespace h3d {

struct FusedArgs {
  HoughDev h;                 // n = points after the NaN strip, n_dev = nullptr
  LoopState *st;              // configuration in, final state out
  h3d_iter *iters;
  unsigned *pool;
  const unsigned *orig;
  const double *dirs;
  unsigned long long *gbest;  // one zeroed slot per iteration (+1)
  unsigned *bar;              // zeroed grid-barrier counter
  int *overflow;              // set to 1 when an inlier set does not fit the scratch area
  int n_own_max;              // records per CTA (multiple of kResG)
  unsigned long long *prof;   // measurement hook (nullable): [it][cta][kProfSlots] globaltimer stamps
  int prof_iters;
};
constexpr int kProfSlots = 10;

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

struct FusedShared {
  ResShared res;              // first vote (vote_resident)
  double a[3], b[3];          // current line
  LsqOut lsq;
  LsqShared lsq_s;
  unsigned long long key;
  unsigned long long warp_key[kResWarps];
  int slot_base[16];          // ordered-compaction offsets of the point slots
  int warp_cnt[16][kResWarps];
  int ny, nlist, flag;
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
    if ((live >> p0) == 0) break;
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

// pointsCloseToLine over the register points + ordered compaction of the inliers' coordinates into
// shared memory (order = point index = slot-major, thread-minor).  Returns this thread's inlier
// mask; sh.ny = |Y|.  CTA 0 also appends the original indices to the pool.
template <int P>
__device__ __forceinline__ unsigned inlier_pass(const double (&px)[P], const double (&py)[P], const double (&pz)[P],
                                                unsigned live, double dx, FusedShared &sh, double *ypx, double *ypy,
                                                double *ypz, int ycap, bool to_pool, unsigned *pool,
                                                unsigned pool_base, unsigned pool_cap, const unsigned *orig) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const double a[3] = {sh.a[0], sh.a[1], sh.a[2]}, b[3] = {sh.b[0], sh.b[1], sh.b[2]};
  unsigned m = 0;
#pragma unroll
  for (int p = 0; p < P; ++p)
    if (((live >> p) & 1u) && point_near_line(px[p], py[p], pz[p], a, b, dx)) m |= 1u << p;
#pragma unroll
  for (int p = 0; p < P; ++p) {
    const unsigned bal = __ballot_sync(0xFFFFFFFFu, (m >> p) & 1u);
    if (lane == 0) sh.warp_cnt[p][warp] = __popc(bal);
  }
  __syncthreads();
  if (tid < P) {
    int run = 0;
    for (int w = 0; w < kResWarps; ++w) { const int t = sh.warp_cnt[tid][w]; sh.warp_cnt[tid][w] = run; run += t; }
    sh.slot_base[tid] = run;
  }
  __syncthreads();
  if (tid == 0) {
    int run = 0;
    for (int p = 0; p < P; ++p) { const int t = sh.slot_base[p]; sh.slot_base[p] = run; run += t; }
    sh.ny = run;
  }
  __syncthreads();
  const unsigned below = (1u << lane) - 1u;
#pragma unroll
  for (int p = 0; p < P; ++p) {
    const bool in = (m >> p) & 1u;
    const unsigned bal = __ballot_sync(0xFFFFFFFFu, in);
    if (in) {
      const int pos = sh.slot_base[p] + sh.warp_cnt[p][warp] + __popc(bal & below);
      if (pos < ycap) { ypx[pos] = px[p]; ypy[pos] = py[p]; ypz[pos] = pz[p]; }
      if (to_pool && pool_base + (unsigned)pos < pool_cap) pool[pool_base + pos] = orig[tid + p * kResThreads];
    }
  }
  __syncthreads();
  return m;
}


template <int P>
__global__ void __launch_bounds__(kResThreads, 1) fused_loop_kernel(const FusedArgs A) {
  extern __shared__ __align__(16) unsigned planes[];
  __shared__ __align__(16) FusedShared sh;
  constexpr int G = kResG;
  const HoughDev &h = A.h;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int cta = blockIdx.x, n_cta = gridDim.x;
  const unsigned stride = h.plane_stride, quads = stride / 4;
  const unsigned plane_words = 2u * G * stride;
  const unsigned planes_sa = (unsigned)__cvta_generic_to_shared(planes);
  const unsigned dummy_sa = planes_sa + plane_words * 4u;
  // per-CTA tables behind the planes and their pad
  const int n_own = A.n_own_max;
  double *own_coef = reinterpret_cast<double *>(planes + plane_words + 4);    // 5 scaled coefficients per record
  unsigned *rv = reinterpret_cast<unsigned *>(own_coef + 5 * n_own);          // max votes (bound when stale)
  unsigned *rcell = rv + n_own;                                               // first max cell
  unsigned *rstale = rcell + n_own;
  int *rlist = reinterpret_cast<int *>(rstale + n_own);
  // scratch inside the (zero) plane area while no vote is running
  const int ycap = (int)((plane_words * 4u - 6u * kResThreads * 4u) / 24u);
  double *ypx = reinterpret_cast<double *>(planes), *ypy = ypx + ycap, *ypz = ypy + ycap;
  float *prod = reinterpret_cast<float *>(ypz + ycap);

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
  // the points again, for the loop
  double px[P], py[P], pz[P];
  unsigned live = 0;
#pragma unroll
  for (int p = 0; p < P; ++p) {
    const int i = tid + p * kResThreads;
    const bool ok = i < h.n;
    px[p] = ok ? h.px[i] : 0.0; py[p] = ok ? h.py[i] : 0.0; pz[p] = ok ? h.pz[i] : 0.0;
    live |= ok ? (1u << p) : 0u;
  }
  const LoopState cfg = *A.st;
  int n_alive = h.n, n_lines = 0, it = 0, max_scratch = 0;
  unsigned pool_used = 0, bar_target = 0;
  unsigned long long revotes = 0;
  __syncthreads();

  auto stamp = [&](int slot, unsigned long long extra) {
    if (A.prof && tid == 0 && it < A.prof_iters)
      A.prof[((size_t)it * n_cta + cta) * kProfSlots + slot] = extra ? extra : globaltimer_ns();
  };
  for (;;) {
    stamp(0, 0);
    // ---- publish this CTA's best record, meet the others, read the winner (K2) -----------------------
    {
      unsigned long long k = 0;
      for (int r = tid; r < n_own; r += kResThreads) {
        const int l = (cta + (r / G) * n_cta) * G + (r % G);
        if (l < h.dir_count) { const unsigned long long kk = line_key(rv[r], rcell[r], h.dir_begin + l, h.num_b); k = kk > k ? kk : k; }
      }
      k = block_max_u64(k, sh.warp_key);
      if (tid == 0) atomicMax(&A.gbest[it], k);
      stamp(1, 0);
      grid_sync(A.bar, bar_target, (unsigned)n_cta);
      stamp(2, 0);
      if (tid == 0) {
        const unsigned long long key = *reinterpret_cast<volatile unsigned long long *>(&A.gbest[it]);
        const unsigned votes = (unsigned)(key >> 32);
        const unsigned flat = votes ? 0xFFFFFFFFu - (unsigned)key : 0u;
        sh.key = key;
        line_from_index(h, A.dirs, flat, sh.a, sh.b);          // Hough::getLine, lib/hough.cpp:114-132
      }
      __syncthreads();
    }
    const bool rec = cta == 0 && it < cfg.max_iters;
    h3d_iter *R = rec ? &A.iters[it] : nullptr;
    if (rec && tid == 0) {
      const unsigned votes = (unsigned)(sh.key >> 32);
      R->votes = votes; R->status = 0; R->argmax = votes ? 0xFFFFFFFFu - (unsigned)sh.key : 0u;
      for (int c = 0; c < 3; ++c) { R->a_raw[c] = sh.a[c]; R->b_raw[c] = sh.b[c]; R->a1[c] = 0; R->b1[c] = 0; R->a2[c] = 0; R->b2[c] = 0; }
      R->rc1 = 0; R->rc2 = 0; R->n1 = 0; R->n2 = 0; R->off1 = pool_used; R->off2 = pool_used;
    }
    int status = 0;
    // ---- X.pointsCloseToLine + orthogonal_LSQ, twice (K3), wrap.h:245-255 ----------------------------
    (void)inlier_pass<P>(px, py, pz, live, cfg.opt_dx, sh, ypx, ypy, ypz, ycap, cta == 0, A.pool, pool_used, cfg.pool_cap, A.orig);
    const int ny1 = sh.ny;
    max_scratch = max(max_scratch, min(ny1, ycap));
    if (ny1 > ycap) { if (tid == 0) *A.overflow = 1; break; }
    if (rec && tid == 0) { R->n1 = (unsigned)ny1; R->off1 = pool_used; }
    pool_used = min(pool_used + (unsigned)ny1, cfg.pool_cap);
    if (rec && tid == 0) R->off2 = pool_used;
    stamp(3, 0);
    lsq_cta([&](int k, int c) { return (c == 0 ? ypx : c == 1 ? ypy : ypz)[k]; }, ny1, sh.lsq_s, &sh.lsq);
    stamp(4, 0);
    if (tid == 0) for (int c = 0; c < 3; ++c) { sh.a[c] = sh.lsq.a[c]; sh.b[c] = sh.lsq.b[c]; }
    if (rec && tid == 0) { for (int c = 0; c < 3; ++c) { R->a1[c] = sh.lsq.a[c]; R->b1[c] = sh.lsq.b[c]; } R->rc1 = sh.lsq.rc; }
    __syncthreads();
    unsigned m2 = 0;
    int ny2 = 0;
    if (sh.lsq.rc == 0.0) status = 1;                                                      // wrap.h:248
    if (!status) {
      m2 = inlier_pass<P>(px, py, pz, live, cfg.opt_dx, sh, ypx, ypy, ypz, ycap, cta == 0, A.pool, pool_used, cfg.pool_cap, A.orig);
      ny2 = sh.ny;
      max_scratch = max(max_scratch, min(ny2, ycap));
      if (ny2 > ycap) { if (tid == 0) *A.overflow = 1; break; }
      if (rec && tid == 0) { R->n2 = (unsigned)ny2; R->off2 = pool_used; }
      pool_used = min(pool_used + (unsigned)ny2, cfg.pool_cap);
      if ((unsigned)ny2 < (unsigned)cfg.minvotes) status = 2;                              // wrap.h:251-252
    }
    stamp(5, 0);
    if (!status) {
      lsq_cta([&](int k, int c) { return (c == 0 ? ypx : c == 1 ? ypy : ypz)[k]; }, ny2, sh.lsq_s, &sh.lsq);
      if (rec && tid == 0) { for (int c = 0; c < 3; ++c) { R->a2[c] = sh.lsq.a[c]; R->b2[c] = sh.lsq.b[c]; } R->rc2 = sh.lsq.rc; }
      if (sh.lsq.rc == 0.0) status = 3;                                                    // wrap.h:255
    }
    stamp(6, 0);
    if (status) { if (rec && tid == 0) R->status = (unsigned)status; ++it; break; }
    if (rec && tid == 0) for (int c = 0; c < 3; ++c) R->a2[c] = __dadd_rn(sh.lsq.a[c], cfg.shift[c]);   // wrap.h:257
    ++n_lines;
    live &= ~m2;                                                                           // X.removePoints(Y)
    n_alive -= ny2;
    if (!((n_alive > 1) && (cfg.nlines_max == 0 || cfg.nlines_max > n_lines))) { ++it; break; }     // wrap.h:341-342

    // ---- hough->subtract(Y) for the own directions (K4): Y is still in the scratch area ---------------
    {
      const int n_chunks = (ny2 + kUnvoteChunk - 1) / kUnvoteChunk;
      for (int t = tid; t < n_own * n_chunks; t += kResThreads) {
        const int r = t % n_own, chunk = t / n_own;
        if (rstale[r] || rv[r] == 0) continue;
        const int j = h.dir_begin + (cta + (r / G) * n_cta) * G + (r % G);
        FastCoef f;
        f.c0s = own_coef[5 * r]; f.c1n = own_coef[5 * r + 1]; f.c2s = own_coef[5 * r + 2];
        f.bxn = own_coef[5 * r + 3]; f.byn = own_coef[5 * r + 4];
        const unsigned target = rcell[r];
        const int e0 = chunk * kUnvoteChunk, e1 = min(e0 + kUnvoteChunk, ny2);
        bool hit = false;
        for (int e = e0; e < e1; ++e) hit |= vote_cell(h, f, j, ypx[e], ypy[e], ypz[e]) == target;
        if (hit) rstale[r] = 1;
      }
    }
    __syncthreads();
    stamp(7, 0);
    // give the plane area back: zero what the scratch used
    for (int q = tid; q < max_scratch; q += kResThreads) { ypx[q] = 0.0; ypy[q] = 0.0; ypz[q] = 0.0; }
    for (int q = tid; q < 6 * kResThreads; q += kResThreads) prod[q] = 0.0f;
    max_scratch = 0;
    // stale records whose bound reaches this CTA's best exact record must be rebuilt
    {
      unsigned fv = 0;
      for (int r = tid; r < n_own; r += kResThreads) if (!rstale[r]) fv = max(fv, rv[r]);
      fv = __reduce_max_sync(0xFFFFFFFFu, fv);
      if (lane == 0) sh.warp_cnt[0][warp] = (int)fv;
      if (tid == 0) sh.nlist = 0;
      __syncthreads();
      fv = 0;
      for (int w = 0; w < kResWarps; ++w) fv = max(fv, (unsigned)sh.warp_cnt[0][w]);
      for (int r = tid; r < n_own; r += kResThreads)
        if (rstale[r] && rv[r] >= fv) rlist[atomicAdd(&sh.nlist, 1)] = r;
      __syncthreads();
    }
    const int nlist = sh.nlist;
    stamp(8, (unsigned long long)nlist + 1);
    revotes += (unsigned long long)nlist * (unsigned long long)n_alive;
    for (int b0 = 0; b0 < nlist; b0 += 2 * G) {
      const int nb = min(2 * G, nlist - b0);
      for (int bi = 0; bi < nb; ++bi) {
        const int r = rlist[b0 + bi];
        FastCoef f;
        f.c0s = own_coef[5 * r]; f.c1n = own_coef[5 * r + 1]; f.c2s = own_coef[5 * r + 2];
        f.bxn = own_coef[5 * r + 3]; f.byn = own_coef[5 * r + 4];
        const int j = h.dir_begin + (cta + (r / G) * n_cta) * G + (r % G);
        vote_one_direction<P>(h, f, j, px, py, pz, live, planes_sa + (unsigned)bi * stride * 4u, dummy_sa);
      }
      __syncthreads();
      for (int bi = 0; bi < nb; ++bi) {
        const unsigned long long key = scan_plane_fast<kResThreads>(planes + (size_t)bi * stride, quads, tid);
        if (lane == 0) sh.res.warp_best[bi / G][bi % G][warp] = key;
      }
      __syncthreads();
      if (warp < nb) {
        unsigned long long key = lane < kResWarps ? sh.res.warp_best[warp / G][warp % G][lane] : 0ull;
        key = warp_max_u64(key);
        if (lane == 0) {
          const int r = rlist[b0 + warp];
          const unsigned votes = (unsigned)(key >> 32);
          rv[r] = votes; rcell[r] = votes ? 0xFFFFFFFFu - (unsigned)key : 0u; rstale[r] = 0;
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
