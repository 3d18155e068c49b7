// h3d_fit.cuh -- K3: point-to-line distance / inlier mask / ordered compaction and the
// orthogonal least-squares refit (sm_100a).
//   PointCloud::pointsCloseToLine   lib/pointcloud.cpp:94-105
//   PointCloud::meanValue           lib/pointcloud.cpp:29-38
//   orthogonal_LSQ                  wrap.h:121-150 (== lib/hough3dlines.cpp:49-77)
//
// Parity contract.  The inlier test is evaluated with the reference's operation order in
// IEEE double without FMA, so the inlier SETS are bit-exact.  The refit feeds the second
// inlier test, so it must be bit-stable too: the anchor is the reference's sequential double
// sum in point order; the direction follows the float pipeline of wrap.h:129-147 with the
// canonical order fixed by the oracle: every float sum (column means, scatter products) is
// TREE32 -- 32 interleaved partial sums (partial l adds elements l, l+32, l+64, ... in index
// order) combined by a halving tree (l += l+16, l+8, l+4, l+2, l+1), the shape of a vectorised
// reduction like Eigen's own -- followed by Eigen's SelfAdjointEigenSolver<MatrixXf> = scale,
// Householder tridiagonalisation, implicit symmetric QR with Wilkinson shift, ascending sort.
// Eigen is not vendored by the reference: PARITY UNPINNED against Eigen itself, bit-exact
// against oracle/h3d_oracle.c.
//
// The anchor IS pinned (sequential double sum, lib/pointcloud.cpp:29-38).  It is evaluated in
// parallel whenever that is provably the same number: if every coordinate is a multiple of
// 2^lo and n * 2^(hi+1) <= 2^(53+lo) (hi = largest exponent), every partial sum of ANY order is
// an exactly representable multiple of 2^lo, so the sequential sum equals the exact sum, which
// an int64 reduction delivers.  Otherwise the warp runs the sequential chain (loads batched, adds in order).
#ifndef H3D_FIT_CUH_
#define H3D_FIT_CUH_

#include <float.h>

#include "h3d_device.cuh"

namespace h3d {

// ---- pointsCloseToLine: is p within dx of the line (a, b)?  lib/pointcloud.cpp:99-101 ----
__device__ __forceinline__ bool point_near_line(double px, double py, double pz, const double a[3],
                                                const double b[3], double dx) {
  const double ex = __dsub_rn(px, a[0]), ey = __dsub_rn(py, a[1]), ez = __dsub_rn(pz, a[2]);
  // t = b * (p - a): (x.x*y.x + x.y*y.y + x.z*y.z), lib/vector3d.cpp:57-59
  const double t = __dadd_rn(__dadd_rn(__dmul_rn(b[0], ex), __dmul_rn(b[1], ey)), __dmul_rn(b[2], ez));
  // d = p - (a + t*b)
  const double qx = __dadd_rn(a[0], __dmul_rn(t, b[0]));
  const double qy = __dadd_rn(a[1], __dmul_rn(t, b[1]));
  const double qz = __dadd_rn(a[2], __dmul_rn(t, b[2]));
  const double dxv = __dsub_rn(px, qx), dyv = __dsub_rn(py, qy), dzv = __dsub_rn(pz, qz);
  const double nn = __dadd_rn(__dadd_rn(__dmul_rn(dxv, dxv), __dmul_rn(dyv, dyv)), __dmul_rn(dzv, dzv));
  return __dsqrt_rn(nn) <= dx;   // inclusive, lib/pointcloud.cpp:101
}

// ---- 3x3 symmetric float eigen-solve (Eigen::SelfAdjointEigenSolver<MatrixXf> restated) ----
// m: column-major symmetric 3x3 (lower triangle used).  evals ascending, evecs column-major.
// Built with -fmad=false: every product and sum below is rounded separately.
struct Eig3f {
  float evals[3];
  float evecs[9];
  int ok;
};

// IEEE division / square root out of line on the device: eig3f has ~25 of them and runs on one
// thread inside persistent kernels whose loop body must stay small (32 KB instruction cache)
#ifdef __CUDA_ARCH__
__device__ __forceinline__ float eig_div_dev(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float eig_sqrt_dev(float a) { return __fsqrt_rn(a); }
#define H3D_EIG_DIV(a, b) eig_div_dev((a), (b))
#define H3D_EIG_SQRT(a) eig_sqrt_dev((a))
#define H3D_EIG_NOINLINE __noinline__
#else
#define H3D_EIG_DIV(a, b) ((a) / (b))
#define H3D_EIG_SQRT(a) sqrtf((a))
#define H3D_EIG_NOINLINE
#endif

__host__ __device__ H3D_EIG_NOINLINE inline void eig_givens(float p, float q, float *c, float *s) {   // JacobiRotation::makeGivens
  if (q == 0.0f) { *c = p < 0.0f ? -1.0f : 1.0f; *s = 0.0f; return; }
  if (p == 0.0f) { *c = 0.0f; *s = q < 0.0f ? 1.0f : -1.0f; return; }
  if (fabsf(p) > fabsf(q)) {
    const float t = H3D_EIG_DIV(q, p);
    float u = H3D_EIG_SQRT(1.0f + t * t);
    if (p < 0.0f) u = -u;
    *c = H3D_EIG_DIV(1.0f, u);
    *s = -t * *c;
  } else {
    const float t = H3D_EIG_DIV(p, q);
    float u = H3D_EIG_SQRT(1.0f + t * t);
    if (q < 0.0f) u = -u;
    *s = H3D_EIG_DIV(-1.0f, u);
    *c = -t * *s;
  }
}

__host__ __device__ inline float eig_hypot(float x, float y) {   // numext::hypot
  const float ax = fabsf(x), ay = fabsf(y);
  const float big = ax > ay ? ax : ay;
  if (big == 0.0f) return 0.0f;
  const float small = ax > ay ? ay : ax;
  const float r = H3D_EIG_DIV(small, big);
  return big * H3D_EIG_SQRT(1.0f + r * r);
}

// FIXED3: the tridiagonalisation Eigen selects for a fixed-size Matrix3f
// (tridiagonalization_inplace_selector<MatrixType,3,false>, closed form; pcl::PCA's solver):
// eigenvalues only, evecs are not meaningful.
template <bool FIXED3 = false>
__host__ __device__ H3D_EIG_NOINLINE inline Eig3f eig3f(const float m[9]) {
  Eig3f out;
  float a00 = m[0], a10 = m[1], a20 = m[2], a11 = m[4], a21 = m[5], a22 = m[8];
  // scale the matrix by its largest coefficient
  float scale = fabsf(a00);
  if (fabsf(a10) > scale) scale = fabsf(a10);
  if (fabsf(a20) > scale) scale = fabsf(a20);
  if (fabsf(a11) > scale) scale = fabsf(a11);
  if (fabsf(a21) > scale) scale = fabsf(a21);
  if (fabsf(a22) > scale) scale = fabsf(a22);
  if (scale == 0.0f) scale = 1.0f;
  a00 = H3D_EIG_DIV(a00, scale); a10 = H3D_EIG_DIV(a10, scale); a20 = H3D_EIG_DIV(a20, scale);
  a11 = H3D_EIG_DIV(a11, scale); a21 = H3D_EIG_DIV(a21, scale); a22 = H3D_EIG_DIV(a22, scale);

  // tridiagonalisation: one Householder reflection on column 0
  float diag[3], sub[2];
  float q[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  const float tail_sq = a20 * a20;
  if (FIXED3) {
    diag[0] = a00;
    if (tail_sq <= FLT_MIN) {
      diag[1] = a11; diag[2] = a22; sub[0] = a10; sub[1] = a21;
    } else {
      const float beta = H3D_EIG_SQRT(a10 * a10 + tail_sq);
      const float inv_beta = H3D_EIG_DIV(1.0f, beta);
      const float m01 = a10 * inv_beta, m02 = a20 * inv_beta;
      const float qq = 2.0f * m01 * a21 + m02 * (a22 - a11);
      diag[1] = a11 + m02 * qq;
      diag[2] = a22 - m02 * qq;
      sub[0] = beta;
      sub[1] = a21 - m01 * qq;
    }
  } else if (tail_sq <= FLT_MIN) {
    diag[0] = a00; diag[1] = a11; diag[2] = a22; sub[0] = a10; sub[1] = a21;
  } else {
    float beta = H3D_EIG_SQRT(a10 * a10 + tail_sq);
    if (a10 >= 0.0f) beta = -beta;
    const float v1 = H3D_EIG_DIV(a20, a10 - beta);
    const float tau = H3D_EIG_DIV(beta - a10, beta);
    const float u0 = tau, u1 = tau * v1;
    const float p0 = a11 * u0 + a21 * u1;
    const float p1 = a21 * u0 + a22 * u1;
    const float dot = p0 + p1 * v1;
    const float alpha = tau * -0.5f * dot;
    const float w0 = p0 + alpha, w1 = p1 + alpha * v1;
    a11 = a11 - (w0 + w0);
    a21 = a21 - (v1 * w0 + w1);
    a22 = a22 - (v1 * w1 + w1 * v1);
    diag[0] = a00; diag[1] = a11; diag[2] = a22; sub[0] = beta; sub[1] = a21;
    const float tv = tau * v1;
    q[4] = 1.0f - tau; q[7] = 0.0f - tau * v1;
    q[5] = 0.0f - tv;  q[8] = 1.0f - tv * v1;
  }

  // implicit symmetric QR iteration on the tridiagonal matrix (Eigen computeFromTridiagonal_impl
  // + tridiagonal_qr_step), written out for n = 3 on scalars so that everything stays in
  // registers: the active block [start, end] is [0,2], [1,2] or [0,1]
  float d0 = diag[0], d1 = diag[1], d2 = diag[2], s0 = sub[0], s1 = sub[1];
  float q0 = q[0], q1 = q[1], q2 = q[2], q3 = q[3], q4 = q[4], q5 = q[5], q6 = q[6], q7 = q[7], q8 = q[8];
  int end = 2, start = 0, iter = 0;
  const float precision = 2.0f * FLT_EPSILON, consider_zero = FLT_MIN;
  int ok = 1;
  // one Givens rotation of the QR step on rows/columns (k, k+1); dk, dk1, sk by reference
#define H3D_EIG_ROT(dk, dk1, sk, qa0, qa1, qa2, qb0, qb1, qb2, PRE, POST)          \
  {                                                                                \
    float c, sn;                                                                   \
    eig_givens(x, z, &c, &sn);                                                     \
    const float sdk = sn * dk + c * sk;                                            \
    const float dkp1 = sn * sk + c * dk1;                                          \
    dk = c * (c * dk - sn * sk) - sn * (c * sk - sn * dk1);                        \
    dk1 = sn * sdk + c * dkp1;                                                     \
    sk = c * sdk - sn * dkp1;                                                      \
    PRE;                                                                           \
    x = sk;                                                                        \
    POST;                                                                          \
    float xi, yi;                                                                  \
    xi = qa0; yi = qb0; qa0 = c * xi - sn * yi; qb0 = sn * xi + c * yi;            \
    xi = qa1; yi = qb1; qa1 = c * xi - sn * yi; qb1 = sn * xi + c * yi;            \
    xi = qa2; yi = qb2; qa2 = c * xi - sn * yi; qb2 = sn * xi + c * yi;            \
  }
  while (end > 0) {
    if (start <= 0 && end > 0 && (fabsf(s0) <= (fabsf(d0) + fabsf(d1)) * precision || fabsf(s0) <= consider_zero)) s0 = 0.0f;
    if (start <= 1 && end > 1 && (fabsf(s1) <= (fabsf(d1) + fabsf(d2)) * precision || fabsf(s1) <= consider_zero)) s1 = 0.0f;
    if (end == 2 && s1 == 0.0f) end = 1;
    if (end == 1 && s0 == 0.0f) end = 0;
    if (end <= 0) break;
    iter++;
    if (iter > 30 * 3) { ok = 0; break; }
    start = (end == 2 && s0 == 0.0f) ? 1 : 0;

    const float de1 = end == 2 ? d1 : d0, de = end == 2 ? d2 : d1;   // diag[end-1], diag[end]
    const float td = (de1 - de) * 0.5f;
    const float e = end == 2 ? s1 : s0;                              // sub[end-1]
    float mu = de;
    if (td == 0.0f) {
      mu -= fabsf(e);
    } else {
      const float e2 = e * e;
      const float hyp = eig_hypot(td, e);
      if (e2 == 0.0f) mu -= H3D_EIG_DIV(e, td + (td > 0.0f ? 1.0f : -1.0f)) * H3D_EIG_DIV(e, hyp);
      else mu -= H3D_EIG_DIV(e2, td + (td > 0.0f ? hyp : -hyp));
    }
    float x = (start == 0 ? d0 : d1) - mu;
    float z = start == 0 ? s0 : s1;
    if (start == 0) H3D_EIG_ROT(d0, d1, s0, q0, q1, q2, q3, q4, q5, (void)0, if (end == 2) { z = -sn * s1; s1 = c * s1; })
    if (end == 2) H3D_EIG_ROT(d1, d2, s1, q3, q4, q5, q6, q7, q8, if (start == 0) s0 = c * s0 - sn * z, (void)0)
  }
#undef H3D_EIG_ROT
  diag[0] = d0; diag[1] = d1; diag[2] = d2;
  q[0] = q0; q[1] = q1; q[2] = q2; q[3] = q3; q[4] = q4; q[5] = q5; q[6] = q6; q[7] = q7; q[8] = q8;
  if (ok) {   // ascending selection sort, eigenvector columns follow
#pragma unroll
    for (int i = 0; i < 2; i++) {
      int k = i;
      float least = diag[i];
#pragma unroll
      for (int j = i + 1; j < 3; j++) if (diag[j] < least) { least = diag[j]; k = j; }
      if (k > i) {
#pragma unroll
        for (int kk = 1; kk < 3; kk++) {
          if (kk == k && kk > i) {
            const float t = diag[i]; diag[i] = diag[kk]; diag[kk] = t;
#pragma unroll
            for (int r = 0; r < 3; r++) { const float u = q[i * 3 + r]; q[i * 3 + r] = q[kk * 3 + r]; q[kk * 3 + r] = u; }
          }
        }
      }
    }
  }
  for (int i = 0; i < 3; i++) out.evals[i] = diag[i] * scale;
  for (int i = 0; i < 9; i++) out.evecs[i] = q[i];
  out.ok = ok;
  return out;
}

// ---- orthogonal_LSQ of an ordered cloud (wrap.h:121-150) by one CTA --------------------------
struct LsqOut {
  double a[3], b[3], rc;
};
struct LsqShared {
  float mean[3];
  float sc[6];   // xx xy xz yy yz zz
};

__device__ __forceinline__ float tree32_combine(float acc) {
#pragma unroll
  for (int o = 16; o; o >>= 1) acc = __fadd_rn(acc, __shfl_down_sync(0xFFFFFFFFu, acc, o));
  return acc;   // lane 0 holds the canonical sum
}

// meanValue of axis values get(k) (k = 0..n-1 in cloud order) by one warp: the reference's
// sequential double sum, via the exact int64 route when it is provably identical
// U: batches of 32 values fetched together (U loads in flight per lane).  1 when get() reads shared memory; 8 when it
// reads global memory: with one batch of look-ahead every batch of the chain waited for an L2 round trip (~0.35 us),
// 202 us per call for the 11 k inliers of a configs[3] line (ncu launch list, round 2).
template <int U = 1, class GetAxis>
__device__ __forceinline__ double mean_axis_warp(GetAxis get, int n, int lane) {
  int lo = 0x7FFFFFFF, hi = -0x7FFFFFFF;
  bool bad = false;
  const int nbits = 32 - __clz(n);
  // The exact-int64 route needs hi + 1 - lo + bits(n) <= 53 over ALL values; lo only falls and hi only rises, so
  // the scan stops at the first group of 32*U values that already rules the route out (doubles with full
  // mantissas do so at once: the scan was a quarter of the kernel's time for nothing).
  for (int k0 = 0; k0 < n && !bad; k0 += 32 * U) {
    double v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) { const int k = k0 + u * 32 + lane; v[u] = k < n ? get(k) : 0.0; }   // loads together
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long bits = __double_as_longlong(v[u]);
      const int e = (int)((bits >> 52) & 0x7FF);
      const long long mant = bits & 0xFFFFFFFFFFFFFLL;
      if (e == 0x7FF || (e == 0 && mant != 0)) bad = true;          // Inf/NaN/subnormal: sequential path
      else if (e != 0) {                                            // (+-0, also the padding of the last group: no constraint)
        lo = min(lo, e - 1075 + (__ffsll(mant | 0x10000000000000LL) - 1));
        hi = max(hi, e - 1023);
      }
    }
    const int glo = __reduce_min_sync(0xFFFFFFFFu, lo), ghi = __reduce_max_sync(0xFFFFFFFFu, hi);
    bad = __any_sync(0xFFFFFFFFu, bad) || (glo != 0x7FFFFFFF && !(ghi + 1 - glo + nbits <= 53 && glo > -900 && ghi < 900));
  }
  lo = __reduce_min_sync(0xFFFFFFFFu, lo);
  hi = __reduce_max_sync(0xFFFFFFFFu, hi);
  double sum;
  if (lo == 0x7FFFFFFF && !bad) {
    sum = 0.0;                                                     // all zero (or n == 0)
  } else if (!bad && hi + 1 - lo + nbits <= 53 && lo > -900 && hi < 900) {
    const double scale = __longlong_as_double((long long)(1023 - lo) << 52);   // 2^-lo
    long long acc = 0;
#pragma unroll U
    for (int k = lane; k < n; k += 32) acc += __double2ll_rn(__dmul_rn(get(k), scale));
#pragma unroll
    for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xFFFFFFFFu, acc, o);
    sum = __dmul_rn((double)acc, __longlong_as_double((long long)(1023 + lo) << 52));
  } else {
    // the sequential chain itself, with the loads taken off it: the warp fetches 32 values at a time
    // (the next batch is in flight while this one is added) and every lane adds them in order from
    // the shuffles, so only the 32 dependent DADDs remain per batch (a lone lane reading get(k)
    // paid two dependent L2 round trips per element: 2 ms for the 11 k inliers of a configs[3] line)
    //
    // Most batches do not even need the chain.  While the running sum S stays inside one binade
    // [2^e, 2^(e+1)), every partial sum is a multiple of u = 2^(e-52), and fl(S + x) = S + rne_u(x) whenever
    // x/u is not exactly half-way between two integers (adding the integer S/u does not move a non-tie
    // rounding).  So the warp rounds its 32 values to multiples of u, prefix-sums them as int64 on top of
    // S/u, and accepts the batch iff no value was a tie and every partial sum P_j stays STRICTLY inside
    // (2^52, 2^53) in magnitude -- then every intermediate result of the sequential chain had spacing u and
    // S + sum is the chain's result bit for bit.  Otherwise (binade crossings, ~log2(n) of them, the first
    // batches, sums near zero) this batch runs the chain.  ~45 instructions instead of 32 dependent DADDs.
    sum = 0.0;
    double buf[U];
#pragma unroll
    for (int u = 0; u < U; ++u) { const int k = u * 32 + lane; buf[u] = k < n ? get(k) : 0.0; }
    for (int k0 = 0; k0 < n; k0 += 32 * U) {
      double nbuf[U];
#pragma unroll
      for (int u = 0; u < U; ++u) { const int k = k0 + (U + u) * 32 + lane; nbuf[u] = k < n ? get(k) : 0.0; }
      // U batches at once under the same argument: if all 32*U partial sums stay strictly inside the binade S is
      // in when the group starts (and no value is a tie), the whole chain through the group had spacing u.  The U
      // warp scans are independent (one batch at a time, each batch waited for the previous one's S: ~400 cycles of
      // dependent conversions, shuffles and a vote per batch, 70 us for 11 k values).
      bool group_done = false;
      if (U > 1) {
        const int es = (int)((__double_as_longlong(sum) >> 52) & 0x7FF);
        if (es > 100 && es < 1900) {
          const double inv_u = __longlong_as_double((long long)(1023 + 52 + 1023 - es) << 52);
          long long pj[U];
          bool ok = true;
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const double q = __dmul_rn(buf[u], inv_u), r = rint(q);
            ok = ok && fabs(q) < 1152921504606846976.0 && fabs(__dsub_rn(q, r)) != 0.5;
            pj[u] = (long long)r;
          }
#pragma unroll
          for (int o = 1; o < 32; o <<= 1) {
#pragma unroll
            for (int u = 0; u < U; ++u) {
              const long long up = __shfl_up_sync(0xFFFFFFFFu, pj[u], o);
              if (lane >= o) pj[u] += up;
            }
          }
          long long tot[U];
#pragma unroll
          for (int u = 0; u < U; ++u) tot[u] = __shfl_sync(0xFFFFFFFFu, pj[u], 31);
          long long base = (long long)__dmul_rn(sum, inv_u);                   // S/u: an integer in (2^52, 2^53)
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const long long v = pj[u] + base, mag = v < 0 ? -v : v;
            ok = ok && mag > 4503599627370496LL && mag < 9007199254740992LL;
            base += tot[u];
          }
          if (__all_sync(0xFFFFFFFFu, ok)) {
            sum = __dmul_rn((double)base, __longlong_as_double((long long)(es - 52) << 52));   // P * u, exact
            group_done = true;
          }
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int kb = k0 + u * 32;
        if (group_done || kb >= n) break;                                      // uniform
        const double cur = buf[u];
        const int m = n - kb < 32 ? n - kb : 32;
        bool done = false;
        const int es = (int)((__double_as_longlong(sum) >> 52) & 0x7FF);      // biased exponent of S (uniform)
        if (es > 100 && es < 1900) {
          const double inv_u = __longlong_as_double((long long)(1023 + 52 + 1023 - es) << 52);   // 2^(52-e)
          const double q = __dmul_rn(cur, inv_u);                              // exact: a power of two, no underflow
          const double r = rint(q);                                            // lanes >= m hold 0
          // |q| < 2^60 (int64 range; also rejects Inf/NaN), not a tie (q - r is exact).  A q that underflows
          // is far below 1/2 and rounds to 0, as the chain's addition does.
          const bool lane_ok = fabs(q) < 1152921504606846976.0 && fabs(__dsub_rn(q, r)) != 0.5;
          long long pj = (long long)r;
#pragma unroll
          for (int o = 1; o < 32; o <<= 1) {
            const long long up = __shfl_up_sync(0xFFFFFFFFu, pj, o);
            if (lane >= o) pj += up;
          }
          pj += (long long)__dmul_rn(sum, inv_u);                              // S/u: an integer in (2^52, 2^53)
          const long long mag = pj < 0 ? -pj : pj;
          const bool inside = mag > 4503599627370496LL && mag < 9007199254740992LL;
          if (__all_sync(0xFFFFFFFFu, lane_ok && inside)) {
            const long long last = __shfl_sync(0xFFFFFFFFu, pj, 31);
            sum = __dmul_rn((double)last, __longlong_as_double((long long)(es - 52) << 52));   // P * u, exact
            done = true;
          }
        }
        if (!done) {
#pragma unroll 4
          for (int j = 0; j < m; ++j) sum = __dadd_rn(sum, __shfl_sync(0xFFFFFFFFu, cur, j));
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) buf[u] = nbuf[u];
    }
  }
  return n > 0 ? __ddiv_rn(sum, (double)n) : sum;
}

// x, y, z: the cloud's coordinate arrays (n values each, global memory).  blockDim.x a multiple of 32; all threads
// of the CTA call it; ends with a barrier (out and s are shared memory).
// The axis of a task is picked ONCE, outside the loops: with an accessor get(k, c) and a run-time c the compiler
// turned the choice of the array into branches inside the loop body, the loads of an unrolled body could no
// longer be issued together, and every element paid its own L2 round trip (ncu SASS, round 2: the scatter sums
// took 124 us for 11 k points, 700 cycles per element).
// stage / cap: shared memory for 3 * cap floats (nullptr / 0: none).  The warps that sum the float columns keep
// the converted values of the first `cap` points there, and the six scatter sums read them back instead of
// converting every coordinate four more times inside their dependent chains (F2F.F32.F64 on inexact values
// made those chains 3x slower than on float-exact data: 78 vs 38 us for 11 k points, tools/lsq_bench.py).
constexpr int kLsqStageCap = 16384;
__device__ __forceinline__ void lsq_cta(const double *__restrict__ x, const double *__restrict__ y,
                                        const double *__restrict__ z, int n, LsqShared &s, LsqOut *out,
                                        unsigned long long *dbg = nullptr, float *stage = nullptr, int cap = 0) {
  auto dbg_stamp = [&](int k) {
    if (dbg && threadIdx.x == 0) { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t) : : "memory"); dbg[k] = t; }
  };
  auto axis = [&](int c) { return c == 0 ? x : c == 1 ? y : z; };
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarps = blockDim.x >> 5;
  for (int q = warp; q < 6; q += nwarps) {
    if (q < 3) {                      // colwise().mean() of the float copy (wrap.h:129-137)
      const double *__restrict__ p = axis(q);
      const int ns = n < cap ? n : cap;          // staged part (a multiple of 32 or all of n: lane order is kept)
      float *__restrict__ sq = stage + (size_t)q * cap;
      float acc = 0.0f;
#pragma unroll 8
      for (int k = lane; k < ns; k += 32) { const float v = (float)p[k]; sq[k] = v; acc = __fadd_rn(acc, v); }
#pragma unroll 8
      for (int k = ns + lane; k < n; k += 32) acc = __fadd_rn(acc, (float)p[k]);   // loads run ahead of the chain
      acc = tree32_combine(acc);
      if (lane == 0) s.mean[q] = __fdiv_rn(acc, (float)n);
      if (dbg && q == 0 && lane == 0) { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t) : : "memory"); dbg[5] = t; }
    } else {                          // meanValue (lib/pointcloud.cpp:29-38)
      const double *__restrict__ p = axis(q - 3);
      const double m = mean_axis_warp<8>([&](int k) { return p[k]; }, n, lane);
      if (lane == 0) out->a[q - 3] = m;
      if (dbg && q == 3 && lane == 0) { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t) : : "memory"); dbg[4] = t; }
    }
  }
  __syncthreads();
  dbg_stamp(0);
  for (int q = warp; q < 6; q += nwarps) {   // centred^T * centred (wrap.h:137-138)
    const int c0 = q < 3 ? 0 : q < 5 ? 1 : 2, c1 = q < 3 ? q : q < 5 ? q - 2 : 2;
    const double *__restrict__ p0 = axis(c0), *__restrict__ p1 = axis(c1);
    const float m0 = s.mean[c0], m1 = s.mean[c1];
    const int ns = n < cap ? n : cap;
    const float *__restrict__ s0 = stage + (size_t)c0 * cap, *__restrict__ s1 = stage + (size_t)c1 * cap;
    float acc = 0.0f;
#pragma unroll 8
    for (int k = lane; k < ns; k += 32) acc = __fadd_rn(acc, __fmul_rn(__fsub_rn(s0[k], m0), __fsub_rn(s1[k], m1)));
#pragma unroll 8
    for (int k = ns + lane; k < n; k += 32)
      acc = __fadd_rn(acc, __fmul_rn(__fsub_rn((float)p0[k], m0), __fsub_rn((float)p1[k], m1)));
    acc = tree32_combine(acc);
    if (lane == 0) s.sc[q] = acc;
  }
  __syncthreads();
  dbg_stamp(1);
  if (tid == 0) {
    const float *sc = s.sc;
    const float m[9] = {sc[0], sc[1], sc[2], sc[1], sc[3], sc[4], sc[2], sc[4], sc[5]};
    const Eig3f e = eig3f(m);
    out->b[0] = e.evecs[6]; out->b[1] = e.evecs[7]; out->b[2] = e.evecs[8];   // last column, wrap.h:146
    out->rc = (double)e.evals[2];
  }
  dbg_stamp(2);
  __syncthreads();
}

}  // namespace h3d
#endif  // H3D_FIT_CUH_
