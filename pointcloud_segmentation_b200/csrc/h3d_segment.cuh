// h3d_segment.cuh -- segment post-tests of every accepted line, on the device.
//
// Reference: wrap.h:261-337, the block of hough3dlines() that runs on the
// final inlier set Y of a line right after the second orthogonal_LSQ: find_proj (wrap.h:78-85),
// find_t with its sorted insertion (wrap.h:41-68), largest gap between neighbours (wrap.h:286-292),
// radius match (wrap.h:295-309), segPCA (wrap.h:94-110), density test (wrap.h:320-322).
// ("wrap.h" = include/pointcloud_segmentation/hough_3d_lines.h of the reference.)
//
// The reference copies Y.points into Eigen vectors line by line on the host; here the inlier
// index lists of the loop are still in device memory (the index pool), so one launch with one
// CTA per iteration record tests every line of the frame and only the 200-byte records go back.
//
// Arithmetic restated (same as oracle/h3d_oracle.c h3do_segment_test):
//   * Eigen fixed-size Vector3d: sums of three terms reduce as x0 + (x1 + x2); no FMA contraction
//     (this file is compiled with -fmad=false);
//   * the sorted insertion of find_t is a stable sort by t (equal t keep Y order): a bitonic sort of
//     (order-preserving 64-bit key of t, index) pairs in shared memory for up to 2048 inliers, else
//     rank(i) = #{j : t_j < t_i or (t_j == t_i and j < i)} by counting; only max/min-type results
//     depend on the order, no float sum does (NaN t -- a non-finite anchor -- is not reproduced);
//   * pcl::PCA<PointXYZ>: float centroid (sequential sum), float demeaned products summed in the
//     canonical TREE32 order, SelfAdjointEigenSolver<Matrix3f> (eig3f<true>).  PARITY UNPINNED
//     for this part (PCL and Eigen are not vendored in the reference).
#ifndef H3D_SEGMENT_CUH_
#define H3D_SEGMENT_CUH_

#include "../../include/h3d.h"
#include "h3d_fit.cuh"

namespace h3d {

constexpr int kSegThreads = 256;
constexpr unsigned kSegNone = 4;        // the iteration did not yield a line
constexpr unsigned kSegTruncated = 5;   // the inlier list did not fit the index pool
constexpr unsigned kSegSortCap = 2048;  // inliers per line sorted in shared memory

struct SegArgs {
  const h3d_iter *iters;     // iteration records of the frame (device)
  const int *n_iters;        // records written by the loop (device; nullptr: every CTA has one)
  const unsigned *pool;      // index pool (device); nullptr: the list is 0..n2-1
  unsigned pool_cap;
  const void *src;           // the cloud the pool indexes: f32 or f64, `stride` elements apart
  int stride, src_is_float;
  double ns[3];              // what shiftToOrigin subtracted (lib/pointcloud.cpp:18-26)
  double shift[3];           // X.shift
  const double *ns_dev;      // not null: both of the above are read from here (3 doubles in device memory,
                             // FrameDev::ns of the frame fast path: the host never sees the bounding box)
  h3d_segment_config cfg;
  h3d_pose pose;             // drone -> world (node.cpp:429-470)
  int pose_on;
  double *scr;               // 5 * scr_n doubles: t, |a + t b|, sorted |a + t b|, then 3 * scr_n floats (points)
  size_t scr_n;
  h3d_segment *out;
};

__device__ __forceinline__ double seg_dot3(double u0, double u1, double u2, double v0, double v1, double v2) {
  return __dadd_rn(__dmul_rn(u0, v0), __dadd_rn(__dmul_rn(u1, v1), __dmul_rn(u2, v2)));
}

// point k of the source cloud in the frame of the reported line: (p - ns) + shift, wrap.h:271
__device__ __forceinline__ void seg_point(const SegArgs &A, unsigned k, double p[3]) {
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const size_t e = (size_t)k * A.stride + c;
    const double v = A.src_is_float ? (double)static_cast<const float *>(A.src)[e] : static_cast<const double *>(A.src)[e];
    const double nsc = A.ns_dev ? A.ns_dev[c] : A.ns[c], shc = A.ns_dev ? __dadd_rn(0.0, A.ns_dev[c]) : A.shift[c];
    p[c] = __dadd_rn(__dsub_rn(v, nsc), shc);
  }
}

// drone2WorldSeg + heighSegmentCutoff (node.cpp:429-470) on one kept segment.
// R = Eigen::Quaterniond::toRotationMatrix(); R*v per row as x0 + (x1 + x2), then + position.
__device__ __forceinline__ void seg_to_world(const h3d_pose &P, h3d_segment &s) {
  const double w = P.q[0], x = P.q[1], y = P.q[2], z = P.q[3];
  const double tx = __dmul_rn(2.0, x), ty = __dmul_rn(2.0, y), tz = __dmul_rn(2.0, z);
  const double twx = __dmul_rn(tx, w), twy = __dmul_rn(ty, w), twz = __dmul_rn(tz, w);
  const double txx = __dmul_rn(tx, x), txy = __dmul_rn(ty, x), txz = __dmul_rn(tz, x);
  const double tyy = __dmul_rn(ty, y), tyz = __dmul_rn(tz, y), tzz = __dmul_rn(tz, z);
  const double R[3][3] = {{__dsub_rn(1.0, __dadd_rn(tyy, tzz)), __dsub_rn(txy, twz), __dadd_rn(txz, twy)},
                          {__dadd_rn(txy, twz), __dsub_rn(1.0, __dadd_rn(txx, tzz)), __dsub_rn(tyz, twx)},
                          {__dsub_rn(txz, twy), __dadd_rn(tyz, twx), __dsub_rn(1.0, __dadd_rn(txx, tyy))}};
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    s.a_world[i] = __dadd_rn(seg_dot3(R[i][0], R[i][1], R[i][2], s.a[0], s.a[1], s.a[2]), P.position[i]);
    s.b_world[i] = seg_dot3(R[i][0], R[i][1], R[i][2], s.b[0], s.b[1], s.b[2]);
  }
  const double z1 = __dadd_rn(__dmul_rn(s.t_min, s.b_world[2]), s.a_world[2]);
  const double z2 = __dadd_rn(__dmul_rn(s.t_max, s.b_world[2]), s.a_world[2]);
  s.world_status = (z1 > P.floor_trim_height || z2 > P.floor_trim_height) ? 1u : 2u;
}

__device__ __forceinline__ double seg_block_max(double v, double *red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int o = 16; o; o >>= 1) v = fmax(v, __shfl_xor_sync(0xFFFFFFFFu, v, o));
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  double m = red[0];
  for (int w = 1; w < kSegThreads / 32; ++w) m = fmax(m, red[w]);
  return m;
}

__global__ void __launch_bounds__(kSegThreads) segment_tests_kernel(const SegArgs A) {
  __shared__ double red[kSegThreads / 32];
  __shared__ double sh_t[2], sh_r[2];
  __shared__ float sh_cen[3], sh_sc[6];
  __shared__ unsigned long long sh_key[kSegSortCap];
  __shared__ unsigned sh_idx[kSegSortCap];
  const int it = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const h3d_iter &R = A.iters[it];
  h3d_segment *S = &A.out[it];
  const unsigned n = R.n2;
  unsigned status = 1;
  if ((A.n_iters && it >= *A.n_iters) || R.status != 0 || n == 0) status = kSegNone;
  else if (A.pool && (size_t)R.off2 + n > A.pool_cap) status = kSegTruncated;
  // the final inlier sets of the accepted lines are disjoint (removePoints, wrap.h:339), so
  // their sizes sum to at most the cloud: scratch base = sizes of the earlier lines
  size_t base = 0;
  if (status == 1)
    for (int k = 0; k < it; ++k) if (A.iters[k].status == 0) base += A.iters[k].n2;
  if (status == 1 && base + n > A.scr_n) status = kSegTruncated;
  double a[3], b[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) { a[c] = R.a2[c]; b[c] = R.b2[c]; }
  if (status == 1 && b[0] == 0.0) status = 3;                       // find_t: no solution, wrap.h:43-45
  if (status != 1) {
    if (tid == 0) {
      h3d_segment s;
      memset(&s, 0, sizeof s);
      s.iter = (uint32_t)it; s.status = status; s.n_points = status == kSegNone ? 0u : n;
      for (int c = 0; c < 3; ++c) { s.a[c] = a[c]; s.b[c] = b[c]; }
      *S = s;
    }
    return;
  }
  const unsigned *list = A.pool ? A.pool + R.off2 : nullptr;
  double *t_arr = A.scr + base, *n_arr = A.scr + A.scr_n + base, *s_arr = A.scr + 2 * A.scr_n + base;
  float *pf = reinterpret_cast<float *>(A.scr + 3 * A.scr_n) + 3 * base;          // pcl::PointXYZ(x, y, z), wrap.h:99
  unsigned np2 = 1;
  while (np2 < n) np2 <<= 1;
  const bool in_smem = np2 <= kSegSortCap;
  // ---- per point: projection, radius, t, |a + t b| ---------------------------- wrap.h:270-283
  double d[3];
#pragma unroll
  for (int c = 0; c < 3; ++c) d[c] = __dsub_rn(__dadd_rn(a[c], b[c]), a[c]);      // p_B - p_A
  const double dd = seg_dot3(d[0], d[1], d[2], d[0], d[1], d[2]);
  for (unsigned i = tid; i < n; i += kSegThreads) {
    double p[3], proj[3], e[3], q[3];
    seg_point(A, list ? list[i] : i, p);
    const double num = seg_dot3(__dsub_rn(p[0], a[0]), __dsub_rn(p[1], a[1]), __dsub_rn(p[2], a[2]), d[0], d[1], d[2]);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      proj[c] = __dadd_rn(a[c], __ddiv_rn(__dmul_rn(d[c], num), dd));
      e[c] = __dsub_rn(proj[c], p[c]);
    }
    if (i == 0 || i == n - 1) {
      const double r = __dsqrt_rn(seg_dot3(e[0], e[1], e[2], e[0], e[1], e[2]));
      if (i == 0) sh_r[0] = r;
      if (i == n - 1) sh_r[1] = r;
    }
    const double t = __ddiv_rn(__dsub_rn(proj[0], a[0]), b[0]);
#pragma unroll
    for (int c = 0; c < 3; ++c) q[c] = __dadd_rn(a[c], __dmul_rn(t, b[c]));
    t_arr[i] = t;
    n_arr[i] = __dsqrt_rn(seg_dot3(q[0], q[1], q[2], q[0], q[1], q[2]));
    pf[3 * i] = (float)p[0]; pf[3 * i + 1] = (float)p[1]; pf[3 * i + 2] = (float)p[2];
    if (in_smem) {
      const long long bits = __double_as_longlong(__dadd_rn(t, 0.0));             // -0.0 == +0.0 for upper_bound
      sh_key[i] = (unsigned long long)(bits < 0 ? ~bits : (bits | (long long)0x8000000000000000ull));
      sh_idx[i] = i;
    }
  }
  if (in_smem)
    for (unsigned i = n + tid; i < np2; i += kSegThreads) { sh_key[i] = ~0ull; sh_idx[i] = 0xFFFFFFFFu; }
  __syncthreads();
  // ---- find_t's upper_bound insertion = stable sort by t; norms follow ------------------------
  if (in_smem) {
    for (unsigned k = 2; k <= np2; k <<= 1)
      for (unsigned j = k >> 1; j > 0; j >>= 1) {
        for (unsigned i = tid; i < np2; i += kSegThreads) {
          const unsigned l = i ^ j;
          if (l > i) {
            const unsigned long long ka = sh_key[i], kb = sh_key[l];
            const unsigned ia = sh_idx[i], ib = sh_idx[l];
            const bool gt = (ka > kb) | ((ka == kb) & (ia > ib));
            if (gt == ((i & k) == 0)) { sh_key[i] = kb; sh_key[l] = ka; sh_idx[i] = ib; sh_idx[l] = ia; }
          }
        }
        __syncthreads();
      }
    for (unsigned k = tid; k < n; k += kSegThreads) s_arr[k] = n_arr[sh_idx[k]];
    if (tid == 0) { sh_t[0] = t_arr[sh_idx[0]]; sh_t[1] = t_arr[sh_idx[n - 1]]; }
  } else {
    for (unsigned i = tid; i < n; i += kSegThreads) {
      const double ti = t_arr[i];
      unsigned rank = 0;
      for (unsigned j = 0; j < n; ++j) {
        const double tj = t_arr[j];
        rank += (tj < ti) | ((tj == ti) & (j < i));
      }
      s_arr[rank] = n_arr[i];
      if (rank == 0) sh_t[0] = ti;
      if (rank == n - 1) sh_t[1] = ti;
    }
  }
  __syncthreads();
  // ---- largest gap between neighbours on the line, wrap.h:286-292 ----------------------------
  double gap = 0.0;
  for (unsigned k = tid; k + 1 < n; k += kSegThreads) gap = fmax(gap, fabs(__dsub_rn(s_arr[k + 1], s_arr[k])));
  const double max_dist = seg_block_max(gap, red);
  // ---- radius match, wrap.h:295-312 ----------------------------------------------------------
  const double radius = fmax(sh_r[0], sh_r[1]);
  double closest = A.cfg.radius_sizes[0];
  double min_diff = fabs(__dsub_rn(radius, A.cfg.radius_sizes[0]));
  double max_radius = min_diff;                                                   // sic, wrap.h:298
  for (unsigned k = 0; k < A.cfg.n_radius; ++k) {
    const double r = A.cfg.radius_sizes[k];
    const double cur = fabs(__dsub_rn(radius, r));
    if (cur < min_diff) { min_diff = cur; closest = r; }
    if (r > max_radius) max_radius = r;
  }
  const bool pass = min_diff < A.cfg.diag_voxel && max_radius <= closest && max_dist < __dmul_rn(2.0, A.cfg.diag_voxel);
  h3d_segment s;
  if (tid == 0) {
    memset(&s, 0, sizeof s);
    s.iter = (uint32_t)it; s.n_points = n;
    for (int c = 0; c < 3; ++c) { s.a[c] = a[c]; s.b[c] = b[c]; }
    s.t_min = sh_t[0]; s.t_max = sh_t[1];
    s.radius_raw = radius; s.radius = closest; s.max_dist = max_dist;
    s.status = 1;
  }
  if (!pass) {
    if (tid == 0) *S = s;
    return;
  }
  // ---- segPCA (wrap.h:94-110): pcl::PCA in float ---------------------------------------------
  if (warp == 0 && lane < 3) {            // compute3DCentroid: sequential float sum per axis
    float acc = 0.0f;
    unsigned i = 0;
    for (; i + 8 <= n; i += 8) {            // loads up front, the adds stay in cloud order
      float v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = pf[3 * (i + u) + lane];
#pragma unroll
      for (int u = 0; u < 8; ++u) acc = __fadd_rn(acc, v[u]);
    }
    for (; i < n; ++i) acc = __fadd_rn(acc, pf[3 * i + lane]);
    sh_cen[lane] = __fdiv_rn(acc, (float)n);
  }
  __syncthreads();
  if (warp < 6) {                         // one warp per product: xx xy xz yy yz zz, TREE32
    const int u = warp < 3 ? 0 : warp < 5 ? 1 : 2;
    const int v = warp < 3 ? warp : warp < 5 ? warp - 2 : 2;
    float acc = 0.0f;
    for (unsigned i = lane; i < n; i += 32) {
      const float cu = __fsub_rn(pf[3 * i + u], sh_cen[u]), cv = __fsub_rn(pf[3 * i + v], sh_cen[v]);
      acc = __fadd_rn(acc, __fmul_rn(cu, cv));
    }
    acc = tree32_combine(acc);
    if (lane == 0) sh_sc[warp] = __fmul_rn(__fdiv_rn(1.0f, __fsub_rn((float)n, 1.0f)), acc);
  }
  __syncthreads();
  if (tid == 0) {
    const float m[9] = {sh_sc[0], sh_sc[1], sh_sc[2], sh_sc[1], sh_sc[3], sh_sc[4], sh_sc[2], sh_sc[4], sh_sc[5]};
    bool finite = true;
    for (int c = 0; c < 6; ++c) finite &= isfinite(sh_sc[c]);
    float ev[3] = {NAN, NAN, NAN};
    if (finite) {
      const Eig3f e = eig3f<true>(m);
      ev[0] = e.evals[0]; ev[1] = e.evals[1]; ev[2] = e.evals[2];
    }
    const double e0 = (double)ev[2], e1 = (double)ev[1], e2 = (double)ev[0];      // descending
    s.pca_eigenvalues[0] = e0; s.pca_eigenvalues[1] = e1; s.pca_eigenvalues[2] = e2;
    s.pca_coeff = __ddiv_rn(e0, __dadd_rn(__dadd_rn(e0, e1), e2));                // wrap.h:315
    double dl[3];
    for (int c = 0; c < 3; ++c)                                                   // wrap.h:317-319
      dl[c] = __dsub_rn(__dadd_rn(__dmul_rn(s.t_max, b[c]), a[c]), __dadd_rn(__dmul_rn(s.t_min, b[c]), a[c]));
    s.length = __dsqrt_rn(seg_dot3(dl[0], dl[1], dl[2], dl[0], dl[1], dl[2]));
    const double den = __dmul_rn(__dmul_rn(__dmul_rn(__dmul_rn(A.cfg.rad_2_leaf_ratio, 2.0), A.cfg.diag_voxel), 2.0), A.cfg.diag_voxel);
    const double want = __ddiv_rn(__dmul_rn(__dmul_rn(2.0, closest), s.length), den);   // wrap.h:320
    // static_cast<int> as x86 converts it: out of range or NaN -> INT_MIN
    const int min_nb = (want >= 2147483648.0 || want <= -2147483649.0 || want != want) ? (int)0x80000000 : __double2int_rz(want);
    s.min_nb_points = min_nb;
    // points.size() > min_nb_points_seg: size_t against int, the int converts to unsigned
    const bool keep = s.pca_coeff > A.cfg.min_pca_coeff && (unsigned long long)n > (unsigned long long)(long long)min_nb;
    s.status = keep ? 0u : 2u;
    if (keep && A.pose_on) seg_to_world(A.pose, s);
    *S = s;
  }
}

}  // namespace h3d
#endif  // H3D_SEGMENT_CUH_
