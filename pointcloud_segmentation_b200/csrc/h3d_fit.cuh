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
// canonical order fixed by the oracle (sequential float sums in point order, Eigen's
// SelfAdjointEigenSolver<MatrixXf> = scale, Householder tridiagonalisation, implicit
// symmetric QR with Wilkinson shift, ascending sort).  Eigen is not vendored by the
// reference: PARITY UNPINNED against Eigen itself, bit-exact against oracle/h3d_oracle.c.
//
// Sequential sums are inherently serial; the kernels keep them exact but cheap: a warp
// stages 32 points at a time (loads, float casts, centring and products run 32-wide), then
// one lane per accumulator adds the 32 staged values in order, so the serial chain costs one
// shared-memory read and one add per point.
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

__host__ __device__ inline void eig_givens(float p, float q, float *c, float *s) {   // JacobiRotation::makeGivens
  if (q == 0.0f) { *c = p < 0.0f ? -1.0f : 1.0f; *s = 0.0f; return; }
  if (p == 0.0f) { *c = 0.0f; *s = q < 0.0f ? 1.0f : -1.0f; return; }
  if (fabsf(p) > fabsf(q)) {
    const float t = q / p;
    float u = sqrtf(1.0f + t * t);
    if (p < 0.0f) u = -u;
    *c = 1.0f / u;
    *s = -t * *c;
  } else {
    const float t = p / q;
    float u = sqrtf(1.0f + t * t);
    if (q < 0.0f) u = -u;
    *s = -1.0f / u;
    *c = -t * *s;
  }
}

__host__ __device__ inline float eig_hypot(float x, float y) {   // numext::hypot
  const float ax = fabsf(x), ay = fabsf(y);
  const float big = ax > ay ? ax : ay;
  if (big == 0.0f) return 0.0f;
  const float small = ax > ay ? ay : ax;
  const float r = small / big;
  return big * sqrtf(1.0f + r * r);
}

__host__ __device__ inline Eig3f eig3f(const float m[9]) {
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
  a00 /= scale; a10 /= scale; a20 /= scale; a11 /= scale; a21 /= scale; a22 /= scale;

  // tridiagonalisation: one Householder reflection on column 0
  float diag[3], sub[2];
  float q[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  const float tail_sq = a20 * a20;
  if (tail_sq <= FLT_MIN) {
    diag[0] = a00; diag[1] = a11; diag[2] = a22; sub[0] = a10; sub[1] = a21;
  } else {
    float beta = sqrtf(a10 * a10 + tail_sq);
    if (a10 >= 0.0f) beta = -beta;
    const float v1 = a20 / (a10 - beta);
    const float tau = (beta - a10) / beta;
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

  // implicit symmetric QR iteration on the tridiagonal matrix
  const int n = 3;
  int end = n - 1, start = 0, iter = 0;
  const float precision = 2.0f * FLT_EPSILON, consider_zero = FLT_MIN;
  int ok = 1;
  while (end > 0) {
    for (int i = start; i < end; i++)
      if (fabsf(sub[i]) <= (fabsf(diag[i]) + fabsf(diag[i + 1])) * precision || fabsf(sub[i]) <= consider_zero)
        sub[i] = 0.0f;
    while (end > 0 && sub[end - 1] == 0.0f) end--;
    if (end <= 0) break;
    iter++;
    if (iter > 30 * n) { ok = 0; break; }
    start = end - 1;
    while (start > 0 && sub[start - 1] != 0.0f) start--;

    const float td = (diag[end - 1] - diag[end]) * 0.5f;
    const float e = sub[end - 1];
    float mu = diag[end];
    if (td == 0.0f) {
      mu -= fabsf(e);
    } else {
      const float e2 = e * e;
      const float hyp = eig_hypot(td, e);
      if (e2 == 0.0f) mu -= (e / (td + (td > 0.0f ? 1.0f : -1.0f))) * (e / hyp);
      else mu -= e2 / (td + (td > 0.0f ? hyp : -hyp));
    }
    float x = diag[start] - mu;
    float z = sub[start];
    for (int k = start; k < end; k++) {
      float c, s;
      eig_givens(x, z, &c, &s);
      const float sdk = s * diag[k] + c * sub[k];
      const float dkp1 = s * sub[k] + c * diag[k + 1];
      diag[k] = c * (c * diag[k] - s * sub[k]) - s * (c * sub[k] - s * diag[k + 1]);
      diag[k + 1] = s * sdk + c * dkp1;
      sub[k] = c * sdk - s * dkp1;
      if (k > start) sub[k - 1] = c * sub[k - 1] - s * z;
      x = sub[k];
      if (k < end - 1) { z = -s * sub[k + 1]; sub[k + 1] = c * sub[k + 1]; }
      for (int r = 0; r < 3; r++) {
        const float xi = q[k * 3 + r], yi = q[(k + 1) * 3 + r];
        q[k * 3 + r] = c * xi - s * yi;
        q[(k + 1) * 3 + r] = s * xi + c * yi;
      }
    }
  }
  if (ok) {   // ascending selection sort, eigenvector columns follow
    for (int i = 0; i < n - 1; i++) {
      int k = i;
      for (int j = i + 1; j < n; j++) if (diag[j] < diag[k]) k = j;
      if (k > i) {
        const float t = diag[i]; diag[i] = diag[k]; diag[k] = t;
        for (int r = 0; r < 3; r++) { const float u = q[i * 3 + r]; q[i * 3 + r] = q[k * 3 + r]; q[k * 3 + r] = u; }
      }
    }
  }
  for (int i = 0; i < 3; i++) out.evals[i] = diag[i] * scale;
  for (int i = 0; i < 9; i++) out.evecs[i] = q[i];
  out.ok = ok;
  return out;
}

}  // namespace h3d
#endif  // H3D_FIT_CUH_
