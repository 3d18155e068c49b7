// h3d_device.cuh -- device-side data model and arithmetic helpers (sm_100a).
//
// Accumulator model.  The reference keeps a dense unsigned array
// VotingSpace[x_i][y_i][j] (lib/hough.cpp:41, 88-90).  A vote of point p for
// direction j only ever touches plane j, and getLine only needs the first
// maximum, so on the B200 a plane (num_x^2 u32 cells, <= 24 KB for the ToF
// configuration) is built in SHARED memory by one CTA, reduced to
// (max votes, first max cell) and dropped.  HBM holds, per direction,
//   dvotes[j]  the plane maximum (exact when dstale[j]==0, an upper bound else)
//   dcell[j]   the plane cell x_i*num_x + y_i of its first maximum
//   dstale[j]  1 when a removed point voted for dcell[j] since the last build
// Removing points can only lower counts, so a stale maximum is an upper bound
// and only planes whose bound reaches the best exact maximum are rebuilt
// (h3d_kernels.cuh).  The flat reference index is dcell*num_b + j.
#ifndef H3D_DEVICE_CUH_
#define H3D_DEVICE_CUH_

#include <cuda_runtime.h>
#include <stdint.h>

#include "h3d_internal.h"

namespace h3d {

struct HoughDev {
  // discretisation (lib/hough.cpp:32-40)
  int num_x;              // cells per axis
  unsigned plane_cells;   // num_x * num_x
  unsigned plane_stride;  // plane_cells rounded up to a multiple of 4 (uint4 scan)
  int num_b;              // directions of the whole space (flat-index multiplier)
  int dir_begin;          // first direction owned by this instance
  int dir_count;          // directions owned
  double max_x, dx;
  // fast cell path: n = round((x' + max_x)/dx * 2^20 + 2^19) read from the low
  // mantissa word of fma(x', inv_s, kmagic)
  double inv_s, kmagic;
  int fast_ok;
  const DirCoef *coef;    // whole space, indexed by global direction
  // points (SoA, centred frame) and their weights: +1 added, -1 subtracted
  // without a matching added point, 0 removed
  const double *px, *py, *pz;
  const int *w;
  int n;
  // per-direction results, indexed by LOCAL direction l = j - dir_begin
  unsigned *dvotes, *dcell;
  unsigned char *dstale;
};

// ---- exact reference arithmetic (no FMA anywhere: the reference is built
// without -march / -ffp-contract, lib/Makefile:19, CMakeLists.txt:5) -----------

__device__ __forceinline__ double ref_x_new(const DirCoef &c, double px, double py, double pz) {
  return __dsub_rn(__dsub_rn(__dmul_rn(c.c0, px), __dmul_rn(c.c1, py)), __dmul_rn(c.bx, pz));
}
__device__ __forceinline__ double ref_y_new(const DirCoef &c, double px, double py, double pz) {
  return __dsub_rn(__dsub_rn(__dmul_rn(c.c2, py), __dmul_rn(c.c1, px)), __dmul_rn(c.by, pz));
}

// roundToNearest((v + max_x)/dx) of lib/hough.cpp:22-24,86-87 evaluated exactly,
// as a signed integer.  Values beyond +-2^40 (and NaN) return 2^40: such a vote
// can never land inside the accumulator.
__device__ __noinline__ long long ref_cell_slow(double v, double max_x, double dx) {
  const double q = __ddiv_rn(__dadd_rn(v, max_x), dx);
  const double r = (q > 0.0) ? floor(__dadd_rn(q, 0.5)) : ceil(__dsub_rn(q, 0.5));
  if (!(fabs(r) < 1099511627776.0)) return 1099511627776LL;
  return (long long)r;
}

// Plane cell x_i*num_x + y_i of one vote, or 0xFFFFFFFF when the reference's
// flat-index test `index < VotingSpace.size()` (lib/hough.cpp:92) drops it.
// Only the FLAT bound is tested, so y_i == num_x aliases into the next row
// exactly as the reference does.
__device__ __forceinline__ unsigned vote_cell(const HoughDev &h, const DirCoef &c, double px, double py,
                                              double pz) {
  const double xn = ref_x_new(c, px, py, pz);
  const double yn = ref_y_new(c, px, py, pz);
  if (h.fast_ok) {
    // One DFMA per axis.  sx = 1.5*2^52 + n with n = round(t*2^20), t = (x'+max_x)/dx + 0.5,
    // |error| < 2 units of 2^-20.  If the 20 fraction bits of n are >= 4 units away
    // from a cell boundary, floor(t) is certain and equals the reference's
    // floor(fl(fl(x'+max_x)/dx) + 0.5) (its rounding slack is ~1e-12 cells).
    const double sx = __fma_rn(xn, h.inv_s, h.kmagic);
    const double sy = __fma_rn(yn, h.inv_s, h.kmagic);
    const unsigned lx = (unsigned)__double2loint(sx), ly = (unsigned)__double2loint(sy);
    const bool sure = (__double2hiint(sx) == 0x43380000) & (__double2hiint(sy) == 0x43380000) &
                      (((lx & 0xFFFFFu) - 4u) < 0xFFFF8u) & (((ly & 0xFFFFFu) - 4u) < 0xFFFF8u);
    if (sure) {
      const unsigned cell = (lx >> 20) * (unsigned)h.num_x + (ly >> 20);
      return cell < h.plane_cells ? cell : 0xFFFFFFFFu;
    }
  }
  const long long xi = ref_cell_slow(xn, h.max_x, h.dx);
  const long long yi = ref_cell_slow(yn, h.max_x, h.dx);
  const long long cell = xi * (long long)h.num_x + yi;   // |.| < 2^40 * 2^31: no overflow
  return (cell >= 0 && cell < (long long)h.plane_cells) ? (unsigned)cell : 0xFFFFFFFFu;
}

// packed argmax key: larger votes first, then the LOWEST flat index
// (lib/hough.cpp:107-112 strict `>` scan from index 0)
__device__ __forceinline__ unsigned long long line_key(unsigned votes, unsigned cell, int j, int num_b) {
  const unsigned flat = cell * (unsigned)num_b + (unsigned)j;
  return ((unsigned long long)votes << 32) | (unsigned long long)(0xFFFFFFFFu - flat);
}

__device__ __forceinline__ unsigned long long warp_max_u64(unsigned long long v) {
  const unsigned hi = __reduce_max_sync(0xFFFFFFFFu, (unsigned)(v >> 32));
  const unsigned lo = __reduce_max_sync(0xFFFFFFFFu, (unsigned)(v >> 32) == hi ? (unsigned)v : 0u);
  return ((unsigned long long)hi << 32) | lo;
}

}  // namespace h3d
#endif  // H3D_DEVICE_CUH_
