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
  unsigned plane_stride;  // plane_cells + 2*num_x + 4 slack cells, rounded up to a multiple of 4 (uint4 scan);
                          // the slack absorbs the out-of-range votes of the BOUNDED vote form
  int num_b;              // directions of the whole space (flat-index multiplier)
  int dir_begin;          // first direction owned by this instance
  int dir_count;          // directions owned
  double max_x, dx;
  // fast cell path: n = round((x' + max_x)/dx * 2^20 + 2^19) read from the low
  // mantissa word of fma(x', inv_s, kmagic)
  double inv_s, kmagic;
  int fast_ok;
  int bounded;            // box symmetric about the origin (shiftToOrigin) and num_x < 2000: with every point
                          // inside the box, 0 <= x_i, y_i <= num_x + 1 and the fast path needs no range test
  const DirCoef *coef;    // whole space, indexed by global direction
  // points (SoA, centred frame) and their weights: +1 added, -1 subtracted
  // without a matching added point, 0 removed
  const double *px, *py, *pz;
  const int *w;
  int n;                  // point count known to the host at launch ...
  const int *n_dev;       // ... or, when not null, the live count in device memory (loop)
  // per-direction results, indexed by LOCAL direction l = j - dir_begin
  unsigned *dvotes, *dcell;
  unsigned char *dstale;
};

__device__ __forceinline__ int point_count(const HoughDev &h) { return h.n_dev ? *h.n_dev : h.n; }

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
__device__ __forceinline__ long long ref_cell_slow(double v, double max_x, double dx) {
  const double q = __ddiv_rn(__dadd_rn(v, max_x), dx);
  const double r = (q > 0.0) ? floor(__dadd_rn(q, 0.5)) : ceil(__dsub_rn(q, 0.5));
  if (!(fabs(r) < 1099511627776.0)) return 1099511627776LL;
  return (long long)r;
}

// Exact plane cell x_i*num_x + y_i of one vote, or 0xFFFFFFFF when the reference's
// flat-index test `index < VotingSpace.size()` (lib/hough.cpp:92) drops it.
// Only the FLAT bound is tested, so y_i == num_x aliases into the next row
// exactly as the reference does.  This is the rarely taken exact path.
// (scalar arguments only: a by-reference HoughDev would force the kernel parameters
// into local memory at every call site)
__device__ __noinline__ unsigned vote_cell_exact(const DirCoef *coef_j, double max_x, double dx, int num_x,
                                                 double px, double py, double pz) {
  const DirCoef c = *coef_j;
  const long long xi = ref_cell_slow(ref_x_new(c, px, py, pz), max_x, dx);
  const long long yi = ref_cell_slow(ref_y_new(c, px, py, pz), max_x, dx);
  const long long cell = xi * (long long)num_x + yi;   // |.| < 2^40 * 2^31: no overflow
  return (cell >= 0 && cell < (long long)num_x * num_x) ? (unsigned)cell : 0xFFFFFFFFu;
}

// Fast path.  Per direction five coefficients pre-scaled by inv_s = 2^20/dx:
//   sx = fma(c0s,px, fma(c1n,py, fma(bxn,pz, K)))     K = 1.5*2^52 + round(max_x*inv_s + 2^19)
// so the low mantissa word of sx is n = round(t * 2^20) with t = (x'+max_x)/dx + 0.5 and an
// error below 2.5 units of 2^-20 cell (three roundings to integer units, the rounding of K,
// ~1e-6 units from the scaled coefficients; the reference's own rounding slack is ~1e-8
// units).  When the 20 fraction bits of n are at least 8 units away from a cell boundary,
// floor(t) is certain and equals the reference's roundToNearest; otherwise (probability
// ~3e-5 per vote) the exact path decides.  6 DFMA per vote instead of 12 FP64 ops + 2 divisions.
struct FastCoef {
  double c0s, c1n, c2s, bxn, byn;
};
__device__ __forceinline__ FastCoef make_fast(const DirCoef &c, double inv_s) {
  FastCoef f;
  f.c0s = c.c0 * inv_s; f.c1n = -(c.c1 * inv_s); f.c2s = c.c2 * inv_s;
  f.bxn = -(c.bx * inv_s); f.byn = -(c.by * inv_s);
  return f;
}

// branch-free part: candidate cell and whether it is certain
__device__ __forceinline__ unsigned vote_cell_fast(const FastCoef &f, double kmagic, unsigned num_x, double px,
                                                   double py, double pz, bool *sure) {
  const double sx = __fma_rn(f.c0s, px, __fma_rn(f.c1n, py, __fma_rn(f.bxn, pz, kmagic)));
  const double sy = __fma_rn(f.c1n, px, __fma_rn(f.c2s, py, __fma_rn(f.byn, pz, kmagic)));
  const unsigned lx = (unsigned)__double2loint(sx), ly = (unsigned)__double2loint(sy);
  const unsigned hbad = ((unsigned)__double2hiint(sx) ^ 0x43380000u) | ((unsigned)__double2hiint(sy) ^ 0x43380000u);
  *sure = (hbad == 0) & (((lx << 12) - 0x8000u) < 0xFFFF0000u) & (((ly << 12) - 0x8000u) < 0xFFFF0000u);
  return (lx >> 20) * num_x + (ly >> 20);
}

// same, for the unrolled hot loops: instead of a per-vote flag the guards are ACCUMULATED
// (gmin = min over votes of the distance-to-boundary word, hacc = OR of the exponent
// mismatches) and tested once per batch; the rare unsure batch is then re-examined vote by
// vote with vote_cell_fast.  4 integer ops of guard per vote instead of ~8.
__device__ __forceinline__ unsigned vote_cell_acc(const FastCoef &f, double kmagic, unsigned num_x, double px,
                                                  double py, double pz, unsigned &gmin, unsigned &hacc) {
  const double sx = __fma_rn(f.c0s, px, __fma_rn(f.c1n, py, __fma_rn(f.bxn, pz, kmagic)));
  const double sy = __fma_rn(f.c1n, px, __fma_rn(f.c2s, py, __fma_rn(f.byn, pz, kmagic)));
  const unsigned lx = (unsigned)__double2loint(sx), ly = (unsigned)__double2loint(sy);
  gmin = min(gmin, min((lx << 12) + 0x8000u, (ly << 12) + 0x8000u));   // < 0x10000: within 8 units
  hacc |= ((unsigned)__double2hiint(sx) ^ 0x43380000u) | ((unsigned)__double2hiint(sy) ^ 0x43380000u);
  return (lx >> 20) * num_x + (ly >> 20);
}
// BOUNDED form (HoughDev::bounded, all points inside the box): byte offset of the cell inside its
// plane, 4*(x_i*num_x + y_i) with row4 = 4*num_x.  No exponent test (0 < n < 2^32 is guaranteed) and
// no range test: x_i, y_i <= num_x + 1, so the offset stays inside the plane's slack cells, which the
// scan ignores.  14 instructions per vote instead of 19.
__device__ __forceinline__ unsigned vote_off_bounded(const FastCoef &f, double kmagic, unsigned row4, double px,
                                                     double py, double pz, unsigned &gmin) {
  const double sx = __fma_rn(f.c0s, px, __fma_rn(f.c1n, py, __fma_rn(f.bxn, pz, kmagic)));
  const double sy = __fma_rn(f.c1n, px, __fma_rn(f.c2s, py, __fma_rn(f.byn, pz, kmagic)));
  const unsigned lx = (unsigned)__double2loint(sx), ly = (unsigned)__double2loint(sy);
  gmin = min(gmin, min((lx << 12) + 0x8000u, (ly << 12) + 0x8000u));   // < 0x10000: within 8 units
  return (lx >> 20) * row4 + ((ly >> 20) << 2);
}

__device__ __forceinline__ bool batch_unsure(unsigned gmin, unsigned hacc) { return (gmin < 0x10000u) | (hacc != 0u); }

// shared-memory increment by 32-bit shared address, predicated, no return value
__device__ __forceinline__ void red_shared_inc(unsigned addr) {
  asm volatile("red.shared.add.u32 [%0], 1;" :: "r"(addr) : "memory");
}
// shared-memory increment that returns the old count (ATOMS.ADD): the last voter of a cell sees its final
// count - 1, so max over all votes of (old + 1) is the plane maximum -- no scan of the plane needed
__device__ __forceinline__ unsigned atom_shared_inc_ret(unsigned addr) {
  unsigned old;
  asm volatile("atom.shared.add.u32 %0, [%1], 1;" : "=r"(old) : "r"(addr) : "memory");
  return old;
}
__device__ __forceinline__ void red_shared_add(unsigned addr, unsigned val, bool pred) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %2, 0;\n\t@p red.shared.add.u32 [%0], %1;\n\t}"
      :: "r"(addr), "r"(val), "r"((unsigned)pred) : "memory");
}
__device__ __forceinline__ void cp_async8(unsigned dst_shared, const void *src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" :: "r"(dst_shared), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// Rare path of the unrolled loops, deliberately out of line and self-contained (it shares
// no intermediate with the fast path, so the fast path keeps nothing alive for it):
// recompute one vote, return the certain fast cell or the exact one.
__device__ __noinline__ unsigned vote_cell_checked(const DirCoef *coef_j, double inv_s, double kmagic,
                                                   double max_x, double dx, int num_x, int fast_ok, double px,
                                                   double py, double pz) {
  bool sure;
  const unsigned cell = vote_cell_fast(make_fast(*coef_j, inv_s), kmagic, (unsigned)num_x, px, py, pz, &sure);
  if (sure && fast_ok) return cell;
  return vote_cell_exact(coef_j, max_x, dx, num_x, px, py, pz);
}

// same, reading everything it needs from the kernel's __grid_constant__ HoughDev: the call site
// passes one pointer and three coordinates, so the hot loop keeps no register or constant load
// alive for the rare path
__device__ __noinline__ unsigned vote_cell_checked_h(const HoughDev *h, int l, double px, double py, double pz) {
  return vote_cell_checked(h->coef + h->dir_begin + l, h->inv_s, h->kmagic, h->max_x, h->dx, h->num_x, h->fast_ok,
                           px, py, pz);
}

// cell of one vote; the caller tests cell < plane_cells (0xFFFFFFFF from the exact path fails it)
__device__ __forceinline__ unsigned vote_cell(const HoughDev &h, const FastCoef &f, int j, double px,
                                              double py, double pz) {
  bool sure;
  unsigned cell = vote_cell_fast(f, h.kmagic, (unsigned)h.num_x, px, py, pz, &sure);
  if (!sure | !h.fast_ok) cell = vote_cell_exact(h.coef + j, h.max_x, h.dx, h.num_x, px, py, pz);
  return cell;
}

// Which points of Y can vote for cell (cx, cy) of the direction with scaled coefficients f?  Y lies in the
// slab { a + t b + e : |e| <= radius, e orthogonal to b }.  The map to the (x', y') plane is an orthogonal
// projection, so in cell units a point of parameter t lands within radius/dx of P(t) = (xa + t gx, ya + t gy)
// (relative to the cell's centre), and it votes for the cell only if it lands within sqrt(2)/2 of the centre:
// |P(t)| <= reach, reach^2 = reach2 >= (sqrt(2)/2 + radius/dx)^2.  Returns false when no t qualifies, else the
// interval [*tlo, *thi] (conservative, never exact; +-inf when the line is parallel to the direction).
__device__ __forceinline__ bool strip_interval(const FastCoef &f, const double (&a)[3], const double (&b)[3],
                                               double cell_off, double reach2, unsigned cx, unsigned cy, double *tlo,
                                               double *thi) {
  const double s = 1.0 / 1048576.0;   // the coefficients carry 2^20/dx
  const double xa = (f.c0s * a[0] + f.c1n * a[1] + f.bxn * a[2]) * s + cell_off - ((double)cx + 0.5);
  const double ya = (f.c1n * a[0] + f.c2s * a[1] + f.byn * a[2]) * s + cell_off - ((double)cy + 0.5);
  const double gx = (f.c0s * b[0] + f.c1n * b[1] + f.bxn * b[2]) * s;
  const double gy = (f.c1n * b[0] + f.c2s * b[1] + f.byn * b[2]) * s;
  const double gg = gx * gx + gy * gy, bq = xa * gx + ya * gy, cq = xa * xa + ya * ya - reach2;
  if (!(gg > 1e-12)) {                // the whole line projects (almost) to one spot
    *tlo = -INFINITY; *thi = INFINITY;
    return cq <= 1e-6;
  }
  const double disc = bq * bq - gg * cq;
  if (disc < 0.0) return false;
  const double sq = sqrt(disc);
  *tlo = (-bq - sq) / gg; *thi = (-bq + sq) / gg;
  return true;
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
