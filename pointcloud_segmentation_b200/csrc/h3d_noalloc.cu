// h3d_noalloc.cu -- PointCloud::pointsCloseToLine (lib/pointcloud.cpp:94-105) for callers that must not
// allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.
//
// h3d_points_close_to_line() uploads the cloud into the library's growing scratch arena (cudaMalloc on first
// use and whenever a larger cloud arrives) and keeps its one-byte-per-point inlier flags there.  The variant
// below takes DEVICE pointers for its input and outputs, keeps the flags in a file-scope __device__ array of
// fixed size and therefore calls no cudaMalloc / cudaFree / allocating library routine, ever.
//
// The array is part of the library image: every GPU that loads libh3d.so holds kNoAllocMaxPoints bytes
// (1 MiB) of it until the process exits.  Every call on a device shares it and it keeps the last call's
// contents; nothing in it has to start at zero, because the mask kernel writes flags[0, n) before the
// compaction kernel (same stream, launched after it) reads exactly that range.  Two calls whose kernels overlap
// on the GPU (different streams, or different host threads on different streams) race on it: serialise them.
#include "../../include/h3d.h"

#include <cuda_runtime.h>

#include <cstdio>
#include <mutex>
#include <string>

#include "h3d_internal.h"   // set_error only: the device headers hold out-of-line device functions that belong to
                            // h3d_capi.cu's translation unit

namespace h3d {

constexpr size_t kNoAllocMaxPoints = H3D_NOALLOC_MAX_POINTS;
constexpr int kNoAllocMaxDevices = 64;

// In host code this name is only a stand-in for the symbol (not a device address): the address a kernel can use
// comes from cudaGetSymbolAddress, below.  Device code could name it directly.
static __device__ __align__(256) unsigned char g_noalloc_flags[kNoAllocMaxPoints];

// is p within dx of the line (a, b)?  lib/pointcloud.cpp:99-101, the reference's expression tree in IEEE double
// without FMA (same operations in the same order as point_near_line of h3d_fit.cuh)
__device__ __forceinline__ bool noalloc_near_line(double px, double py, double pz, const double a[3],
                                                  const double b[3], double dx) {
  const double ex = __dsub_rn(px, a[0]), ey = __dsub_rn(py, a[1]), ez = __dsub_rn(pz, a[2]);
  const double t = __dadd_rn(__dadd_rn(__dmul_rn(b[0], ex), __dmul_rn(b[1], ey)), __dmul_rn(b[2], ez));   // b * (p - a)
  const double fx = __dsub_rn(px, __dadd_rn(a[0], __dmul_rn(t, b[0])));                                   // p - (a + t*b)
  const double fy = __dsub_rn(py, __dadd_rn(a[1], __dmul_rn(t, b[1])));
  const double fz = __dsub_rn(pz, __dadd_rn(a[2], __dmul_rn(t, b[2])));
  const double nn = __dadd_rn(__dadd_rn(__dmul_rn(fx, fx), __dmul_rn(fy, fy)), __dmul_rn(fz, fz));
  return __dsqrt_rn(nn) <= dx;                                                                            // inclusive
}

// inlier flags of every point
__global__ void __launch_bounds__(256) noalloc_mask_kernel(const double *__restrict__ p, int n, double a0, double a1,
                                                           double a2, double b0, double b1, double b2, double dx,
                                                           unsigned char *__restrict__ flags) {
  const double a[3] = {a0, a1, a2}, b[3] = {b0, b1, b2};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    flags[i] = noalloc_near_line(p[3 * (size_t)i], p[3 * (size_t)i + 1], p[3 * (size_t)i + 2], a, b, dx) ? 1 : 0;
}

// indices of the flagged points in input order (one CTA, 1024 points per trip, running base in a register of
// every thread); *n_out = how many
__global__ void __launch_bounds__(1024, 1) noalloc_compact_kernel(const unsigned char *__restrict__ flags, int n,
                                                                  unsigned *__restrict__ idx, int *n_out) {
  __shared__ int warp_sum[2][32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  int base = 0, buf = 0;
  for (int c0 = 0; c0 < n; c0 += 1024, buf ^= 1) {
    const int i = c0 + tid;
    const bool keep = i < n && flags[i];
    const unsigned bal = __ballot_sync(0xFFFFFFFFu, keep);
    if (lane == 0) warp_sum[buf][warp] = __popc(bal);
    __syncthreads();                                  // one barrier per trip: the sums are double-buffered
    const int mine = warp_sum[buf][lane];
    int incl = mine;                                  // inclusive prefix over the 32 warp sums
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xFFFFFFFFu, incl, o);
      if (lane >= o) incl += v;
    }
    const int before = __shfl_sync(0xFFFFFFFFu, incl - mine, warp);
    if (keep) idx[base + before + __popc(bal & ((1u << lane) - 1u))] = (unsigned)i;
    base += __shfl_sync(0xFFFFFFFFu, incl, 31);
  }
  if (tid == 0) *n_out = base;
}

// set-up, once per device, with that device current: the device address of the scratch array
static int noalloc_scratch(unsigned char **out) {
  static std::mutex mu;
  static unsigned char *addr[kNoAllocMaxDevices] = {};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kNoAllocMaxDevices) {
    set_error(std::string("h3d_points_close_to_line_noalloc: no usable CUDA device (") +
              (e != cudaSuccess ? cudaGetErrorString(e) : "device ordinal out of range") + ")");
    return H3D_ERR_CUDA;
  }
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[dev]) {
    void *p = nullptr;
    e = cudaGetSymbolAddress(&p, g_noalloc_flags);
    if (e != cudaSuccess || !p) {
      set_error(std::string("h3d_points_close_to_line_noalloc: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e));
      return H3D_ERR_CUDA;
    }
    addr[dev] = static_cast<unsigned char *>(p);
  }
  *out = addr[dev];
  return H3D_OK;
}

}  // namespace h3d

extern "C" int h3d_points_close_to_line_noalloc(const double *xyz_dev, size_t n, const double a[3], const double b[3],
                                                double dx, uint32_t *idx_out_dev, int *n_out_dev, void *stream) {
  using namespace h3d;
  if (!a || !b || !n_out_dev || (n && (!xyz_dev || !idx_out_dev))) {
    set_error("h3d_points_close_to_line_noalloc: bad argument");
    return H3D_ERR_ARG;
  }
  if (n > kNoAllocMaxPoints) {
    char msg[200];
    snprintf(msg, sizeof msg, "h3d_points_close_to_line_noalloc: %zu points exceed the fixed scratch of %zu points "
             "(H3D_NOALLOC_MAX_POINTS); use h3d_points_close_to_line", n, (size_t)kNoAllocMaxPoints);
    set_error(msg);
    return H3D_ERR_RANGE;
  }
  unsigned char *flags = nullptr;
  const int rc = noalloc_scratch(&flags);
  if (rc != H3D_OK) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (n) {
    const unsigned blocks = (unsigned)((n + 255) / 256);
    noalloc_mask_kernel<<<blocks < 2368u ? blocks : 2368u, 256, 0, st>>>(xyz_dev, (int)n, a[0], a[1], a[2], b[0], b[1],
                                                                        b[2], dx, flags);
  }
  noalloc_compact_kernel<<<1, 1024, 0, st>>>(flags, (int)n, idx_out_dev, n_out_dev);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error(std::string("h3d_points_close_to_line_noalloc: launch failed: ") + cudaGetErrorString(e));
    return H3D_ERR_CUDA;
  }
  return H3D_OK;
}
