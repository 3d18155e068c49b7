// h3d_capi.cu -- the extern "C" boundary of libh3d.so (include/h3d.h): contexts,
// device memory, kernel launches.  No CPU fallback: every compute entry point
// needs a CUDA device and fails with H3D_ERR_CUDA otherwise.
#include "../../include/h3d.h"

#include <dlfcn.h>
#include <nccl.h>   // types and enums only: NCCL is bound with dlopen, no symbol of it is linked

#include <atomic>
#include <cfloat>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <map>
#include <mutex>
#include <new>
#include <thread>

#include "h3d_fused.cuh"
#include "h3d_prefilter.cuh"
#include "h3d_segment.cuh"

namespace h3d {

static thread_local std::string g_error;
static std::atomic<unsigned long long> g_launches{0};   // kernels launched by this library (bench.py: gpu_launches)
static inline void note_launches(unsigned n) { g_launches += n; }
void set_error(const std::string &msg) { g_error = msg; }

#define H3D_CUDA(call)                                                                       \
  do {                                                                                       \
    cudaError_t e_ = (call);                                                                 \
    if (e_ != cudaSuccess) {                                                                 \
      char buf_[512];                                                                        \
      snprintf(buf_, sizeof buf_, "%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
      set_error(buf_);                                                                       \
      return e_ == cudaErrorMemoryAllocation ? H3D_ERR_NOMEM : H3D_ERR_CUDA;                 \
    }                                                                                        \
  } while (0)

#define H3D_TRY(call)             \
  do {                            \
    int rc_ = (call);             \
    if (rc_ != H3D_OK) return rc_; \
  } while (0)

static int fail(int code, const char *msg) {
  set_error(msg);
  return code;
}

// ---- device facts -----------------------------------------------------------
struct DeviceFacts {
  int ok = 0, sm_count = 0, cc = 0;
  size_t smem_optin = 0;
};
static int device_facts(DeviceFacts *out) {
  static std::mutex mu;
  static std::map<int, DeviceFacts> cache;
  int dev = 0;
  H3D_CUDA(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lock(mu);
  auto it = cache.find(dev);
  if (it == cache.end()) {
    cudaDeviceProp p;
    H3D_CUDA(cudaGetDeviceProperties(&p, dev));
    DeviceFacts f;
    f.ok = 1; f.sm_count = p.multiProcessorCount; f.cc = p.major * 10 + p.minor;
    f.smem_optin = p.sharedMemPerBlockOptin;
    it = cache.emplace(dev, f).first;
  }
  *out = it->second;
  return H3D_OK;
}

// cudaFuncAttributeMaxDynamicSharedMemorySize is a property of (device, kernel), not of the calling
// thread: one host thread may drive pipelines on several GPUs (h3d_pipeline_run sets the device), and
// many threads share one device.  The opt-in is cached per (device ordinal, kernel address).
static int ensure_dyn_smem(const void *kernel, size_t want, size_t smem_optin) {
  static std::mutex mu;
  static std::map<std::pair<int, const void *>, size_t> configured;
  int dev = 0;
  H3D_CUDA(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lock(mu);
  size_t &have = configured[std::make_pair(dev, kernel)];
  if (want <= have) return H3D_OK;
  cudaFuncAttributes fa;
  H3D_CUDA(cudaFuncGetAttributes(&fa, kernel));
  if (smem_optin < fa.sharedSizeBytes) return fail(H3D_ERR_RANGE, "kernel's static shared memory exceeds the opt-in limit");
  const size_t dyn_max = smem_optin - fa.sharedSizeBytes;   // the opt-in limit covers static + dynamic
  if (want > dyn_max) return fail(H3D_ERR_RANGE, "dynamic shared memory request exceeds the device limit");
  H3D_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn_max));
  have = dyn_max;
  return H3D_OK;
}

// ---- direction spaces (lib/hough.cpp:15-20: one global sphere) ---------------
static std::mutex g_space_mu;
static std::map<std::pair<int, unsigned>, std::shared_ptr<const Space>> g_spaces;   // (device, g)
static std::shared_ptr<const Space> g_current;

int get_space(unsigned g, std::shared_ptr<const Space> *out) {
  if (g > 8) return fail(H3D_ERR_ARG, "granularity must be 0..8");
  int dev = 0;
  H3D_CUDA(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lock(g_space_mu);
  auto key = std::make_pair(dev, g);
  auto it = g_spaces.find(key);
  if (it != g_spaces.end()) { *out = it->second; return H3D_OK; }
  auto sp = std::make_shared<Space>();
  sp->granularity = g;
  build_sphere((int)g, sp->dirs, sp->triangles);
  sp->num_b = sp->dirs.size() / 3;
  std::vector<DirCoef> coef(sp->num_b);
  for (size_t j = 0; j < sp->num_b; ++j) {
    const double bx = sp->dirs[3 * j], by = sp->dirs[3 * j + 1], bz = sp->dirs[3 * j + 2];
    const double beta = 1 / (1 + bz);                 // lib/hough.cpp:73
    coef[j].c0 = 1 - (beta * (bx * bx));              // :76
    coef[j].c1 = beta * (bx * by);                    // :77, :81
    coef[j].c2 = 1 - (beta * (by * by));              // :82
    coef[j].bx = bx; coef[j].by = by;
  }
  H3D_CUDA(cudaMalloc(&sp->coef_dev, sizeof(DirCoef) * sp->num_b));
  H3D_CUDA(cudaMalloc(&sp->dirs_dev, sizeof(double) * 3 * sp->num_b));
  H3D_CUDA(cudaMemcpy(sp->coef_dev, coef.data(), sizeof(DirCoef) * sp->num_b, cudaMemcpyHostToDevice));
  H3D_CUDA(cudaMemcpy(sp->dirs_dev, sp->dirs.data(), sizeof(double) * 3 * sp->num_b, cudaMemcpyHostToDevice));
  g_spaces[key] = sp;
  *out = sp;
  return H3D_OK;
}

std::shared_ptr<const Space> current_space() {
  std::lock_guard<std::mutex> lock(g_space_mu);
  return g_current;
}

}  // namespace h3d

using namespace h3d;

// ---- the handle ----------------------------------------------------------------
struct h3d_hough {
  std::shared_ptr<const Space> space;
  HoughDev d{};
  DeviceFacts dev;
  cudaStream_t stream = nullptr;
  size_t cap = 0;                 // point capacity
  double *px = nullptr, *py = nullptr, *pz = nullptr;
  int *w = nullptr;
  double *stage = nullptr;        // AoS staging for uploads
  size_t stage_cap = 0;           // in points
  bool all_dirty = true;          // no usable per-direction record
  bool has_negative = false;      // a subtract() hit no added point: planes may wrap
  unsigned long long *keys = nullptr;   // [0] best exact, [1] best overall
  int *list = nullptr, *count = nullptr;
  int *idx = nullptr;             // scratch: indices of removed points
  size_t idx_cap = 0;
  int vote_d = 1;                 // directions per work item of the full vote
  unsigned *gplanes = nullptr;    // sliced vote form: dir_count global planes (few directions, many points)
  size_t gplanes_cells = 0;
  unsigned *rplanes = nullptr;    // point-sliced list re-votes of the streaming vote: rplanes_cap global planes, kept zero
  int rplanes_cap = 0, *rflag = nullptr;
  size_t rplanes_cells = 0;       // plane_cells the buffer was sized for
  uint64_t votes_cast = 0;
};

extern "C" const char *h3d_last_error(void) { return g_error.c_str(); }
extern "C" unsigned long long h3d_launch_count(void) { return g_launches.load(); }

extern "C" int h3d_device_info(int *sm_count, int *cc, size_t *smem_optin) {
  DeviceFacts f;
  H3D_TRY(device_facts(&f));
  if (sm_count) *sm_count = f.sm_count;
  if (cc) *cc = f.cc;
  if (smem_optin) *smem_optin = f.smem_optin;
  return H3D_OK;
}

extern "C" int h3d_init_space(unsigned granularity) {
  std::shared_ptr<const Space> sp;
  H3D_TRY(get_space(granularity, &sp));
  std::lock_guard<std::mutex> lock(g_space_mu);
  g_current = sp;
  return H3D_OK;
}

extern "C" int h3d_space_info(unsigned *granularity, size_t *num_b) {
  auto sp = current_space();
  if (!sp) return fail(H3D_ERR_ARG, "h3d_init_space() has not been called");
  if (granularity) *granularity = sp->granularity;
  if (num_b) *num_b = sp->num_b;
  return H3D_OK;
}

extern "C" int h3d_sphere(int g, double *vertices, size_t cap_v, size_t *n_v, unsigned *triangles,
                          size_t cap_t, size_t *n_t) {
  if (g < 0 || g > 8) return fail(H3D_ERR_ARG, "granularity must be 0..8");
  std::vector<double> v;
  std::vector<unsigned> t;
  build_sphere(g, v, t);   // pure host geometry, no device needed
  if (n_v) *n_v = v.size() / 3;
  if (n_t) *n_t = t.size();
  if (vertices) {
    if (cap_v < v.size() / 3) return fail(H3D_ERR_ARG, "vertex buffer too small");
    memcpy(vertices, v.data(), v.size() * sizeof(double));
  }
  if (triangles) {
    if (cap_t < t.size()) return fail(H3D_ERR_ARG, "triangle buffer too small");
    memcpy(triangles, t.data(), t.size() * sizeof(unsigned));
  }
  return H3D_OK;
}

// ---- Hough: lib/hough.cpp:26-135 ---------------------------------------------------
static int ensure_points(h3d_hough *h, size_t need) {
  if (need <= h->cap) return H3D_OK;
  size_t cap = h->cap ? h->cap : 4096;
  while (cap < need) cap *= 2;
  double *nx, *ny, *nz; int *nw;
  H3D_CUDA(cudaMalloc(&nx, cap * sizeof(double)));
  H3D_CUDA(cudaMalloc(&ny, cap * sizeof(double)));
  H3D_CUDA(cudaMalloc(&nz, cap * sizeof(double)));
  H3D_CUDA(cudaMalloc(&nw, cap * sizeof(int)));
  if (h->d.n) {
    H3D_CUDA(cudaMemcpyAsync(nx, h->px, h->d.n * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
    H3D_CUDA(cudaMemcpyAsync(ny, h->py, h->d.n * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
    H3D_CUDA(cudaMemcpyAsync(nz, h->pz, h->d.n * sizeof(double), cudaMemcpyDeviceToDevice, h->stream));
    H3D_CUDA(cudaMemcpyAsync(nw, h->w, h->d.n * sizeof(int), cudaMemcpyDeviceToDevice, h->stream));
    H3D_CUDA(cudaStreamSynchronize(h->stream));
  }
  cudaFree(h->px); cudaFree(h->py); cudaFree(h->pz); cudaFree(h->w);
  h->px = nx; h->py = ny; h->pz = nz; h->w = nw; h->cap = cap;
  h->d.px = nx; h->d.py = ny; h->d.pz = nz; h->d.w = nw;
  return H3D_OK;
}

static int ensure_stage(h3d_hough *h, size_t n) {
  if (n <= h->stage_cap) return H3D_OK;
  cudaFree(h->stage); h->stage = nullptr;
  cudaFree(h->idx); h->idx = nullptr;
  size_t cap = h->stage_cap ? h->stage_cap : 4096;
  while (cap < n) cap *= 2;
  H3D_CUDA(cudaMalloc(&h->stage, cap * 3 * sizeof(double)));
  H3D_CUDA(cudaMalloc(&h->idx, cap * sizeof(int)));
  h->stage_cap = cap; h->idx_cap = cap;
  return H3D_OK;
}

// AoS staging -> SoA append, weight +1
__global__ void append_points_kernel(const double *__restrict__ aos, int n, double *px, double *py, double *pz,
                                     int *w, int at) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    px[at + i] = aos[3 * i]; py[at + i] = aos[3 * i + 1]; pz[at + i] = aos[3 * i + 2];
    w[at + i] = 1;
  }
}

// Hough::subtract bookkeeping: y_k is matched to the (r+1)-th live point with the
// same coordinates, r = number of earlier identical y.  idx[k] = matched point or
// -1.  Brute force: |Y| threads x |X| compares.
__global__ void match_points_kernel(const double *__restrict__ y, int ny, const double *__restrict__ px,
                                    const double *__restrict__ py, const double *__restrict__ pz,
                                    const int *__restrict__ w, int n, int *idx, int *unmatched) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= ny) return;
  const double x0 = y[3 * k], y0 = y[3 * k + 1], z0 = y[3 * k + 2];
  int rank = 0;
  for (int q = 0; q < k; ++q) rank += (y[3 * q] == x0 && y[3 * q + 1] == y0 && y[3 * q + 2] == z0);
  int found = -1;
  for (int i = 0; i < n; ++i) {
    if (w[i] > 0 && px[i] == x0 && py[i] == y0 && pz[i] == z0) {
      if (rank-- == 0) { found = i; break; }
    }
  }
  idx[k] = found;
  if (found < 0) atomicAdd(unmatched, 1);
}

__global__ void kill_points_kernel(const int *__restrict__ idx, int n_idx, int *w) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < n_idx && idx[k] >= 0) w[idx[k]] = 0;
}

// stable in-place compaction of the live points (weight != 0); one CTA walks the array in
// chunks, so a chunk is entirely in registers before anything at or below it is rewritten
__global__ void __launch_bounds__(1024, 1)
compact_points_kernel(double *px, double *py, double *pz, int *w, int n) {
  __shared__ int warp_sum[32];
  __shared__ int base_s;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) base_s = 0;
  __syncthreads();
  for (int c0 = 0; c0 < n; c0 += 1024) {
    const int i = c0 + tid;
    const int wv = i < n ? w[i] : 0;
    const bool keep = wv != 0;
    double x = 0, y = 0, z = 0;
    if (keep) { x = px[i]; y = py[i]; z = pz[i]; }
    const unsigned bal = __ballot_sync(0xFFFFFFFFu, keep);
    if (lane == 0) warp_sum[warp] = __popc(bal);
    __syncthreads();
    int off = base_s;
    for (int q = 0; q < warp; ++q) off += warp_sum[q];
    int total = 0;
    if (tid == 0) for (int q = 0; q < 32; ++q) total += warp_sum[q];
    const int pos = off + __popc(bal & ((1u << lane) - 1u));
    if (keep) { px[pos] = x; py[pos] = y; pz[pos] = z; w[pos] = wv; }
    __syncthreads();
    if (tid == 0) base_s += total;
    __syncthreads();
  }
}

// unmatched subtract: append as weight -1 points (unsigned cells wrap like the reference's)
__global__ void append_negative_kernel(const double *__restrict__ y, const int *__restrict__ idx, int ny,
                                       double *px, double *py, double *pz, int *w, int at, int *cursor) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= ny || idx[k] >= 0) return;
  const int slot = at + atomicAdd(cursor, 1);
  px[slot] = y[3 * k]; py[slot] = y[3 * k + 1]; pz[slot] = y[3 * k + 2];
  w[slot] = -1;
}

template <int D, int ABLATE>
static int launch_vote_t(h3d_hough *h, const int *list, const int *count_ptr, int count_imm, unsigned *dense,
                         int grid) {
  const size_t smem = (size_t)D * h->d.plane_stride * sizeof(unsigned);
  H3D_TRY(ensure_dyn_smem(reinterpret_cast<const void *>(vote_kernel<D, ABLATE>), smem, h->dev.smem_optin));
  vote_kernel<D, ABLATE><<<grid, kVoteThreads, smem, h->stream>>>(h->d, list, count_ptr, count_imm, dense);
  note_launches(1);
  H3D_CUDA(cudaGetLastError());
  return H3D_OK;
}

template <int P, int ABLATE>
static int launch_resident_t(h3d_hough *h, const int *list, const int *count_ptr, int count_imm) {
  const size_t smem = ((size_t)2 * kResG * h->d.plane_stride + 4) * sizeof(unsigned);
  H3D_TRY(ensure_dyn_smem(reinterpret_cast<const void *>(vote_resident_kernel<P, ABLATE>), smem, h->dev.smem_optin));
  vote_resident_kernel<P, ABLATE><<<h->dev.sm_count, kResThreads, smem, h->stream>>>(h->d, list, count_ptr,
                                                                                    count_imm);
  note_launches(1);
  H3D_CUDA(cudaGetLastError());
  return H3D_OK;
}

// first vote of the fused path: BOUNDED form, `grid` CTAs (the pipeline's SM share)
template <int P>
static int launch_first_vote_t(h3d_hough *h, int grid) {
  const size_t smem = ((size_t)2 * kResG * h->d.plane_stride + 4) * sizeof(unsigned);
  H3D_TRY(ensure_dyn_smem(reinterpret_cast<const void *>(vote_resident_kernel<P, 0, true>), smem, h->dev.smem_optin));
  vote_resident_kernel<P, 0, true><<<grid, kResThreads, smem, h->stream>>>(h->d, nullptr, nullptr, h->d.dir_count);
  note_launches(1);
  H3D_CUDA(cudaGetLastError());
  return H3D_OK;
}

constexpr int kResidentMaxP = 12;
static bool resident_ok(const h3d_hough *h) {
  return !h->has_negative && h->d.n <= kResThreads * kResidentMaxP &&
         ((size_t)2 * kResG * h->d.plane_stride + 4) * sizeof(unsigned) <= h->dev.smem_optin - 2048;
}

static int launch_resident(h3d_hough *h, int ablate, const int *list, const int *count_ptr, int count_imm) {
  const int need = (h->d.n + kResThreads - 1) / kResThreads;
#define H3D_RES_CASE(PP)                                                                     \
  if (need <= PP) {                                                                          \
    if (ablate == 0) return launch_resident_t<PP, 0>(h, list, count_ptr, count_imm);         \
    if (ablate == 1) return launch_resident_t<PP, 1>(h, list, count_ptr, count_imm);         \
    return launch_resident_t<PP, 2>(h, list, count_ptr, count_imm);                          \
  }
  H3D_RES_CASE(4)
  H3D_RES_CASE(8)
  H3D_RES_CASE(10)
  H3D_RES_CASE(12)
#undef H3D_RES_CASE
  return fail(H3D_ERR_ARG, "too many points for the resident vote kernel");
}

// TMA-staged streaming vote (vote_stream_kernel): D <= 8 planes + the point-tile ring in shared memory
// h3d_set_stream_mode: 0 (default) tile loop with coalesced LDGs, 1 the same loop fed by TMA bulk copies through a
// producer warp, 2 the round-1 vote_kernel.  Measured on B200 (profiles/r02_stream_vote_ab_v2.txt, 1 M points, g6):
// LDG 20.2 ms (1012 Gvotes/s), TMA 21.9 ms (935), round 1 32.9 ms (622): the kernel is bound by the shared-memory
// pipe (the atomics), and staging tiles in shared memory adds their 24 B per point to that same pipe.
static std::atomic<int> g_stream_mode{0};

constexpr int kStreamSlicePlanes = 2048;   // list re-votes of up to this many planes may be point-sliced (33 MB at num_x 64)

template <int D, bool BOUNDED, bool TMA>
static int launch_stream_t(h3d_hough *h, const int *list, const int *count_ptr, int count_imm, unsigned *dense, size_t smem) {
  H3D_TRY(ensure_dyn_smem(reinterpret_cast<const void *>(vote_stream_kernel<D, BOUNDED, TMA>), smem, h->dev.smem_optin));
  // a re-vote LIST over a large cloud may be point-sliced on the device (see the kernel): global planes for it
  const bool may_slice = !TMA && list && !dense && h->d.n >= 8 * kStreamTile;
  if (may_slice && (h->rplanes_cells != h->d.plane_cells || !h->rplanes)) {
    H3D_CUDA(cudaStreamSynchronize(h->stream));
    cudaFree(h->rplanes); h->rplanes = nullptr; h->rplanes_cap = 0;
    const size_t cells = (size_t)kStreamSlicePlanes * h->d.plane_cells;
    H3D_CUDA(cudaMalloc(&h->rplanes, cells * sizeof(unsigned)));
    H3D_CUDA(cudaMemsetAsync(h->rplanes, 0, cells * sizeof(unsigned), h->stream));
    if (!h->rflag) H3D_CUDA(cudaMalloc(&h->rflag, sizeof(int)));
    h->rplanes_cap = kStreamSlicePlanes; h->rplanes_cells = h->d.plane_cells;
  }
  vote_stream_kernel<D, BOUNDED, TMA><<<h->dev.sm_count, kStreamThreads + (TMA ? 32 : 0), smem, h->stream>>>(
      h->d, list, count_ptr, count_imm, dense, may_slice ? h->rplanes : nullptr, may_slice ? h->rplanes_cap : 0,
      may_slice ? h->rflag : nullptr);
  note_launches(1);
  if (may_slice) {
    stream_merge_kernel<<<h->dev.sm_count * 2, 256, 0, h->stream>>>(h->d, list, h->rplanes, h->rflag);
    note_launches(1);
  }
  H3D_CUDA(cudaGetLastError());
  return H3D_OK;
}

template <int D>
static int launch_stream_d(h3d_hough *h, const int *list, const int *count_ptr, int count_imm, unsigned *dense, size_t smem, bool tma) {
  if (h->d.bounded) return tma ? launch_stream_t<D, true, true>(h, list, count_ptr, count_imm, dense, smem)
                               : launch_stream_t<D, true, false>(h, list, count_ptr, count_imm, dense, smem);
  return tma ? launch_stream_t<D, false, true>(h, list, count_ptr, count_imm, dense, smem)
             : launch_stream_t<D, false, false>(h, list, count_ptr, count_imm, dense, smem);
}

// returns H3D_OK with *launched = false when the planes do not fit beside the tile ring
static int launch_stream(h3d_hough *h, const int *list, const int *count_ptr, int count_imm, unsigned *dense, bool tma,
                         bool *launched) {
  *launched = false;
  cudaFuncAttributes fa;
  H3D_CUDA(cudaFuncGetAttributes(&fa, vote_stream_kernel<1, false, true>));
  const size_t ring = (size_t)kStreamStages * 3 * kStreamTile * sizeof(double);
  const size_t plane_bytes = (size_t)h->d.plane_stride * sizeof(unsigned);
  if (h->dev.smem_optin < fa.sharedSizeBytes + 512 + ring + plane_bytes + 16) return H3D_OK;
  const size_t room = h->dev.smem_optin - fa.sharedSizeBytes - 512 - ring - 16;
  int D = (int)(room / plane_bytes);
  if (D > kStreamMaxD) D = kStreamMaxD;
  // no more planes than keep every SM busy: items = ceil(directions / D) >= 2 per CTA when possible
  const int dirs = count_ptr ? h->d.dir_count : count_imm;
  while (D > 1 && (dirs + D - 1) / D < 2 * h->dev.sm_count) --D;
  const size_t smem = ring + (size_t)D * plane_bytes + 16;
  *launched = true;
  switch (D) {
    case 1: return launch_stream_d<1>(h, list, count_ptr, count_imm, dense, smem, tma);
    case 2: return launch_stream_d<2>(h, list, count_ptr, count_imm, dense, smem, tma);
    case 3: return launch_stream_d<3>(h, list, count_ptr, count_imm, dense, smem, tma);
    case 4: return launch_stream_d<4>(h, list, count_ptr, count_imm, dense, smem, tma);
    case 5: return launch_stream_d<5>(h, list, count_ptr, count_imm, dense, smem, tma);
    case 6: return launch_stream_d<6>(h, list, count_ptr, count_imm, dense, smem, tma);
    case 7: return launch_stream_d<7>(h, list, count_ptr, count_imm, dense, smem, tma);
    default: return launch_stream_d<8>(h, list, count_ptr, count_imm, dense, smem, tma);
  }
}

// D planes per CTA, two CTAs per SM when they fit.  D == 0: pick the kernel form
// (register-resident points when they fit, else streaming with D = 2).
static bool g_sliced_vote = true;   // h3d_set_sliced_vote
static int launch_vote(h3d_hough *h, int D, int ablate, const int *list, const int *count_ptr, int count_imm,
                       unsigned *dense) {
  const size_t plane_bytes = (size_t)h->d.plane_stride * sizeof(unsigned);
  const size_t budget = h->dev.smem_optin - 2048;   // static warp_best + per-CTA reserve
  if (plane_bytes > budget) return fail(H3D_ERR_RANGE, "one (x',y') plane does not fit in shared memory");
  if (D == 0) {
    if (!dense && resident_ok(h)) return launch_resident(h, ablate, list, count_ptr, count_imm);
    // few directions, many points: slice the points of a direction over CTAs (full votes only)
    if (g_sliced_vote && !dense && !list && !count_ptr && ablate == 0 && count_imm == h->d.dir_count &&
        h->d.dir_count > 0 && h->d.dir_count <= h->dev.sm_count && h->d.n >= 65536 &&   // g0-g2, narrow shards (at 321 directions it lost at 4 M points)
        (size_t)h->d.dir_count * h->d.plane_cells <= ((size_t)64 << 20) / sizeof(unsigned)) {
      const size_t cells = (size_t)h->d.dir_count * h->d.plane_cells;
      if (cells > h->gplanes_cells) {
        cudaFree(h->gplanes); h->gplanes = nullptr; h->gplanes_cells = 0;
        H3D_CUDA(cudaMalloc(&h->gplanes, cells * sizeof(unsigned)));
        H3D_CUDA(cudaMemsetAsync(h->gplanes, 0, cells * sizeof(unsigned), h->stream));
        h->gplanes_cells = cells;
      }
      const int grid = h->dev.sm_count * 2;
      int slices = (4 * grid + h->d.dir_count - 1) / h->d.dir_count;   // ~4 items per CTA: balanced tail
      const int max_slices = h->d.n / 8192;                 // a slice is worth a CTA from 8 k points on
      if (slices > max_slices) slices = max_slices;
      const int min_slices = (h->d.n + 262143) / 262144;    // <= 256 k points (7 MB) per slice: measured, 4 M points
      if (slices < min_slices) slices = min_slices;
      if (slices < 1) slices = 1;
      const size_t smem = (size_t)h->d.plane_stride * sizeof(unsigned);
      H3D_TRY(ensure_dyn_smem(reinterpret_cast<const void *>(vote_sliced_kernel), smem, h->dev.smem_optin));
      vote_sliced_kernel<<<grid, kVoteThreads, smem, h->stream>>>(h->d, slices, h->gplanes);
      sliced_scan_kernel<<<h->d.dir_count, 256, 0, h->stream>>>(h->d, h->gplanes);
      note_launches(2);
      H3D_CUDA(cudaGetLastError());
      return H3D_OK;
    }
    const int mode = g_stream_mode.load(std::memory_order_relaxed);
    if (mode != 2 && ablate == 0 && !h->has_negative) {   // unit weights: the TMA-staged streaming form
      bool launched = false;
      H3D_TRY(launch_stream(h, list, count_ptr, count_imm, dense, mode == 1, &launched));
      if (launched) return H3D_OK;
    }
    D = 2;
  }
  while (D > 1 && (size_t)D * plane_bytes > budget / 2) D >>= 1;
  const int per_sm = ((size_t)D * plane_bytes <= budget / 2) ? 2 : 1;
  const int grid = h->dev.sm_count * per_sm;
#define H3D_VOTE_CASE(DD)                                                                               \
  case DD:                                                                                              \
    if (ablate == 0) return launch_vote_t<DD, 0>(h, list, count_ptr, count_imm, dense, grid);          \
    if (ablate == 1) return launch_vote_t<DD, 1>(h, list, count_ptr, count_imm, dense, grid);          \
    return launch_vote_t<DD, 2>(h, list, count_ptr, count_imm, dense, grid);
  switch (D) {
    H3D_VOTE_CASE(1)
    H3D_VOTE_CASE(2)
    H3D_VOTE_CASE(4)
  }
#undef H3D_VOTE_CASE
  return fail(H3D_ERR_ARG, "bad D");
}

// Hough::Hough discretisation (lib/hough.cpp:32-40) into the device descriptor
static int configure_hough(h3d_hough *h, const double minP[3], const double maxP[3], double var_dx) {
  const Space &sp = *h->space;
  const HoughGeom g = hough_geometry(minP, maxP, var_dx, (unsigned long long)sp.num_b);   // h3d_frame.cuh: host == device
  if (g.rc == 1) return fail(H3D_ERR_RANGE, "num_x out of range");
  if (g.rc == 2)
    return fail(H3D_ERR_RANGE, "accumulator has >= 2^32 cells (the reference's unsigned argmax index wraps)");
  HoughDev &d = h->d;
  d.num_x = g.num_x;
  d.plane_cells = g.plane_cells;
  d.plane_stride = g.plane_stride;
  d.num_b = (int)sp.num_b;
  d.max_x = g.max_x; d.dx = g.dx;
  d.inv_s = g.inv_s; d.kmagic = g.kmagic;
  d.fast_ok = g.fast_ok;
  d.bounded = g.bounded;
  d.coef = sp.coef_dev;
  return H3D_OK;
}

// handle + per-direction buffers for directions [dir_begin, dir_end) of the current space
static int alloc_hough(size_t dir_begin, size_t dir_end, bool whole, h3d_hough **out) {
  auto sp = current_space();
  if (!sp) return fail(H3D_ERR_ARG, "h3d_init_space() has not been called");
  if (whole) { dir_begin = 0; dir_end = sp->num_b; }
  if (dir_begin > dir_end || dir_end > sp->num_b) return fail(H3D_ERR_ARG, "bad direction range");
  h3d_hough *h = new (std::nothrow) h3d_hough();
  if (!h) return fail(H3D_ERR_NOMEM, "host allocation failed");
  std::unique_ptr<h3d_hough, void (*)(h3d_hough *)> guard(h, h3d_destroy);
  h->space = sp;
  H3D_TRY(device_facts(&h->dev));
  HoughDev &d = h->d;
  d.dir_begin = (int)dir_begin;
  d.dir_count = (int)(dir_end - dir_begin);
  H3D_CUDA(cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
  const size_t nd = d.dir_count ? d.dir_count : 1;
  H3D_CUDA(cudaMalloc(&d.dvotes, nd * sizeof(unsigned)));
  H3D_CUDA(cudaMalloc(&d.dcell, nd * sizeof(unsigned)));
  H3D_CUDA(cudaMalloc(&d.dstale, nd));
  H3D_CUDA(cudaMemsetAsync(d.dvotes, 0, nd * sizeof(unsigned), h->stream));
  H3D_CUDA(cudaMemsetAsync(d.dcell, 0, nd * sizeof(unsigned), h->stream));
  H3D_CUDA(cudaMemsetAsync(d.dstale, 0, nd, h->stream));
  H3D_CUDA(cudaMalloc(&h->keys, 2 * sizeof(unsigned long long)));
  H3D_CUDA(cudaMalloc(&h->list, nd * sizeof(int)));
  H3D_CUDA(cudaMalloc(&h->count, 4 * sizeof(int)));
  H3D_TRY(ensure_points(h, 4096));
  *out = guard.release();
  return H3D_OK;
}

static int create_impl(const double minP[3], const double maxP[3], double var_dx, size_t dir_begin,
                       size_t dir_end, bool whole, h3d_hough **out) {
  if (!minP || !maxP || !out) return fail(H3D_ERR_ARG, "null argument");
  h3d_hough *h = nullptr;
  H3D_TRY(alloc_hough(dir_begin, dir_end, whole, &h));
  const int rc = configure_hough(h, minP, maxP, var_dx);
  if (rc != H3D_OK) { h3d_destroy(h); return rc; }
  h->all_dirty = false;   // an empty accumulator: every record is exactly zero
  *out = h;
  return H3D_OK;
}

extern "C" int h3d_create(const double minP[3], const double maxP[3], double dx, h3d_hough **out) {
  return create_impl(minP, maxP, dx, 0, 0, true, out);
}
extern "C" int h3d_create_sharded(const double minP[3], const double maxP[3], double dx, size_t dir_begin,
                                  size_t dir_end, h3d_hough **out) {
  return create_impl(minP, maxP, dx, dir_begin, dir_end, false, out);
}

extern "C" void h3d_destroy(h3d_hough *h) {
  if (!h) return;
  if (h->stream) cudaStreamSynchronize(h->stream);
  cudaFree(h->px); cudaFree(h->py); cudaFree(h->pz); cudaFree(h->w);
  cudaFree(h->stage); cudaFree(h->idx);
  cudaFree(h->d.dvotes); cudaFree(h->d.dcell); cudaFree(h->d.dstale);
  cudaFree(h->keys); cudaFree(h->list); cudaFree(h->count); cudaFree(h->gplanes); cudaFree(h->rplanes); cudaFree(h->rflag);
  if (h->stream) cudaStreamDestroy(h->stream);
  delete h;
}

extern "C" int h3d_info(const h3d_hough *h, size_t *num_x, size_t *num_b, double *dx, double *max_x) {
  if (!h) return fail(H3D_ERR_ARG, "null handle");
  if (num_x) *num_x = (size_t)h->d.num_x;
  if (num_b) *num_b = (size_t)h->d.num_b;
  if (dx) *dx = h->d.dx;
  if (max_x) *max_x = h->d.max_x;
  return H3D_OK;
}

extern "C" int h3d_add(h3d_hough *h, const double *xyz, size_t n) {
  if (!h || (!xyz && n)) return fail(H3D_ERR_ARG, "null argument");
  if (n == 0) return H3D_OK;
  if (h->d.n + n > 0x7FFFFFF0u) return fail(H3D_ERR_RANGE, "too many points");
  H3D_TRY(ensure_points(h, h->d.n + n));
  H3D_TRY(ensure_stage(h, n));
  H3D_CUDA(cudaMemcpyAsync(h->stage, xyz, n * 3 * sizeof(double), cudaMemcpyHostToDevice, h->stream));
  append_points_kernel<<<(unsigned)((n + 255) / 256), 256, 0, h->stream>>>(h->stage, (int)n, h->px, h->py,
                                                                          h->pz, h->w, h->d.n);
  H3D_CUDA(cudaGetLastError());
  H3D_CUDA(cudaStreamSynchronize(h->stream));   // xyz may be pageable and reused by the caller
  h->d.n += (int)n;
  h->all_dirty = true;   // new votes can raise any plane: rebuild on the next getLine
  return H3D_OK;
}

extern "C" int h3d_subtract(h3d_hough *h, const double *xyz, size_t n) {
  if (!h || (!xyz && n)) return fail(H3D_ERR_ARG, "null argument");
  if (n == 0) return H3D_OK;
  H3D_TRY(ensure_stage(h, n));
  H3D_TRY(ensure_points(h, h->d.n + n));
  H3D_CUDA(cudaMemcpyAsync(h->stage, xyz, n * 3 * sizeof(double), cudaMemcpyHostToDevice, h->stream));
  H3D_CUDA(cudaMemsetAsync(h->count, 0, 4 * sizeof(int), h->stream));
  const unsigned blocks = (unsigned)((n + 127) / 128);
  match_points_kernel<<<blocks, 128, 0, h->stream>>>(h->stage, (int)n, h->px, h->py, h->pz, h->w, h->d.n,
                                                     h->idx, h->count + 1);
  // records stay usable only while they describe the votes still present
  if (!h->all_dirty) {
    unvote_kernel<<<h->dev.sm_count * 4, 256, 0, h->stream>>>(h->d, h->idx, (int)n);
  }
  kill_points_kernel<<<blocks, 128, 0, h->stream>>>(h->idx, (int)n, h->w);
  append_negative_kernel<<<blocks, 128, 0, h->stream>>>(h->stage, h->idx, (int)n, h->px, h->py, h->pz, h->w,
                                                        h->d.n, h->count + 2);
  H3D_CUDA(cudaGetLastError());
  int counts[4];
  H3D_CUDA(cudaMemcpyAsync(counts, h->count, sizeof counts, cudaMemcpyDeviceToHost, h->stream));
  H3D_CUDA(cudaStreamSynchronize(h->stream));
  if (counts[1] > 0) {   // subtracting points that were never added: exact, but no bounds any more
    h->d.n += counts[2];
    h->has_negative = true;
    h->all_dirty = true;
  }
  if (!h->has_negative) {   // keep the live points dense (the resident vote kernel relies on it)
    compact_points_kernel<<<1, 1024, 0, h->stream>>>(h->px, h->py, h->pz, h->w, h->d.n);
    H3D_CUDA(cudaGetLastError());
    h->d.n -= (int)n - counts[1];
  }
  return H3D_OK;
}

// bring every per-direction record needed for the argmax up to date, then reduce
static int refresh_and_reduce(h3d_hough *h, unsigned long long *key_out) {
  if (h->d.dir_count == 0) { *key_out = 0; return H3D_OK; }
  const int blocks = (h->d.dir_count + 255) / 256;
  H3D_CUDA(cudaMemsetAsync(h->keys, 0, 2 * sizeof(unsigned long long), h->stream));
  if (h->all_dirty) {
    H3D_TRY(launch_vote(h, 0, 0, nullptr, nullptr, h->d.dir_count, nullptr));
    h->votes_cast += (uint64_t)h->d.n * h->d.dir_count;
    h->all_dirty = h->has_negative;
  } else {
    H3D_CUDA(cudaMemsetAsync(h->count, 0, sizeof(int), h->stream));
    argmax_kernel<<<blocks, 256, 0, h->stream>>>(h->d, 1, h->keys);
    worklist_kernel<<<blocks, 256, 0, h->stream>>>(h->d, h->keys, h->list, h->count);
    H3D_CUDA(cudaGetLastError());
    H3D_TRY(launch_vote(h, 0, 0, h->list, h->count, 0, nullptr));
  }
  argmax_kernel<<<blocks, 256, 0, h->stream>>>(h->d, 0, h->keys + 1);
  H3D_CUDA(cudaGetLastError());
  H3D_CUDA(cudaMemcpyAsync(key_out, h->keys + 1, sizeof(unsigned long long), cudaMemcpyDeviceToHost, h->stream));
  H3D_CUDA(cudaStreamSynchronize(h->stream));
  return H3D_OK;
}

extern "C" int h3d_line_from_index(const h3d_hough *h, uint64_t flat_index, double a[3], double b[3]) {
  if (!h || !a || !b) return fail(H3D_ERR_ARG, "null argument");
  // lib/hough.cpp:114-132, same integer/double conversions
  const size_t num_x = (size_t)h->d.num_x, num_b = (size_t)h->d.num_b;
  const double dx = h->d.dx, max_x = h->d.max_x;
  unsigned int index = (unsigned int)flat_index;
  if (num_x == 0 || num_b == 0) return fail(H3D_ERR_ARG, "empty accumulator");
  double x = (int)(index / (num_x * num_b));
  index -= (int)x * num_x * num_b;
  x = x * dx - max_x;
  double y = (int)index / num_b;
  index -= (int)y * num_b;
  y = y * dx - max_x;
  const double *dir = &h->space->dirs[3 * (size_t)index];
  const double bx = dir[0], by = dir[1], bz = dir[2];
  b[0] = bx; b[1] = by; b[2] = bz;
  a[0] = x * (1 - ((bx * bx) / (1 + bz))) - y * ((bx * by) / (1 + bz));
  a[1] = x * (-((bx * by) / (1 + bz))) + y * (1 - ((by * by) / (1 + bz)));
  a[2] = -x * bx - y * by;
  return H3D_OK;
}

extern "C" int h3d_get_line_key(h3d_hough *h, uint64_t *key) {
  if (!h || !key) return fail(H3D_ERR_ARG, "null argument");
  unsigned long long k = 0;
  H3D_TRY(refresh_and_reduce(h, &k));
  *key = k;
  return H3D_OK;
}

extern "C" int h3d_get_line(h3d_hough *h, double a[3], double b[3], unsigned *votes, uint64_t *flat_index) {
  if (!h || !a || !b) return fail(H3D_ERR_ARG, "null argument");
  unsigned long long k = 0;
  H3D_TRY(refresh_and_reduce(h, &k));
  const unsigned v = (unsigned)(k >> 32);
  // an all-zero accumulator (or an empty shard) reports index 0 like the reference
  const uint64_t flat = (h->d.dir_count && (v || h->d.dir_begin == 0)) ? 0xFFFFFFFFu - (unsigned)k : 0;
  if (votes) *votes = v;
  if (flat_index) *flat_index = v ? flat : 0;
  return h3d_line_from_index(h, v ? flat : 0, a, b);
}

extern "C" int h3d_download_votes(h3d_hough *h, unsigned *votes, size_t cap) {
  if (!h || !votes) return fail(H3D_ERR_ARG, "null argument");
  if (h->d.dir_begin != 0 || h->d.dir_count != h->d.num_b)
    return fail(H3D_ERR_ARG, "dense download needs an unsharded instance");
  const size_t cells = (size_t)h->d.plane_cells * h->d.num_b;
  if (cap < cells) return fail(H3D_ERR_ARG, "vote buffer too small");
  if (cells == 0) return H3D_OK;
  unsigned *dense = nullptr;
  H3D_CUDA(cudaMalloc(&dense, cells * sizeof(unsigned)));
  int rc = H3D_OK;
  do {
    if (cudaMemsetAsync(dense, 0, cells * sizeof(unsigned), h->stream) != cudaSuccess) { rc = H3D_ERR_CUDA; break; }
    rc = launch_vote(h, 2, 0, nullptr, nullptr, h->d.dir_count, dense);
    if (rc) break;
    h->all_dirty = h->has_negative;   // the same pass rebuilt every record
    if (cudaMemcpyAsync(votes, dense, cells * sizeof(unsigned), cudaMemcpyDeviceToHost, h->stream) != cudaSuccess ||
        cudaStreamSynchronize(h->stream) != cudaSuccess) { rc = H3D_ERR_CUDA; set_error("dense download failed"); }
  } while (0);
  cudaFree(dense);
  return rc;
}

extern "C" int h3d_bench_vote(h3d_hough *h, int reps, int ablate, float *us_per_launch) {
  int D = (ablate >> 8) & 0xFF;   // optional: directions per work item in bits 8..15
  ablate &= 0xFF;
  if (!h || !us_per_launch || reps < 1 || ablate < 0 || ablate > 2 || (D != 0 && D != 1 && D != 2 && D != 4))
    return fail(H3D_ERR_ARG, "bad argument");
  cudaEvent_t e0, e1;
  H3D_CUDA(cudaEventCreate(&e0));
  H3D_CUDA(cudaEventCreate(&e1));
  int rc = launch_vote(h, D, ablate, nullptr, nullptr, h->d.dir_count, nullptr);   // warm-up
  if (rc == H3D_OK) {
    cudaEventRecord(e0, h->stream);
    for (int r = 0; r < reps && rc == H3D_OK; ++r)
      rc = launch_vote(h, D, ablate, nullptr, nullptr, h->d.dir_count, nullptr);
    cudaEventRecord(e1, h->stream);
    if (cudaEventSynchronize(e1) != cudaSuccess) { rc = H3D_ERR_CUDA; set_error("vote bench failed"); }
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    *us_per_launch = ms * 1000.0f / reps;
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  if (ablate != 0) h->all_dirty = true;   // ablated launches leave meaningless records
  return rc;
}

// =====================================================================================
// The whole loop on the device: hough3dlines() of wrap.h:167-350 / lib/hough3dlines.cpp:200-317
// =====================================================================================
struct PrepOut {
  int n, pad;
  double mn[3], mx[3];
};

// wrap.h:175-189 NaN/Inf strip (ordered) + AoS -> SoA + getMinMax3D (lib/pointcloud.cpp:41-59).
// One CTA walks the cloud in chunks.  T = double (stride 3) or float (sensor frame, any stride).
template <typename T>
__global__ void __launch_bounds__(1024, 1)
prep_kernel(const T *__restrict__ in, int n_imm, const int *__restrict__ n_ptr, int stride, double *px, double *py,
            double *pz, unsigned *orig, PrepOut *out) {
  const int n = n_ptr ? *n_ptr : n_imm;
  __shared__ int warp_sum[32];
  __shared__ double red[6][32];
  __shared__ int base_s;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) base_s = 0;
  __syncthreads();
  double mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (int c0 = 0; c0 < n; c0 += 1024) {
    const int i = c0 + tid;
    double x = 0, y = 0, z = 0;
    bool keep = false;
    if (i < n) {
      x = (double)in[(size_t)i * stride]; y = (double)in[(size_t)i * stride + 1]; z = (double)in[(size_t)i * stride + 2];
      keep = isfinite(x) && isfinite(y) && isfinite(z);
    }
    const unsigned bal = __ballot_sync(0xFFFFFFFFu, keep);
    if (lane == 0) warp_sum[warp] = __popc(bal);
    __syncthreads();
    int off = base_s;
    for (int q = 0; q < warp; ++q) off += warp_sum[q];
    if (keep) {
      const int p = off + __popc(bal & ((1u << lane) - 1u));
      px[p] = x; py[p] = y; pz[p] = z; orig[p] = (unsigned)i;
      mn[0] = fmin(mn[0], x); mn[1] = fmin(mn[1], y); mn[2] = fmin(mn[2], z);
      mx[0] = fmax(mx[0], x); mx[1] = fmax(mx[1], y); mx[2] = fmax(mx[2], z);
    }
    __syncthreads();
    if (tid == 0) { int t = 0; for (int q = 0; q < 32; ++q) t += warp_sum[q]; base_s += t; }
    __syncthreads();
  }
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    for (int o = 16; o; o >>= 1) {
      mn[c] = fmin(mn[c], __shfl_xor_sync(0xFFFFFFFFu, mn[c], o));
      mx[c] = fmax(mx[c], __shfl_xor_sync(0xFFFFFFFFu, mx[c], o));
    }
    if (lane == 0) { red[c][warp] = mn[c]; red[3 + c][warp] = mx[c]; }
  }
  __syncthreads();
  if (warp == 0) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      double a = red[c][lane], b = red[3 + c][lane];
      for (int o = 16; o; o >>= 1) {
        a = fmin(a, __shfl_xor_sync(0xFFFFFFFFu, a, o));
        b = fmax(b, __shfl_xor_sync(0xFFFFFFFFu, b, o));
      }
      if (lane == 0) { out->mn[c] = base_s ? a : 0.0; out->mx[c] = base_s ? b : 0.0; }
    }
    if (lane == 0) out->n = base_s;
  }
}

// PointCloud::shiftToOrigin (lib/pointcloud.cpp:18-26): p = p - newshift, component-wise
__global__ void shift_points_kernel(double *px, double *py, double *pz, int *w, int n, double sx, double sy,
                                    double sz) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    px[i] = __dsub_rn(px[i], sx); py[i] = __dsub_rn(py[i], sy); pz[i] = __dsub_rn(pz[i], sz);
    w[i] = 1;
  }
}

namespace {

// ---- NCCL, bound at run time -------------------------------------------------------------------
// The direction-sharded loop needs one 8-byte MAX all-reduce per iteration, in stream order.  NCCL is
// opened with dlopen when a communicator is asked for (a process that already loaded a libnccl.so.2,
// e.g. through torch, gets that one), so libh3d.so itself does not depend on it.
}  // namespace
struct h3d_comm {
  ncclComm_t comm = nullptr;
  int rank = 0, world = 1;
};
namespace {
struct NcclApi {
  void *lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  const char *(*GetErrorString)(ncclResult_t) = nullptr;
};
NcclApi *nccl_api() {
  static std::mutex mu;
  static NcclApi api;
  std::lock_guard<std::mutex> lock(mu);
  if (api.lib) return &api;
  void *lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  if (!lib) return nullptr;
  api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(dlsym(lib, "ncclGetUniqueId"));
  api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(dlsym(lib, "ncclCommInitRank"));
  api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(dlsym(lib, "ncclCommDestroy"));
  api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(dlsym(lib, "ncclAllReduce"));
  api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(dlsym(lib, "ncclGetErrorString"));
  if (!api.GetUniqueId || !api.CommInitRank || !api.CommDestroy || !api.AllReduce || !api.GetErrorString) {
    dlclose(lib);
    return nullptr;
  }
  api.lib = lib;
  return &api;
}
int nccl_fail(NcclApi *api, ncclResult_t r, const char *what) {
  set_error(std::string(what) + ": " + (api ? api->GetErrorString(r) : "NCCL not available"));
  return H3D_ERR_NCCL;
}

struct FrameCtx {
  h3d_hough *hg = nullptr;
  unsigned granularity = ~0u;
  size_t cap = 0, iters_cap = 0, pool_cap = 0, in_bytes = 0;
  void *in_dev = nullptr;
  unsigned *orig = nullptr;
  unsigned char *flags = nullptr;
  int *yidx = nullptr;
  double *ypx = nullptr, *ypy = nullptr, *ypz = nullptr;
  h3d_iter *iters_dev = nullptr;
  unsigned *pool_dev = nullptr;
  LoopState *st_dev = nullptr;
  PrepOut *prep_dev = nullptr;
  unsigned long long *gbest = nullptr;   // fused loop: winner key per iteration
  size_t gbest_cap = 0;
  unsigned *sync_dev = nullptr;          // [0] grid barrier, [1] overflow flag
  unsigned *sync_host = nullptr;         // pinned
  unsigned long long *prof_dev = nullptr;   // measurement hook: fused-loop phase stamps (h3d_pipeline_profile)
  int prof_iters = 0, prof_ctas = 0, prof_grid = 0;   // allocated for prof_ctas CTAs; the last launch used prof_grid
  LoopState *st_host = nullptr;   // pinned
  PrepOut *prep_host = nullptr;   // pinned
  cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};   // start, prep done, first vote done, end, first vote start, loop kernels done
  cudaEvent_t ev_block = nullptr;   // blocking-sync event of frame_wait()
  float us[4] = {0, 0, 0, 0};
  bool timed = false;   // the last frame ran to the end (all events recorded)
  bool fused = false;   // the last frame ran in the fused persistent kernel
  int fused_ctas = 0;   // CTAs (= SMs) the fused kernel may use; 0 = all (h3d_pipeline_set_sm_share)
  // segment post-tests (h3d_pipeline_enable_segments)
  bool seg_on = false;
  h3d_segment_config seg_cfg{};
  h3d_pose seg_pose{};
  bool seg_pose_on = false;
  h3d_segment *seg_dev = nullptr, *seg_host = nullptr;   // iters_cap records each; host copy pinned
  size_t seg_cap = 0, seg_n = 0, seg_grid = 0;
  double seg_ns[3] = {0, 0, 0};          // what shiftToOrigin subtracted in the current frame
  const double *seg_ns_dev = nullptr;    // frame fast path: the same three doubles in device memory (FrameDev::ns)
  unsigned seg_pool_cap = 0;
  double *seg_scr = nullptr;             // 5 * seg_scr_n doubles
  size_t seg_scr_n = 0;
  const void *src = nullptr;             // the cloud the pool indexes (set by the caller of frame_run)
  h3d_comm *comm = nullptr;              // direction-sharded loop (h3d_hough3dlines_sharded): this frame only
  // large clouds (>= kBigCloud points): chunked multi-CTA compaction / removal (h3d_loop.cuh)
  bool big = false;
  size_t big_n = 0;                      // points of the frame when the loop started
  int *ccount = nullptr, *coff = nullptr;   // cap / kChunk + 2 entries each
  double *sx = nullptr, *sy = nullptr, *sz = nullptr;
  unsigned *so = nullptr;
  size_t scr_cap = 0;
  int src_stride = 3, src_is_float = 0;

  ~FrameCtx() { release(); }
  void release() {
    if (hg) { h3d_destroy(hg); hg = nullptr; }
    cudaFree(ccount); cudaFree(coff); cudaFree(sx); cudaFree(sy); cudaFree(sz); cudaFree(so);
    ccount = coff = nullptr; sx = sy = sz = nullptr; so = nullptr; scr_cap = 0;
    cudaFree(in_dev); cudaFree(orig); cudaFree(flags); cudaFree(yidx);
    cudaFree(ypx); cudaFree(ypy); cudaFree(ypz);
    cudaFree(iters_dev); cudaFree(pool_dev); cudaFree(st_dev); cudaFree(prep_dev);
    cudaFree(gbest); cudaFree(sync_dev); cudaFreeHost(sync_host); cudaFree(prof_dev);
    cudaFree(seg_dev); cudaFreeHost(seg_host); cudaFree(seg_scr);
    seg_dev = seg_host = nullptr; seg_scr = nullptr; seg_cap = seg_n = seg_scr_n = 0;
    prof_dev = nullptr; prof_iters = prof_ctas = 0;
    gbest = nullptr; sync_dev = nullptr; sync_host = nullptr; gbest_cap = 0;
    cudaFreeHost(st_host); cudaFreeHost(prep_host);
    for (auto &e : ev) if (e) { cudaEventDestroy(e); e = nullptr; }
    if (ev_block) { cudaEventDestroy(ev_block); ev_block = nullptr; }
    in_dev = nullptr; orig = nullptr; flags = nullptr; yidx = nullptr; ypx = ypy = ypz = nullptr;
    iters_dev = nullptr; pool_dev = nullptr; st_dev = nullptr; prep_dev = nullptr;
    st_host = nullptr; prep_host = nullptr;
    cap = iters_cap = pool_cap = in_bytes = 0; granularity = ~0u;
  }
};

// how a host thread waits for its frame's stream: 0 = cudaStreamSynchronize (spins: lowest
// latency), 1 = a blocking event (the thread sleeps: for more pipelines than host cores),
// 2 = event polling with sched_yield between polls
std::atomic<int> g_wait_mode{0};
std::atomic<int> g_replay_threads{0};   // h3d_set_replay_threads

cudaError_t frame_wait(FrameCtx &f) {
  cudaStream_t s = f.hg->stream;
  if (g_wait_mode.load(std::memory_order_relaxed) == 0) return cudaStreamSynchronize(s);
  if (!f.ev_block) {
    const cudaError_t e = cudaEventCreateWithFlags(&f.ev_block, cudaEventBlockingSync | cudaEventDisableTiming);
    if (e != cudaSuccess) return e;
  }
  cudaError_t e = cudaEventRecord(f.ev_block, s);
  if (e != cudaSuccess) return e;
  if (g_wait_mode.load(std::memory_order_relaxed) == 1) return cudaEventSynchronize(f.ev_block);
  // mode 2: poll and give the core away between polls -- as quick as spinning while cores are
  // free, and the other pipelines' threads get to run when they are not
  while ((e = cudaEventQuery(f.ev_block)) == cudaErrorNotReady) std::this_thread::yield();
  return e;
}

int frame_ensure(FrameCtx &f, unsigned g, size_t n, size_t in_bytes, size_t max_iters, size_t pool_cap) {
  if (!f.hg || f.granularity != g) {
    if (f.hg) { h3d_destroy(f.hg); f.hg = nullptr; }
    H3D_TRY(h3d_init_space(g));
    H3D_TRY(alloc_hough(0, 0, true, &f.hg));
    f.granularity = g;
  }
  if (!f.st_dev) {
    H3D_CUDA(cudaMalloc(&f.st_dev, sizeof(LoopState)));
    H3D_CUDA(cudaMalloc(&f.prep_dev, sizeof(PrepOut)));
    H3D_CUDA(cudaMallocHost(&f.st_host, sizeof(LoopState)));
    H3D_CUDA(cudaMallocHost(&f.prep_host, sizeof(PrepOut)));
    H3D_CUDA(cudaMalloc(&f.sync_dev, 4 * sizeof(unsigned)));
    H3D_CUDA(cudaMallocHost(&f.sync_host, 4 * sizeof(unsigned)));
    for (auto &e : f.ev) H3D_CUDA(cudaEventCreate(&e));
  }
  if (n > f.cap) {
    size_t cap = f.cap ? f.cap : 8192;
    while (cap < n) cap *= 2;
    cudaFree(f.orig); cudaFree(f.flags); cudaFree(f.yidx); cudaFree(f.ypx); cudaFree(f.ypy); cudaFree(f.ypz);
    f.orig = nullptr; f.flags = nullptr; f.yidx = nullptr; f.ypx = f.ypy = f.ypz = nullptr; f.cap = 0;
    H3D_CUDA(cudaMalloc(&f.orig, cap * sizeof(unsigned)));
    H3D_CUDA(cudaMalloc(&f.flags, cap));
    H3D_CUDA(cudaMalloc(&f.yidx, cap * sizeof(int)));
    H3D_CUDA(cudaMalloc(&f.ypx, cap * sizeof(double)));
    H3D_CUDA(cudaMalloc(&f.ypy, cap * sizeof(double)));
    H3D_CUDA(cudaMalloc(&f.ypz, cap * sizeof(double)));
    cudaFree(f.ccount); cudaFree(f.coff); f.ccount = f.coff = nullptr;
    H3D_CUDA(cudaMalloc(&f.ccount, (cap / kChunk + 2) * sizeof(int)));
    H3D_CUDA(cudaMalloc(&f.coff, (cap / kChunk + 2) * sizeof(int)));
    f.cap = cap;
  }
  if (n >= (size_t)kBigCloud && f.scr_cap < f.cap) {
    cudaFree(f.sx); cudaFree(f.sy); cudaFree(f.sz); cudaFree(f.so);
    f.sx = f.sy = f.sz = nullptr; f.so = nullptr; f.scr_cap = 0;
    H3D_CUDA(cudaMalloc(&f.sx, f.cap * sizeof(double)));
    H3D_CUDA(cudaMalloc(&f.sy, f.cap * sizeof(double)));
    H3D_CUDA(cudaMalloc(&f.sz, f.cap * sizeof(double)));
    H3D_CUDA(cudaMalloc(&f.so, f.cap * sizeof(unsigned)));
    f.scr_cap = f.cap;
  }
  H3D_TRY(ensure_points(f.hg, n ? n : 1));
  if (in_bytes > f.in_bytes) {
    cudaFree(f.in_dev); f.in_dev = nullptr; f.in_bytes = 0;
    H3D_CUDA(cudaMalloc(&f.in_dev, in_bytes));
    f.in_bytes = in_bytes;
  }
  if (max_iters > f.iters_cap) {
    cudaFree(f.iters_dev); f.iters_dev = nullptr; f.iters_cap = 0;
    H3D_CUDA(cudaMalloc(&f.iters_dev, max_iters * sizeof(h3d_iter)));
    f.iters_cap = max_iters;
  }
  if (pool_cap > f.pool_cap) {
    cudaFree(f.pool_dev); f.pool_dev = nullptr; f.pool_cap = 0;
    H3D_CUDA(cudaMalloc(&f.pool_dev, pool_cap * sizeof(unsigned)));
    f.pool_cap = pool_cap;
  }
  return H3D_OK;
}

// one iteration of wrap.h:237-342 as a train of kernels on the frame's stream
int enqueue_iteration(FrameCtx &f, const HoughDev &hd, const LoopBuffers &lb) {
  h3d_hough *h = f.hg;
  cudaStream_t s = h->stream;
  const int sm = h->dev.sm_count;
  const int dblocks = (hd.dir_count + 255) / 256;
  loop_begin_kernel<<<1, 1, 0, s>>>(lb);
  loop_unvote_kernel<<<sm * 4, 256, 0, s>>>(hd, lb);                    // hough->subtract(Y)   wrap.h:241
  loop_argmax_kernel<<<dblocks, 256, 0, s>>>(hd, lb, 1);
  loop_worklist_kernel<<<dblocks, 256, 0, s>>>(hd, lb);
  H3D_CUDA(cudaGetLastError());
  H3D_TRY(launch_vote(h, 0, 0, lb.list, &lb.st->work_count, 0, nullptr));
  loop_argmax_kernel<<<dblocks, 256, 0, s>>>(hd, lb, 0);                // hough->getLine       wrap.h:243
  if (f.comm) {   // direction shards: the global winner is the MAX of the packed keys (global flat indices)
    NcclApi *api = nccl_api();
    const ncclResult_t r = api->AllReduce(&lb.st->best_key, &lb.st->best_key, 1, ncclUint64, ncclMax, f.comm->comm, s);
    if (r != ncclSuccess) return nccl_fail(api, r, "ncclAllReduce");
  }
  loop_line_kernel<<<1, 32, 0, s>>>(hd, lb);
  if (f.big) {   // large cloud: every 1024-point chunk is a CTA's work item (same order, same results)
    const int nchunks = (int)((f.big_n + kChunk - 1) / kChunk);
    const int cgrid = nchunks < sm * 2 ? nchunks : sm * 2;
    big_mask_kernel<<<cgrid, kChunk, 0, s>>>(hd, lb);                   // pointsCloseToLine    wrap.h:245
    big_scan_kernel<<<1, 1024, 0, s>>>(lb, 1);
    big_scatter_kernel<<<cgrid, kChunk, 0, s>>>(lb);
    loop_lsq_kernel<<<1, 192, 0, s>>>(hd, lb, 1);                        // orthogonal_LSQ       wrap.h:247
    big_mask_kernel<<<cgrid, kChunk, 0, s>>>(hd, lb);                   // pointsCloseToLine    wrap.h:250
    big_scan_kernel<<<1, 1024, 0, s>>>(lb, 2);
    big_scatter_kernel<<<cgrid, kChunk, 0, s>>>(lb);
    loop_lsq_kernel<<<1, 192, 0, s>>>(hd, lb, 2);                        // minvotes, LSQ        wrap.h:251-259
    big_remove_kernel<<<cgrid, kChunk, 0, s>>>(lb);                     // removePoints         wrap.h:339
    big_copyback_kernel<<<sm * 4, 256, 0, s>>>(lb);
    big_finish_kernel<<<1, 1, 0, s>>>(lb);                              // while (...)          wrap.h:341-342
    note_launches(17);
    H3D_CUDA(cudaGetLastError());
    return H3D_OK;
  }
  loop_mask_kernel<<<sm * 2, 256, 0, s>>>(hd, lb);                      // pointsCloseToLine    wrap.h:245
  loop_compact_kernel<<<1, 1024, 0, s>>>(hd, lb, 1);
  loop_lsq_kernel<<<1, 192, 0, s>>>(hd, lb, 1);                          // orthogonal_LSQ       wrap.h:247
  loop_mask_kernel<<<sm * 2, 256, 0, s>>>(hd, lb);                      // pointsCloseToLine    wrap.h:250
  loop_compact_kernel<<<1, 1024, 0, s>>>(hd, lb, 2);
  loop_lsq_kernel<<<1, 192, 0, s>>>(hd, lb, 2);                          // minvotes, LSQ        wrap.h:251-259
  loop_remove_kernel<<<1, 1024, 0, s>>>(lb);                            // removePoints, while  wrap.h:339-342
  note_launches(13);   // + the vote kernel counted in launch_vote
  H3D_CUDA(cudaGetLastError());
  return H3D_OK;
}

int segments_ensure(FrameCtx &f, size_t n_rec, size_t n_points) {
  if (n_rec > f.seg_cap) {
    cudaFree(f.seg_dev); cudaFreeHost(f.seg_host); f.seg_dev = f.seg_host = nullptr; f.seg_cap = 0;
    const size_t cap = n_rec > f.iters_cap ? n_rec : f.iters_cap;
    H3D_CUDA(cudaMalloc(&f.seg_dev, cap * sizeof(h3d_segment)));
    H3D_CUDA(cudaMallocHost(&f.seg_host, cap * sizeof(h3d_segment)));
    f.seg_cap = cap;
  }
  if (n_points > f.seg_scr_n) {
    cudaFree(f.seg_scr); f.seg_scr = nullptr; f.seg_scr_n = 0;
    const size_t cap = n_points > f.cap ? n_points : f.cap;
    H3D_CUDA(cudaMalloc(&f.seg_scr, 5 * cap * sizeof(double)));
    f.seg_scr_n = cap;
  }
  return H3D_OK;
}

// wrap.h:261-337 on every accepted line: enqueued right behind the loop's kernels (no host round
// trip in between), one CTA per iteration record the loop may have written
int enqueue_segments(FrameCtx &f, size_t grid) {
  if (!f.seg_on || grid == 0) return H3D_OK;
  if (grid > f.iters_cap) grid = f.iters_cap;
  H3D_TRY(segments_ensure(f, grid, (size_t)f.hg->d.n));
  SegArgs sa;
  sa.iters = f.iters_dev; sa.n_iters = &f.st_dev->n_iters; sa.pool = f.pool_dev; sa.pool_cap = f.seg_pool_cap;
  sa.src = f.src; sa.stride = f.src_stride; sa.src_is_float = f.src_is_float;
  for (int c = 0; c < 3; ++c) { sa.ns[c] = f.seg_ns[c]; sa.shift[c] = f.seg_ns[c]; }   // X.shift = 0 + newshift
  sa.ns_dev = f.seg_ns_dev;
  sa.cfg = f.seg_cfg; sa.scr = f.seg_scr; sa.scr_n = f.seg_scr_n; sa.out = f.seg_dev;
  sa.pose = f.seg_pose; sa.pose_on = f.seg_pose_on ? 1 : 0;
  segment_tests_kernel<<<(unsigned)grid, kSegThreads, 0, f.hg->stream>>>(sa);
  note_launches(1);
  H3D_CUDA(cudaGetLastError());
  f.seg_grid = grid;
  return H3D_OK;
}

static bool g_use_fused = true;
static std::atomic<int> g_debug_poison{0};   // h3d_set_debug_poison: records start from 0xFFFFFFFF every frame
static bool g_big_kernels = true;   // h3d_set_big_kernels: chunked K3 / removal kernels for large clouds

int launch_fused(FrameCtx &f, const FusedArgs &args, size_t smem, int grid) {
  H3D_TRY(ensure_dyn_smem(reinterpret_cast<const void *>(fused_loop_kernel), smem, f.hg->dev.smem_optin));
  void *params[] = {const_cast<FusedArgs *>(&args)};
  H3D_CUDA(cudaLaunchCooperativeKernel(reinterpret_cast<const void *>(fused_loop_kernel), dim3(grid),
                                       dim3(kResThreads), params, smem, f.hg->stream));
  note_launches(1);
  return H3D_OK;
}

// largest x with sqrt_rn(x) <= dx: pointsCloseToLine's `sqrt(nn) <= dx` becomes `nn <= x`
// (sqrt_rn is monotonic, so the two tests accept exactly the same nn)
static double sqrt_le_threshold(double dx) {
  if (!(dx >= 0.0)) return -1.0;
  double c = dx * dx;
  if (std::isinf(c)) return DBL_MAX;
  while (c > 0.0 && sqrt(c) > dx) c = nextafter(c, 0.0);
  for (;;) {
    const double up = nextafter(c, INFINITY);
    if (std::isinf(up) || sqrt(up) > dx) break;
    c = up;
  }
  return c;
}

// returns H3D_OK with *launched = false when the frame does not qualify for the fused kernel
int try_launch_fused(FrameCtx &f, const LoopBuffers &lb, int opt_nlines, double opt_dx, bool *launched) {
  *launched = false;
  h3d_hough *h = f.hg;
  const HoughDev &hd = h->d;
  if (!g_use_fused || f.comm || !resident_ok(h) || hd.n < 1 || hd.dir_begin != 0 || !hd.bounded) return H3D_OK;
  if (hd.plane_stride >= 65536u || hd.n >= 65535) return H3D_OK;   // re-vote keys pack (votes, cell) in 16 + 16 bits
  const int groups = (hd.dir_count + kResG - 1) / kResG;
  const int grid = (f.fused_ctas > 0 && f.fused_ctas < h->dev.sm_count) ? f.fused_ctas : h->dev.sm_count;
  const int n_own = ((groups + grid - 1) / grid) * kResG;
  const size_t plane_bytes = (size_t)hd.plane_stride * sizeof(unsigned);
  const size_t tables = (size_t)n_own * 16;
  cudaFuncAttributes fa;
  H3D_CUDA(cudaFuncGetAttributes(&fa, fused_loop_kernel));
  if (h->dev.smem_optin < fa.sharedSizeBytes + 1024) return H3D_OK;
  const size_t smem = ((h->dev.smem_optin - fa.sharedSizeBytes - 256) / 16) * 16;   // everything the SM offers
  if (smem < kFusedHdr + 1024) return H3D_OK;
  const size_t body = smem - kFusedHdr;            // behind the kernel's HoughDev + FusedShared header
  // first vote: 2*G planes + pad; loop: nb planes + pad + the live points (24 B) + the inlier list (4 B) + its t values (4 B)
  if (tables + 2 * kResG * plane_bytes + 16 > body) return H3D_OK;
  const size_t n_cap = ((size_t)hd.n + 1) & ~(size_t)1;
  if (tables + 16 + 32 * n_cap + 2 * plane_bytes > body) return H3D_OK;
  size_t nb = (body - tables - 16 - 32 * n_cap) / plane_bytes;
  if (nb > 2 * kResG) nb = 2 * kResG;
  nb &= ~(size_t)1;
  const size_t iters = (size_t)(opt_nlines > 0 ? opt_nlines : hd.n) + 2;
  if (2 * iters > f.gbest_cap) {
    cudaFree(f.gbest); f.gbest = nullptr; f.gbest_cap = 0;
    H3D_CUDA(cudaMalloc(&f.gbest, 2 * iters * sizeof(unsigned long long)));
    f.gbest_cap = 2 * iters;
  }
  cudaStream_t s = h->stream;
  H3D_CUDA(cudaMemsetAsync(f.gbest, 0, 2 * iters * sizeof(unsigned long long), s));
  H3D_CUDA(cudaMemsetAsync(f.sync_dev, 0, 4 * sizeof(unsigned), s));
  FusedArgs a;
  a.h = hd; a.h.n_dev = nullptr;
  a.st = lb.st; a.iters = lb.iters; a.pool = lb.pool; a.orig = lb.orig; a.dirs = lb.dirs;
  a.xchg = f.gbest; a.bar = f.sync_dev;
  a.n_own_max = n_own;
  a.nb_planes = (int)nb;
  a.n_cap = (int)n_cap;
  a.near_t = sqrt_le_threshold(opt_dx);
  a.cell_off = hd.max_x / hd.dx + 0.5;
  a.prof = f.prof_dev; a.prof_iters = f.prof_iters;
  a.fd = nullptr;
  f.prof_grid = grid;
  if (f.prof_dev)
    H3D_CUDA(cudaMemsetAsync(f.prof_dev, 0, (size_t)f.prof_iters * f.prof_ctas * kProfSlots * sizeof(unsigned long long), s));
  const int need = (hd.n + kResThreads - 1) / kResThreads;
  // Hough::add: every plane once, points in registers (its own kernel: 128 registers per thread
  // for the vote alone); then the loop kernel picks the records up from global memory
  int rc;
  H3D_CUDA(cudaEventRecord(f.ev[4], s));
  switch (need) {   // one instantiation per slot count: no register or issue slot spent on empty slots
    case 1: rc = launch_first_vote_t<1>(h, grid); break;
    case 2: rc = launch_first_vote_t<2>(h, grid); break;
    case 3: rc = launch_first_vote_t<3>(h, grid); break;
    case 4: rc = launch_first_vote_t<4>(h, grid); break;
    case 5: rc = launch_first_vote_t<5>(h, grid); break;
    case 6: rc = launch_first_vote_t<6>(h, grid); break;
    case 7: rc = launch_first_vote_t<7>(h, grid); break;
    case 8: rc = launch_first_vote_t<8>(h, grid); break;
    case 9: rc = launch_first_vote_t<9>(h, grid); break;
    case 10: rc = launch_first_vote_t<10>(h, grid); break;
    case 11: rc = launch_first_vote_t<11>(h, grid); break;
    default: rc = launch_first_vote_t<12>(h, grid); break;
  }
  if (rc != H3D_OK) return rc;
  H3D_CUDA(cudaEventRecord(f.ev[2], s));   // "first vote done" (h3d_pipeline_timing [1])
  rc = launch_fused(f, a, smem, grid);
  if (rc != H3D_OK) return rc;
  H3D_CUDA(cudaEventRecord(f.ev[5], s));
  H3D_TRY(enqueue_segments(f, opt_nlines > 0 ? (size_t)opt_nlines : f.iters_cap));
  H3D_CUDA(cudaMemcpyAsync(f.sync_host, f.sync_dev, 4 * sizeof(unsigned), cudaMemcpyDeviceToHost, s));
  H3D_CUDA(cudaMemcpyAsync(f.st_host, f.st_dev, sizeof(LoopState), cudaMemcpyDeviceToHost, s));
  *launched = true;
  return H3D_OK;
}

// everything after the points are on the device (unshifted SoA + orig + PrepOut): bbox checks,
// Hough ctor, first vote, loop, results.  timing events: ev[1] after prep, ev[2] after the first
// vote, ev[3] at the end.
int frame_run(FrameCtx &f, int opt_minvotes, int opt_nlines, double opt_dx, int mode, h3d_iter *iters,
              size_t max_iters, uint32_t *pool, size_t pool_cap, h3d_result *res) {
  h3d_hough *h = f.hg;
  cudaStream_t s = h->stream;
  H3D_CUDA(cudaMemcpyAsync(f.prep_host, f.prep_dev, sizeof(PrepOut), cudaMemcpyDeviceToHost, s));
  H3D_CUDA(cudaEventRecord(f.ev[1], s));
  H3D_CUDA(frame_wait(f));
  const PrepOut P = *f.prep_host;
  const size_t n = (size_t)P.n;
  res->n_input = (uint32_t)n;
  res->num_b = h->space->num_b;
  if (mode == 1) {                                       // lib/hough3dlines.cpp:170-173, 196-199
    if (opt_minvotes < 2) opt_minvotes = 2;
    if (n < 2) { res->rc = 1; return H3D_OK; }
  }
  const double ex = P.mx[0] - P.mn[0], ey = P.mx[1] - P.mn[1], ez = P.mx[2] - P.mn[2];
  const double d = sqrt((ex * ex) + (ey * ey) + (ez * ez));   // (maxP-minP).norm(), wrap.h:201
  res->d = d;
  if (d == 0.0) { res->rc = 1; return H3D_OK; }         // wrap.h:202-205
  double ns[3], minPs[3], maxPs[3];
  for (int c = 0; c < 3; ++c) {                          // shiftToOrigin + second bbox, wrap.h:206-207
    ns[c] = (P.mn[c] + P.mx[c]) / 2.0;
    res->shift[c] = 0.0 + ns[c];
    minPs[c] = P.mn[c] - ns[c];                          // min(p - s) == min(p) - s: rounding is monotonic
    maxPs[c] = P.mx[c] - ns[c];
  }
  if (mode == 1 && opt_dx == 0.0) opt_dx = d / 64.0;     // lib/hough3dlines.cpp:211-213
  else if (opt_dx >= d) { res->rc = 1; return H3D_OK; }  // wrap.h:209-212
  H3D_TRY(configure_hough(h, minPs, maxPs, opt_dx));     // new Hough(...), wrap.h:220
  HoughDev &hd = h->d;
  res->dx = hd.dx; res->max_x = hd.max_x; res->num_x = (uint64_t)hd.num_x;
  hd.dir_begin = 0; hd.dir_count = hd.num_b;
  if (f.comm) {   // this rank's slice of the direction sphere (every rank holds all points)
    const size_t nb = (size_t)hd.num_b, r = (size_t)f.comm->rank, w = (size_t)f.comm->world;
    hd.dir_begin = (int)(nb * r / w);
    hd.dir_count = (int)(nb * (r + 1) / w) - hd.dir_begin;
  }
  hd.n = (int)n;
  hd.n_dev = nullptr;
  h->has_negative = false;
  shift_points_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(h->px, h->py, h->pz, h->w, (int)n, ns[0], ns[1], ns[2]);
  note_launches(2);   // + the prep kernel of the caller

  LoopState st;
  memset(&st, 0, sizeof st);
  st.opt_dx = opt_dx;
  for (int c = 0; c < 3; ++c) st.shift[c] = res->shift[c];
  st.minvotes = opt_minvotes; st.nlines_max = opt_nlines;
  st.max_iters = (int)(max_iters < f.iters_cap ? max_iters : f.iters_cap);
  // the post-tests read the inlier lists on the device: keep them all even if the host takes fewer
  st.pool_cap = (unsigned)((pool_cap < f.pool_cap && !f.seg_on) ? pool_cap : f.pool_cap);
  st.n_alive = (int)n;
  for (int c = 0; c < 3; ++c) f.seg_ns[c] = ns[c];
  f.seg_pool_cap = st.pool_cap; f.seg_grid = 0;
  *f.st_host = st;
  H3D_CUDA(cudaMemcpyAsync(f.st_dev, f.st_host, sizeof(LoopState), cudaMemcpyHostToDevice, s));

  LoopBuffers lb;
  lb.st = f.st_dev; lb.iters = f.iters_dev; lb.pool = f.pool_dev; lb.orig = f.orig;
  lb.flags = f.flags; lb.yidx = f.yidx; lb.ypx = f.ypx; lb.ypy = f.ypy; lb.ypz = f.ypz;
  lb.px = h->px; lb.py = h->py; lb.pz = h->pz; lb.list = h->list; lb.dirs = h->space->dirs_dev;
  lb.ccount = f.ccount; lb.coff = f.coff; lb.sx = f.sx; lb.sy = f.sy; lb.sz = f.sz; lb.so = f.so;
  f.big = n >= (size_t)kBigCloud && f.scr_cap >= n && g_big_kernels;
  f.big_n = n;
  int rc = H3D_OK;
  if (g_debug_poison.load(std::memory_order_relaxed)) {
    // debug aid: a record the first vote fails to write keeps 0xFFFFFFFF votes and wins every argmax,
    // instead of silently holding the previous frame's maximum for that direction
    const size_t nd = (size_t)(hd.dir_count > 0 ? hd.dir_count : 1);
    H3D_CUDA(cudaMemsetAsync(hd.dvotes, 0xFF, nd * sizeof(unsigned), s));
    H3D_CUDA(cudaMemsetAsync(hd.dcell, 0xFF, nd * sizeof(unsigned), s));
  }
  // fast path: first vote + the whole loop in one persistent cooperative kernel
  bool fused = false;
  H3D_TRY(try_launch_fused(f, lb, opt_nlines, opt_dx, &fused));
  if (fused) {
    H3D_CUDA(frame_wait(f));
    if (f.sync_host[1]) {   // an inlier set larger than the shared-memory scratch: redo with the general loop
      fused = false;
      *f.st_host = st;
      H3D_CUDA(cudaMemcpyAsync(f.st_dev, f.st_host, sizeof(LoopState), cudaMemcpyHostToDevice, s));
    }
  }
  f.fused = fused;
  if (!fused) {
  // hough->add(X), wrap.h:228: every plane once
  H3D_CUDA(cudaEventRecord(f.ev[4], s));
  H3D_TRY(launch_vote(h, 0, 0, nullptr, nullptr, hd.dir_count, nullptr));
  H3D_CUDA(cudaEventRecord(f.ev[2], s));

  hd.n_dev = &f.st_dev->n_alive;                         // the loop shrinks X on the device
  if (opt_nlines > 0) {
    for (int it = 0; it < opt_nlines && rc == H3D_OK; ++it) rc = enqueue_iteration(f, hd, lb);
  } else {                                               // unbounded: look at `done` every few iterations
    for (;;) {
      for (int it = 0; it < 8 && rc == H3D_OK; ++it) rc = enqueue_iteration(f, hd, lb);
      if (rc != H3D_OK) break;
      if (cudaMemcpyAsync(f.st_host, f.st_dev, sizeof(LoopState), cudaMemcpyDeviceToHost, s) != cudaSuccess ||
          frame_wait(f) != cudaSuccess) { rc = fail(H3D_ERR_CUDA, "loop state read-back failed"); break; }
      if (f.st_host->done) break;
    }
  }
  if (rc == H3D_OK) {
    H3D_CUDA(cudaEventRecord(f.ev[5], s));
    H3D_TRY(enqueue_segments(f, opt_nlines > 0 ? (size_t)opt_nlines : f.iters_cap));
    H3D_CUDA(cudaMemcpyAsync(f.st_host, f.st_dev, sizeof(LoopState), cudaMemcpyDeviceToHost, s));
    H3D_CUDA(frame_wait(f));
  }
  }
  hd.n_dev = nullptr;
  if (rc != H3D_OK) return rc;
  const LoopState &fin = *f.st_host;
  res->n_iters = (uint32_t)fin.n_iters; res->n_lines = (uint32_t)fin.n_lines; res->pool_used = fin.pool_used;
  res->votes_cast = (uint64_t)n * hd.dir_count + fin.votes_cast;
  const size_t n_rec = (size_t)fin.n_iters < (size_t)st.max_iters ? (size_t)fin.n_iters : (size_t)st.max_iters;
  if (n_rec && iters) H3D_CUDA(cudaMemcpyAsync(iters, f.iters_dev, n_rec * sizeof(h3d_iter), cudaMemcpyDeviceToHost, s));
  const size_t n_pool = (size_t)fin.pool_used < pool_cap ? (size_t)fin.pool_used : pool_cap;
  if (n_pool && pool) H3D_CUDA(cudaMemcpyAsync(pool, f.pool_dev, n_pool * sizeof(unsigned), cudaMemcpyDeviceToHost, s));
  f.seg_n = 0;
  if (f.seg_on && n_rec) {                                // records of enqueue_segments()
    const size_t m = n_rec < f.seg_grid ? n_rec : f.seg_grid;
    H3D_CUDA(cudaMemcpyAsync(f.seg_host, f.seg_dev, m * sizeof(h3d_segment), cudaMemcpyDeviceToHost, s));
    f.seg_n = m;
  }
  H3D_CUDA(cudaEventRecord(f.ev[3], s));
  H3D_CUDA(frame_wait(f));
  f.timed = true;
  hd.n = 0;
  return H3D_OK;
}

}  // namespace

static thread_local float g_last_call_us[4] = {0, 0, 0, 0};   // h3d_hough3dlines_timing

extern "C" int h3d_hough3dlines(const double *xyz, size_t n, int granularity, double opt_dx, int opt_minvotes,
                                int opt_nlines, int mode, h3d_iter *iters, size_t max_iters, uint32_t *pool,
                                size_t pool_cap, h3d_result *res) {
  return h3d_hough3dlines_sharded(xyz, n, granularity, opt_dx, opt_minvotes, opt_nlines, mode, nullptr, iters,
                                  max_iters, pool, pool_cap, res);
}

extern "C" int h3d_comm_unique_id(unsigned char id[H3D_COMM_ID_BYTES]) {
  if (!id) return fail(H3D_ERR_ARG, "bad argument");
  NcclApi *api = nccl_api();
  if (!api) return nccl_fail(nullptr, ncclSystemError, "dlopen(libnccl.so.2)");
  ncclUniqueId u;
  const ncclResult_t r = api->GetUniqueId(&u);
  if (r != ncclSuccess) return nccl_fail(api, r, "ncclGetUniqueId");
  static_assert(sizeof u == H3D_COMM_ID_BYTES, "ncclUniqueId size");
  memcpy(id, &u, sizeof u);
  return H3D_OK;
}

extern "C" int h3d_comm_create(const unsigned char id[H3D_COMM_ID_BYTES], int rank, int world, h3d_comm **out) {
  if (!id || !out || world < 1 || rank < 0 || rank >= world) return fail(H3D_ERR_ARG, "bad argument");
  NcclApi *api = nccl_api();
  if (!api) return nccl_fail(nullptr, ncclSystemError, "dlopen(libnccl.so.2)");
  int dev_count = 0;
  if (cudaGetDeviceCount(&dev_count) != cudaSuccess || dev_count == 0) { cudaGetLastError(); return fail(H3D_ERR_CUDA, "no CUDA device"); }
  ncclUniqueId u;
  memcpy(&u, id, sizeof u);
  h3d_comm *c = new (std::nothrow) h3d_comm();
  if (!c) return fail(H3D_ERR_NOMEM, "host allocation failed");
  c->rank = rank; c->world = world;
  const ncclResult_t r = api->CommInitRank(&c->comm, world, u, rank);
  if (r != ncclSuccess) { delete c; return nccl_fail(api, r, "ncclCommInitRank"); }
  *out = c;
  return H3D_OK;
}

extern "C" void h3d_comm_destroy(h3d_comm *c) {
  if (!c) return;
  NcclApi *api = nccl_api();
  if (api && c->comm) api->CommDestroy(c->comm);
  delete c;
}

extern "C" int h3d_hough3dlines_sharded(const double *xyz, size_t n, int granularity, double opt_dx,
                                        int opt_minvotes, int opt_nlines, int mode, h3d_comm *comm,
                                        h3d_iter *iters, size_t max_iters, uint32_t *pool, size_t pool_cap,
                                        h3d_result *res) {
  if (!res || (!xyz && n) || granularity < 0 || granularity > 8) return fail(H3D_ERR_ARG, "bad argument");
  if (n > 0x7FFFFFF0u) return fail(H3D_ERR_RANGE, "too many points");
  memset(res, 0, sizeof *res);
  static thread_local FrameCtx ctx;
  FrameCtx &f = ctx;
  H3D_TRY(frame_ensure(f, (unsigned)granularity, n, n * 3 * sizeof(double), max_iters ? max_iters : 1, pool_cap ? pool_cap : 1));
  cudaStream_t s = f.hg->stream;
  H3D_CUDA(cudaEventRecord(f.ev[0], s));
  if (n) H3D_CUDA(cudaMemcpyAsync(f.in_dev, xyz, n * 3 * sizeof(double), cudaMemcpyHostToDevice, s));
  f.timed = false;
  f.src = f.in_dev; f.src_stride = 3; f.src_is_float = 0;
  prep_kernel<double><<<1, 1024, 0, s>>>(static_cast<const double *>(f.in_dev), (int)n, nullptr, 3, f.hg->px,
                                         f.hg->py, f.hg->pz, f.orig, f.prep_dev);
  H3D_CUDA(cudaGetLastError());
  f.comm = comm;
  const int rc = frame_run(f, opt_minvotes, opt_nlines, opt_dx, mode, iters, max_iters, pool, pool_cap, res);
  f.comm = nullptr;
  for (auto &u : g_last_call_us) u = 0.0f;
  if (rc == H3D_OK && f.timed) {
    cudaEventElapsedTime(&g_last_call_us[0], f.ev[0], f.ev[1]);   // upload + NaN strip / bbox
    cudaEventElapsedTime(&g_last_call_us[1], f.ev[4], f.ev[2]);   // Hough::add (the first vote)
    cudaEventElapsedTime(&g_last_call_us[2], f.ev[2], f.ev[5]);   // the loop's kernels (+ in-stream all-reduces)
    cudaEventElapsedTime(&g_last_call_us[3], f.ev[0], f.ev[3]);   // whole call on the device, result copies included
    for (auto &u : g_last_call_us) u *= 1000.0f;
  }
  return rc;
}

extern "C" int h3d_hough3dlines_timing(float us[4]) {
  if (!us) return fail(H3D_ERR_ARG, "bad argument");
  for (int k = 0; k < 4; ++k) us[k] = g_last_call_us[k];
  return H3D_OK;
}

// =====================================================================================
// K0 host side + the persistent pipeline (pre-filter + loop, buffers kept between frames)
// =====================================================================================
namespace {

struct PrefilterCtx {
  K0State *st = nullptr;
  K0State *st_host = nullptr;   // pinned
  int *count = nullptr, *start = nullptr, *slot = nullptr, *cursor = nullptr, *members = nullptr;
  int2 *block_sums = nullptr;
  size_t grid_cap = 0, member_cap = 0;
  ~PrefilterCtx() {
    cudaFree(st); cudaFreeHost(st_host); cudaFree(count); cudaFree(start); cudaFree(slot); cudaFree(cursor);
    cudaFree(members); cudaFree(block_sums);
  }
};

int prefilter_ensure(PrefilterCtx &p, size_t grid_cells, size_t n_points) {
  if (!p.st) {
    H3D_CUDA(cudaMalloc(&p.st, sizeof(K0State)));
    H3D_CUDA(cudaMallocHost(&p.st_host, sizeof(K0State)));
  }
  if (grid_cells > p.grid_cap) {
    cudaFree(p.count); cudaFree(p.start); cudaFree(p.slot); cudaFree(p.cursor); cudaFree(p.block_sums);
    p.count = p.start = p.slot = p.cursor = nullptr; p.block_sums = nullptr; p.grid_cap = 0;
    H3D_CUDA(cudaMalloc(&p.count, grid_cells * sizeof(int)));
    H3D_CUDA(cudaMalloc(&p.start, grid_cells * sizeof(int)));
    H3D_CUDA(cudaMalloc(&p.slot, grid_cells * sizeof(int)));
    H3D_CUDA(cudaMalloc(&p.cursor, grid_cells * sizeof(int)));
    const size_t nb = (grid_cells + kScanBlock * kScanItems - 1) / (kScanBlock * kScanItems);
    H3D_CUDA(cudaMalloc(&p.block_sums, (nb + 1) * sizeof(int2)));
    p.grid_cap = grid_cells;
  }
  if (n_points > p.member_cap) {
    cudaFree(p.members); p.members = nullptr; p.member_cap = 0;
    H3D_CUDA(cudaMalloc(&p.members, n_points * sizeof(int)));
    p.member_cap = n_points;
  }
  return H3D_OK;
}

K0Window make_window(const float win[6]) {
  K0Window w;
  for (int c = 0; c < 3; ++c) { w.lo[c] = win[2 * c]; w.hi[c] = win[2 * c + 1]; }
  return w;
}

// voxel-grid cells a cloud confined to `win` can need
size_t window_grid_cells(const float win[6], float leaf) {
  const float inv = 1.0f / leaf;
  double cells = 1;
  for (int c = 0; c < 3; ++c) {
    const double lo = floor((double)(win[2 * c] * inv)), hi = floor((double)(win[2 * c + 1] * inv));
    cells *= (hi - lo + 2.0);   // +1 cell of slack per axis for float rounding at the limits
  }
  if (!(cells >= 1.0) || cells > 268435456.0) return 0;
  return (size_t)cells;
}

// phase 1: window + bbox + grid geometry;  phase 2: counting sort + centroids.  All on `s`.
int prefilter_phase1(PrefilterCtx &p, const float *in_dev, int n, int stride, const K0Window &w, float leaf,
                     int sm, cudaStream_t s) {
  k0_init_kernel<<<1, 1, 0, s>>>(p.st);
  if (n) k0_bbox_kernel<<<sm * 2, 256, 0, s>>>(in_dev, n, stride, w, p.st);
  k0_setup_kernel<<<1, 1, 0, s>>>(p.st, leaf, (int)(p.grid_cap < 0x7FFFFFFF ? p.grid_cap : 0x7FFFFFFF));
  note_launches(n ? 3 : 2);
  H3D_CUDA(cudaGetLastError());
  return H3D_OK;
}
int prefilter_phase2(PrefilterCtx &p, const float *in_dev, int n, int stride, const K0Window &w, size_t grid_cells,
                     float *out_dev, int out_cap, int out_stride, int sm, cudaStream_t s) {
  if (n == 0 || grid_cells == 0) return H3D_OK;
  const int nb = (int)((grid_cells + kScanBlock * kScanItems - 1) / (kScanBlock * kScanItems));
  H3D_CUDA(cudaMemsetAsync(p.count, 0, grid_cells * sizeof(int), s));
  k0_count_kernel<<<sm * 2, 256, 0, s>>>(in_dev, n, stride, w, p.st, p.count);
  k0_scan_reduce_kernel<<<nb, kScanBlock, 0, s>>>(p.count, p.st, p.block_sums);
  k0_scan_spine_kernel<<<1, 1024, 0, s>>>(p.block_sums, p.st);
  k0_scan_apply_kernel<<<nb, kScanBlock, 0, s>>>(p.count, p.st, p.block_sums, p.start, p.slot, p.cursor);
  k0_scatter_kernel<<<sm * 2, 256, 0, s>>>(in_dev, n, stride, w, p.st, p.start, p.cursor, p.members);
  k0_centroid_kernel<<<(unsigned)((grid_cells + 255) / 256), 256, 0, s>>>(in_dev, stride, p.st, p.count, p.start,
                                                                          p.slot, p.members, out_dev, out_cap,
                                                                          out_stride);
  note_launches(6);
  H3D_CUDA(cudaGetLastError());
  return H3D_OK;
}

}  // namespace

extern "C" int h3d_prefilter(const float *xyz_in, size_t n, size_t stride, const float win[6], float leaf,
                             float *out_xyz, size_t cap, size_t *n_out) {
  if ((!xyz_in && n) || !win || !n_out || stride < 3 || !(leaf > 0.0f) || (!out_xyz && cap))
    return fail(H3D_ERR_ARG, "bad argument");
  if (n > 0x7FFFFFF0u) return fail(H3D_ERR_RANGE, "too many points");
  *n_out = 0;
  if (n == 0) return H3D_OK;
  DeviceFacts dev;
  H3D_TRY(device_facts(&dev));
  static thread_local PrefilterCtx ctx;
  PrefilterCtx &p = ctx;
  cudaStream_t s = nullptr;   // legacy default stream: this entry point is synchronous
  float *in_dev = nullptr, *out_dev = nullptr;
  H3D_CUDA(cudaMalloc(&in_dev, n * stride * sizeof(float)));
  int rc = H3D_OK;
  do {
    if (cudaMemcpyAsync(in_dev, xyz_in, n * stride * sizeof(float), cudaMemcpyHostToDevice, s) != cudaSuccess) { rc = fail(H3D_ERR_CUDA, "upload failed"); break; }
    if ((rc = prefilter_ensure(p, 1, n)) != H3D_OK) break;
    const K0Window w = make_window(win);
    // the grid is sized from the actual bounding box: geometry first, with an unlimited capacity
    const size_t saved = p.grid_cap;
    p.grid_cap = 0x7FFFFFFF;
    rc = prefilter_phase1(p, in_dev, (int)n, (int)stride, w, leaf, dev.sm_count, s);
    p.grid_cap = saved;
    if (rc != H3D_OK) break;
    if (cudaMemcpyAsync(p.st_host, p.st, sizeof(K0State), cudaMemcpyDeviceToHost, s) != cudaSuccess ||
        cudaStreamSynchronize(s) != cudaSuccess) { rc = fail(H3D_ERR_CUDA, "pre-filter geometry read-back failed"); break; }
    if (p.st_host->overflow) { rc = fail(H3D_ERR_RANGE, "voxel grid index would overflow int32 (leaf too small)"); break; }
    const size_t cells = (size_t)p.st_host->n_vox;
    if (cells == 0) break;   // nothing inside the window
    if (cells > 268435456u) { rc = fail(H3D_ERR_RANGE, "voxel grid larger than 2^28 cells"); break; }
    if ((rc = prefilter_ensure(p, cells, n)) != H3D_OK) break;
    const size_t out_cap = cap < (size_t)p.st_host->n_in_window ? cap : (size_t)p.st_host->n_in_window;
    if (out_cap && cudaMalloc(&out_dev, out_cap * 3 * sizeof(float)) != cudaSuccess) { rc = fail(H3D_ERR_NOMEM, "output allocation failed"); break; }
    if ((rc = prefilter_phase2(p, in_dev, (int)n, (int)stride, w, cells, out_dev, (int)out_cap, 3, dev.sm_count, s)) != H3D_OK) break;
    if (cudaMemcpyAsync(p.st_host, p.st, sizeof(K0State), cudaMemcpyDeviceToHost, s) != cudaSuccess ||
        cudaStreamSynchronize(s) != cudaSuccess) { rc = fail(H3D_ERR_CUDA, "pre-filter failed"); break; }
    *n_out = (size_t)p.st_host->n_out;
    const size_t n_copy = *n_out < out_cap ? *n_out : out_cap;
    if (n_copy && cudaMemcpy(out_xyz, out_dev, n_copy * 3 * sizeof(float), cudaMemcpyDeviceToHost) != cudaSuccess) rc = fail(H3D_ERR_CUDA, "download failed");
  } while (0);
  cudaFree(in_dev); cudaFree(out_dev);
  return rc;
}

// Frame fast path (h3d_frame.cuh): K0 + prep + Hough ctor in one cluster launch, first vote, fused loop,
// post-tests and ONE export kernel that writes the results into mapped pinned host memory; the host
// waits once per frame.
struct FastPath {
  bool on = false;              // the pipeline qualifies (pre-filter with a finite window, bounded line count)
  bool count_clean = false;     // pf.count is all zero (the fast path leaves it so; the general path does not)
  bool last_was_fast = false;   // the last frame ran here (h3d_pipeline_filtered reads the count from the descriptor)
  K0Grid grid{};
  int2 *occ = nullptr;
  int *sorted = nullptr;        // voxel member lists in ascending raw index (max_raw_points ints)
  int *tile_sums = nullptr;
  FrameDev *fd_dev = nullptr;
  size_t smem_vote = 0, smem_loop = 0;
  // host staging (pinned, mapped): [FrameDev][LoopState][iters][pool]; the segment records go to f.seg_host
  unsigned char *stage_host = nullptr, *stage_dev = nullptr;
  size_t off_st = 0, off_iters = 0, off_pool = 0, pool_entries = 0, stage_bytes = 0;
  size_t sub_max_iters = 0;     // of the frame in flight (frame_submit_fast -> frame_complete_fast)
  void release() {
    cudaFree(occ); cudaFree(sorted); sorted = nullptr; cudaFree(tile_sums); cudaFree(fd_dev); cudaFreeHost(stage_host);
    occ = nullptr; tile_sums = nullptr; fd_dev = nullptr; stage_host = stage_dev = nullptr; on = false;
  }
};

struct h3d_pipeline {
  h3d_pipeline_config cfg;
  FrameCtx f;
  PrefilterCtx pf;
  K0Window win;
  size_t grid_cells = 0;
  int device = 0;               // the CUDA device the pipeline was created on
  float *raw_dev = nullptr;     // max_raw_points * 4 floats
  float *filt_dev = nullptr;    // max_raw_points * 3 floats
  size_t raw_cap_floats = 0;
  unsigned char *msg_dev = nullptr;   // PointCloud2 byte buffer of the current frame (h3d_pipeline_run_cloud2)
  size_t msg_cap = 0;
  FastPath fast;
};

static std::atomic<int> g_fast_frames{1};   // h3d_set_fast_frames
static std::atomic<int> g_vote_track{2};    // h3d_set_vote_track: first vote of the fast path tracks maxima instead of scanning planes

// one-time set-up of the fast path for a pipeline; leaves fast.on = false when it does not qualify
static int fast_path_setup(h3d_pipeline *p) {
  FastPath &fp = p->fast;
  FrameCtx &f = p->f;
  if (!p->cfg.prefilter || p->cfg.opt_nlines <= 0 || !(p->cfg.opt_dx > 0.0)) return H3D_OK;
  if (!make_window_grid(p->win, p->cfg.leaf, p->pf.grid_cap, &fp.grid)) return H3D_OK;
  const DeviceFacts &dev = f.hg->dev;
  if (dev.cc < 90) return H3D_OK;   // thread-block clusters
  cudaFuncAttributes fa;
  H3D_CUDA(cudaFuncGetAttributes(&fa, vote_frame_kernel<2>));
  if (dev.smem_optin < fa.sharedSizeBytes + 1024) return H3D_OK;
  fp.smem_vote = ((dev.smem_optin - fa.sharedSizeBytes) / 16) * 16;
  H3D_CUDA(cudaFuncGetAttributes(&fa, fused_loop_kernel));
  if (dev.smem_optin < fa.sharedSizeBytes + 1024) return H3D_OK;
  fp.smem_loop = ((dev.smem_optin - fa.sharedSizeBytes - 256) / 16) * 16;
  if (fp.smem_loop < kFusedHdr + 4096) return H3D_OK;
  const size_t n = p->cfg.max_raw_points;
  H3D_CUDA(cudaMalloc(&fp.occ, n * sizeof(int2)));
  H3D_CUDA(cudaMalloc(&fp.sorted, n * sizeof(int)));
  H3D_CUDA(cudaMalloc(&fp.tile_sums, 2 * kK0Cluster * sizeof(int)));
  H3D_CUDA(cudaMalloc(&fp.fd_dev, sizeof(FrameDev)));
  fp.off_st = (sizeof(FrameDev) + 255) & ~(size_t)255;
  fp.off_iters = fp.off_st + ((sizeof(LoopState) + 255) & ~(size_t)255);
  fp.off_pool = fp.off_iters + ((f.iters_cap * sizeof(h3d_iter) + 255) & ~(size_t)255);
  fp.pool_entries = 2 * f.iters_cap * (size_t)kFrameMaxPoints;
  if (fp.pool_entries > f.pool_cap) fp.pool_entries = f.pool_cap;
  fp.stage_bytes = fp.off_pool + fp.pool_entries * sizeof(unsigned);
  H3D_CUDA(cudaHostAlloc(&fp.stage_host, fp.stage_bytes, cudaHostAllocMapped));
  H3D_CUDA(cudaHostGetDevicePointer(reinterpret_cast<void **>(&fp.stage_dev), fp.stage_host, 0));
  memset(fp.stage_host, 0, fp.off_pool);
  H3D_CUDA(cudaMemsetAsync(p->pf.count, 0, p->pf.grid_cap * sizeof(int), f.hg->stream));
  fp.count_clean = true;
  fp.on = true;
  return H3D_OK;
}

// One frame on the fast path.  *done = false: the frame did not qualify (FrameDev::status 2) and must be
// re-run on the general path (the raw frame is still where it was).
// The fast path in two halves, so that one host thread can keep several pipelines busy (replay_impl): the first
// enqueues the whole frame on the pipeline's stream and returns, the second waits (unless the caller saw the
// frame's last event complete) and copies the results out of the mapped staging area.
static int frame_submit_fast(h3d_pipeline *p, const float *raw_dev, size_t n, size_t stride, size_t pool_cap) {
  FastPath &fp = p->fast;
  FrameCtx &f = p->f;
  h3d_hough *h = f.hg;
  cudaStream_t s = h->stream;
  const int opt_nlines = p->cfg.opt_nlines;
  const int grid = (f.fused_ctas > 0 && f.fused_ctas < h->dev.sm_count) ? f.fused_ctas : h->dev.sm_count;
  HoughDev &hd = h->d;
  hd.num_b = (int)h->space->num_b; hd.coef = h->space->coef_dev;
  hd.dir_begin = 0; hd.dir_count = hd.num_b;
  hd.n = 0; hd.n_dev = nullptr;
  h->has_negative = false;
  const int groups = (hd.dir_count + kResG - 1) / kResG;
  const int n_own = ((groups + grid - 1) / grid) * kResG;
  const size_t xchg_words = 2 * ((size_t)opt_nlines + 2);
  if (xchg_words > f.gbest_cap) {
    H3D_CUDA(cudaStreamSynchronize(s));
    cudaFree(f.gbest); f.gbest = nullptr; f.gbest_cap = 0;
    H3D_CUDA(cudaMalloc(&f.gbest, xchg_words * sizeof(unsigned long long)));
    f.gbest_cap = xchg_words;
  }
  if (!fp.count_clean) {
    H3D_CUDA(cudaMemsetAsync(p->pf.count, 0, p->pf.grid_cap * sizeof(int), s));
    fp.count_clean = true;
  }
  const unsigned st_pool_cap = (unsigned)((pool_cap < f.pool_cap && !f.seg_on) ? pool_cap : f.pool_cap);
  const size_t st_max_iters = p->cfg.max_iters < f.iters_cap ? p->cfg.max_iters : f.iters_cap;

  FrameArgs fa;
  fa.raw = raw_dev; fa.n = (int)n; fa.stride = (int)stride;
  fa.win = p->win; fa.grid = fp.grid;
  fa.count = p->pf.count; fa.start = p->pf.start; fa.ncell = p->pf.slot; fa.members = p->pf.members; fa.sorted = fp.sorted; fa.occ = fp.occ; fa.tile_sums = fp.tile_sums;
  fa.filt = p->filt_dev; fa.out_cap = (int)p->cfg.max_raw_points;
  fa.px = h->px; fa.py = h->py; fa.pz = h->pz; fa.w = h->w; fa.orig = f.orig;
  fa.fd = fp.fd_dev;
  fa.opt_dx = p->cfg.opt_dx; fa.num_b = (unsigned long long)hd.num_b;
  fa.st = f.st_dev; fa.minvotes = p->cfg.opt_minvotes; fa.nlines_max = opt_nlines; fa.max_iters = (int)st_max_iters;
  fa.pool_cap = st_pool_cap;
  fa.xchg = f.gbest; fa.xchg_words = (int)xchg_words; fa.bar = f.sync_dev;
  fa.smem_loop = (unsigned)(fp.smem_loop - kFusedHdr); fa.tables_bytes = (unsigned)n_own * 16u;   // the loop's body behind its header
  fa.smem_vote = (unsigned)(fp.smem_vote - kVoteFramePlanesOff);   // what the planes get behind the kernel's header
  {
    // cluster width: the whole GPU is this frame's -> 8 (latency); frames in flight beside it -> 2
    const unsigned width = (grid >= h->dev.sm_count / 2) ? (unsigned)kK0Cluster : 2u;
    cudaLaunchConfig_t lc = {};
    lc.gridDim = dim3(width); lc.blockDim = dim3(kK0Threads); lc.dynamicSmemBytes = 0; lc.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = width; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    lc.attrs = at; lc.numAttrs = 1;
    H3D_CUDA(cudaLaunchKernelEx(&lc, frame_prefilter_kernel, fa));
  }
  H3D_CUDA(cudaGetLastError());
  H3D_CUDA(cudaEventRecord(f.ev[1], s));
  if (g_debug_poison.load(std::memory_order_relaxed)) {
    const size_t nd = (size_t)(hd.dir_count > 0 ? hd.dir_count : 1);
    H3D_CUDA(cudaMemsetAsync(hd.dvotes, 0xFF, nd * sizeof(unsigned), s));
    H3D_CUDA(cudaMemsetAsync(hd.dcell, 0xFF, nd * sizeof(unsigned), s));
  }
  // Hough::add, wrap.h:228
  const int track = g_vote_track.load(std::memory_order_relaxed);
  const void *vk = track == 0 ? reinterpret_cast<const void *>(vote_frame_kernel<0>)
                 : track == 1 ? reinterpret_cast<const void *>(vote_frame_kernel<1>)
                              : reinterpret_cast<const void *>(vote_frame_kernel<2>);
  H3D_TRY(ensure_dyn_smem(vk, fp.smem_vote, h->dev.smem_optin));
  H3D_CUDA(cudaEventRecord(f.ev[4], s));
  if (track == 0) vote_frame_kernel<0><<<grid, kResThreads, fp.smem_vote, s>>>(hd, fp.fd_dev);
  else if (track == 1) vote_frame_kernel<1><<<grid, kResThreads, fp.smem_vote, s>>>(hd, fp.fd_dev);
  else vote_frame_kernel<2><<<grid, kResThreads, fp.smem_vote, s>>>(hd, fp.fd_dev);
  H3D_CUDA(cudaGetLastError());
  H3D_CUDA(cudaEventRecord(f.ev[2], s));
  // the loop, wrap.h:237-342
  FusedArgs a;
  memset(&a, 0, sizeof a);
  a.h = hd;
  a.st = f.st_dev; a.iters = f.iters_dev; a.pool = f.pool_dev; a.orig = f.orig; a.dirs = h->space->dirs_dev;
  a.xchg = f.gbest; a.bar = f.sync_dev;
  a.n_own_max = n_own;
  a.near_t = sqrt_le_threshold(p->cfg.opt_dx);
  a.prof = f.prof_dev; a.prof_iters = f.prof_iters;
  a.fd = fp.fd_dev;
  f.prof_grid = grid;
  if (f.prof_dev)
    H3D_CUDA(cudaMemsetAsync(f.prof_dev, 0, (size_t)f.prof_iters * f.prof_ctas * kProfSlots * sizeof(unsigned long long), s));
  H3D_TRY(launch_fused(f, a, fp.smem_loop, grid));
  H3D_CUDA(cudaEventRecord(f.ev[5], s));
  // wrap.h:261-337 on every accepted line, then everything back to the host in one piece
  f.src = p->filt_dev; f.src_stride = 3; f.src_is_float = 1;
  f.seg_pool_cap = st_pool_cap; f.seg_grid = 0; f.seg_n = 0;
  f.seg_ns_dev = &fp.fd_dev->ns[0];
  hd.n = (int)p->cfg.max_raw_points;   // sizes the post-tests' scratch (the real count is on the device)
  const int rc_seg = enqueue_segments(f, (size_t)opt_nlines);
  f.seg_ns_dev = nullptr;
  hd.n = 0;
  if (rc_seg != H3D_OK) return rc_seg;
  ExportArgs ea;
  ea.fd = fp.fd_dev; ea.st = f.st_dev; ea.iters = f.iters_dev; ea.pool = f.pool_dev;
  ea.segs = (f.seg_on && f.seg_grid) ? f.seg_dev : nullptr;
  ea.max_iters = (int)st_max_iters; ea.seg_cap = (int)f.seg_grid;
  ea.pool_host_cap = (unsigned)fp.pool_entries;
  ea.fd_out = reinterpret_cast<FrameDev *>(fp.stage_dev);
  ea.st_out = reinterpret_cast<LoopState *>(fp.stage_dev + fp.off_st);
  ea.iters_out = reinterpret_cast<h3d_iter *>(fp.stage_dev + fp.off_iters);
  ea.pool_out = reinterpret_cast<unsigned *>(fp.stage_dev + fp.off_pool);
  ea.segs_out = nullptr;
  if (ea.segs) H3D_CUDA(cudaHostGetDevicePointer(reinterpret_cast<void **>(&ea.segs_out), f.seg_host, 0));
  frame_export_kernel<<<8, 256, 0, s>>>(ea);
  H3D_CUDA(cudaGetLastError());
  note_launches(3);                       // K0 cluster, first vote, export (loop and post-tests count themselves)
  H3D_CUDA(cudaEventRecord(f.ev[3], s));
  fp.sub_max_iters = st_max_iters;
  return H3D_OK;
}

static int frame_complete_fast(h3d_pipeline *p, h3d_iter *iters, uint32_t *pool, size_t pool_cap, h3d_result *res,
                               bool *done, bool already_complete) {
  *done = false;
  FastPath &fp = p->fast;
  FrameCtx &f = p->f;
  h3d_hough *h = f.hg;
  const HoughDev &hd = h->d;
  const size_t st_max_iters = fp.sub_max_iters;
  if (!already_complete) H3D_CUDA(frame_wait(f));   // the frame's only wait
  const FrameDev &F = *reinterpret_cast<const FrameDev *>(fp.stage_host);
  const LoopState &fin = *reinterpret_cast<const LoopState *>(fp.stage_host + fp.off_st);
  if (F.status == 2) return H3D_OK;       // *done stays false: the general path takes the frame
  *done = true;
  fp.last_was_fast = true;
  f.timed = true; f.fused = true;
  res->n_input = (uint32_t)F.n;
  res->num_b = h->space->num_b;
  res->d = F.d;
  if (F.status == 1) { res->rc = 1; return H3D_OK; }
  for (int c = 0; c < 3; ++c) { res->shift[c] = 0.0 + F.ns[c]; f.seg_ns[c] = F.ns[c]; }
  res->dx = F.g.dx; res->max_x = F.g.max_x; res->num_x = (uint64_t)F.g.num_x;
  res->n_iters = (uint32_t)fin.n_iters; res->n_lines = (uint32_t)fin.n_lines; res->pool_used = fin.pool_used;
  res->votes_cast = (uint64_t)F.n * hd.dir_count + fin.votes_cast;
  const size_t n_rec = (size_t)fin.n_iters < st_max_iters ? (size_t)fin.n_iters : st_max_iters;
  if (n_rec && iters) memcpy(iters, fp.stage_host + fp.off_iters, n_rec * sizeof(h3d_iter));
  size_t n_pool = (size_t)fin.pool_used < pool_cap ? (size_t)fin.pool_used : pool_cap;
  if (n_pool > fp.pool_entries) n_pool = fp.pool_entries;
  if (n_pool && pool) memcpy(pool, fp.stage_host + fp.off_pool, n_pool * sizeof(unsigned));
  if (f.seg_on && n_rec) f.seg_n = n_rec < f.seg_grid ? n_rec : f.seg_grid;
  return H3D_OK;
}

static int frame_run_fast(h3d_pipeline *p, const float *raw_dev, size_t n, size_t stride, h3d_iter *iters,
                          uint32_t *pool, size_t pool_cap, h3d_result *res, bool *done) {
  *done = false;
  H3D_TRY(frame_submit_fast(p, raw_dev, n, stride, pool_cap));
  return frame_complete_fast(p, iters, pool, pool_cap, res, done, false);
}

// packs x,y,z of a PointCloud2 byte buffer into (x,y,z,0) float quadruples, message order
__global__ void cloud2_gather_kernel(const unsigned char *__restrict__ data, h3d_cloud2_layout L, float4 *out) {
  const size_t n = (size_t)L.width * L.height;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const unsigned char *pt = data + (i / L.width) * L.row_step + (i % L.width) * L.point_step;
    out[i] = make_float4(*reinterpret_cast<const float *>(pt + L.off_x), *reinterpret_cast<const float *>(pt + L.off_y),
                         *reinterpret_cast<const float *>(pt + L.off_z), 0.0f);
  }
}

extern "C" int h3d_pipeline_create(const h3d_pipeline_config *cfg, h3d_pipeline **out) {
  if (!cfg || !out || cfg->granularity < 0 || cfg->granularity > 8 || cfg->max_raw_points == 0 ||
      cfg->max_raw_points > 0x7FFFFFF0u)
    return fail(H3D_ERR_ARG, "bad pipeline configuration");
  h3d_pipeline *p = new (std::nothrow) h3d_pipeline();
  if (!p) return fail(H3D_ERR_NOMEM, "host allocation failed");
  std::unique_ptr<h3d_pipeline, void (*)(h3d_pipeline *)> guard(p, h3d_pipeline_destroy);
  p->cfg = *cfg;
  H3D_CUDA(cudaGetDevice(&p->device));
  const size_t n = cfg->max_raw_points;
  const size_t max_iters = cfg->max_iters ? cfg->max_iters : (size_t)(cfg->opt_nlines > 0 ? cfg->opt_nlines + 1 : 64);
  p->cfg.max_iters = max_iters;
  H3D_TRY(frame_ensure(p->f, (unsigned)cfg->granularity, n, 0, max_iters, 2 * n * max_iters));
  if (cfg->prefilter) {
    if (!(cfg->leaf > 0.0f)) return fail(H3D_ERR_ARG, "leaf must be positive");
    p->win = make_window(cfg->window);
    p->grid_cells = window_grid_cells(cfg->window, cfg->leaf);
    if (p->grid_cells == 0) return fail(H3D_ERR_RANGE, "window / leaf give a voxel grid beyond 2^28 cells");
    H3D_TRY(prefilter_ensure(p->pf, p->grid_cells, n));
    H3D_CUDA(cudaMalloc(&p->filt_dev, n * 3 * sizeof(float)));
  }
  p->raw_cap_floats = n * 8;
  H3D_CUDA(cudaMalloc(&p->raw_dev, p->raw_cap_floats * sizeof(float)));
  H3D_TRY(fast_path_setup(p));
  *out = guard.release();
  return H3D_OK;
}

extern "C" void h3d_pipeline_destroy(h3d_pipeline *p) {
  if (!p) return;
  if (p->f.hg) cudaStreamSynchronize(p->f.hg->stream);
  cudaFree(p->raw_dev); cudaFree(p->filt_dev); cudaFree(p->msg_dev);
  p->fast.release();
  delete p;
}

static void pipeline_fast_timing(FrameCtx &f) {
  for (auto &u : f.us) u = 0.0f;
  if (f.timed) {
    cudaEventElapsedTime(&f.us[0], f.ev[0], f.ev[1]);   // pre-filter + prep + ctor (one cluster launch)
    cudaEventElapsedTime(&f.us[1], f.ev[4], f.ev[2]);   // the first vote kernel alone (Hough::add)
    cudaEventElapsedTime(&f.us[2], f.ev[2], f.ev[5]);   // the loop kernel
    cudaEventElapsedTime(&f.us[3], f.ev[0], f.ev[3]);   // whole frame, results in host memory
    for (auto &u : f.us) u *= 1000.0f;
  }
}

// the general path of a frame: K0 kernels, widening, host-side Hough constructor, frame_run
static int pipeline_run_general(h3d_pipeline *p, const float *raw_dev, size_t n, size_t stride, h3d_iter *iters,
                                uint32_t *pool, size_t pool_cap, h3d_result *res) {
  FrameCtx &f = p->f;
  h3d_hough *h = f.hg;
  cudaStream_t s = h->stream;
  if (p->cfg.prefilter) {
    p->fast.count_clean = false;
    H3D_TRY(prefilter_phase1(p->pf, raw_dev, (int)n, (int)stride, p->win, p->cfg.leaf, h->dev.sm_count, s));
    H3D_TRY(prefilter_phase2(p->pf, raw_dev, (int)n, (int)stride, p->win, p->grid_cells, p->filt_dev, (int)p->cfg.max_raw_points, 3, h->dev.sm_count, s));
    // wrap.h:175-189: float -> double widening of the filtered cloud (+ NaN/Inf strip, bbox)
    prep_kernel<float><<<1, 1024, 0, s>>>(p->filt_dev, 0, &p->pf.st->n_out, 3, h->px, h->py, h->pz, f.orig, f.prep_dev);
    f.src = p->filt_dev; f.src_stride = 3; f.src_is_float = 1;
  } else {
    f.src = raw_dev; f.src_stride = (int)stride; f.src_is_float = 1;
    prep_kernel<float><<<1, 1024, 0, s>>>(raw_dev, (int)n, nullptr, (int)stride, h->px, h->py, h->pz, f.orig, f.prep_dev);
  }
  H3D_CUDA(cudaGetLastError());
  const int rc = frame_run(f, p->cfg.opt_minvotes, p->cfg.opt_nlines, p->cfg.opt_dx, 0, iters, p->cfg.max_iters, pool,
                           pool_cap, res);
  for (auto &u : f.us) u = 0.0f;
  if (rc == H3D_OK && f.timed) {
    cudaEventElapsedTime(&f.us[0], f.ev[0], f.ev[1]);   // pre-filter + prep
    cudaEventElapsedTime(&f.us[1], f.ev[4], f.ev[2]);   // the first vote kernel alone (Hough::add)
    cudaEventElapsedTime(&f.us[2], f.ev[2], f.ev[5]);   // the loop's kernels
    cudaEventElapsedTime(&f.us[3], f.ev[0], f.ev[3]);   // whole frame
    for (auto &u : f.us) u *= 1000.0f;
  }
  return rc;
}

static bool pipeline_fast_eligible(const h3d_pipeline *p, size_t n) {
  return p->fast.on && g_use_fused && g_fast_frames.load(std::memory_order_relaxed) && n > 0;
}

extern "C" int h3d_pipeline_run_dev(h3d_pipeline *p, const float *raw_dev, size_t n, size_t stride, h3d_iter *iters,
                                    uint32_t *pool, size_t pool_cap, h3d_result *res) {
  if (!p || !res || (!raw_dev && n) || stride < 3) return fail(H3D_ERR_ARG, "bad argument");
  if (n > p->cfg.max_raw_points) return fail(H3D_ERR_ARG, "frame larger than max_raw_points");
  H3D_CUDA(cudaSetDevice(p->device));   // the calling thread may be a fresh one (current device 0)
  memset(res, 0, sizeof *res);
  FrameCtx &f = p->f;
  f.timed = false;
  cudaStream_t s = f.hg->stream;
  H3D_CUDA(cudaEventRecord(f.ev[0], s));
  p->fast.last_was_fast = false;
  if (pipeline_fast_eligible(p, n)) {
    bool done = false;
    const int rc_fast = frame_run_fast(p, raw_dev, n, stride, iters, pool, pool_cap, res, &done);
    if (rc_fast != H3D_OK) return rc_fast;
    if (done) { pipeline_fast_timing(f); return H3D_OK; }
    memset(res, 0, sizeof *res);            // FrameDev::status 2: the general path takes the frame
    f.timed = false;
    H3D_CUDA(cudaEventRecord(f.ev[0], s));
  }
  return pipeline_run_general(p, raw_dev, n, stride, iters, pool, pool_cap, res);
}

// ---- split-phase form for the replay: submit a frame, later complete it ------------------------------------
// *async = false: the pipeline does not qualify for the fast path; the caller runs the frame synchronously.
static int pipeline_submit(h3d_pipeline *p, const float *raw, int on_device, size_t n, size_t stride, size_t pool_cap,
                           const float **raw_dev_out, bool *async) {
  *async = false;
  if (!p || (!raw && n) || stride < 3) return fail(H3D_ERR_ARG, "bad argument");
  if (n > p->cfg.max_raw_points) return fail(H3D_ERR_ARG, "frame larger than max_raw_points");
  if (!on_device && n * stride > p->raw_cap_floats) return fail(H3D_ERR_ARG, "frame larger than the pipeline's raw buffer");
  if (!pipeline_fast_eligible(p, n)) return H3D_OK;
  H3D_CUDA(cudaSetDevice(p->device));
  FrameCtx &f = p->f;
  cudaStream_t s = f.hg->stream;
  const float *raw_dev = raw;
  if (!on_device) {
    H3D_CUDA(cudaMemcpyAsync(p->raw_dev, raw, n * stride * sizeof(float), cudaMemcpyHostToDevice, s));
    raw_dev = p->raw_dev;
  }
  *raw_dev_out = raw_dev;
  f.timed = false;
  H3D_CUDA(cudaEventRecord(f.ev[0], s));
  p->fast.last_was_fast = false;
  H3D_TRY(frame_submit_fast(p, raw_dev, n, stride, pool_cap));
  *async = true;
  return H3D_OK;
}

// 0: still running; 1: the frame's last event has completed; -1: the query failed (the completion's wait reports it)
static int pipeline_ready(h3d_pipeline *p) {
  const cudaError_t e = cudaEventQuery(p->f.ev[3]);
  return e == cudaErrorNotReady ? 0 : e == cudaSuccess ? 1 : -1;
}

static int pipeline_complete(h3d_pipeline *p, const float *raw_dev, size_t n, size_t stride, h3d_iter *iters,
                             uint32_t *pool, size_t pool_cap, h3d_result *res, bool already_complete) {
  H3D_CUDA(cudaSetDevice(p->device));
  memset(res, 0, sizeof *res);
  FrameCtx &f = p->f;
  bool done = false;
  H3D_TRY(frame_complete_fast(p, iters, pool, pool_cap, res, &done, already_complete));
  if (done) { pipeline_fast_timing(f); return H3D_OK; }
  memset(res, 0, sizeof *res);              // FrameDev::status 2: the general path takes the frame
  f.timed = false;
  H3D_CUDA(cudaEventRecord(f.ev[0], f.hg->stream));
  return pipeline_run_general(p, raw_dev, n, stride, iters, pool, pool_cap, res);
}

extern "C" int h3d_pipeline_run(h3d_pipeline *p, const float *raw, size_t n, size_t stride, h3d_iter *iters,
                                uint32_t *pool, size_t pool_cap, h3d_result *res) {
  if (!p || (!raw && n) || stride < 3) return fail(H3D_ERR_ARG, "bad argument");
  if (n * stride > p->raw_cap_floats) return fail(H3D_ERR_ARG, "frame larger than the pipeline's raw buffer");
  H3D_CUDA(cudaSetDevice(p->device));
  if (n) H3D_CUDA(cudaMemcpyAsync(p->raw_dev, raw, n * stride * sizeof(float), cudaMemcpyHostToDevice, p->f.hg->stream));
  return h3d_pipeline_run_dev(p, p->raw_dev, n, stride, iters, pool, pool_cap, res);
}

static bool segment_config_ok(const h3d_segment_config *c) {
  return c && c->n_radius >= 1 && c->n_radius <= H3D_MAX_RADIUS_SIZES;
}

extern "C" int h3d_pipeline_enable_segments(h3d_pipeline *p, const h3d_segment_config *cfg) {
  if (!p) return fail(H3D_ERR_ARG, "bad argument");
  if (!cfg) { p->f.seg_on = false; p->f.seg_n = 0; return H3D_OK; }
  if (!segment_config_ok(cfg)) return fail(H3D_ERR_ARG, "radius_sizes needs 1..8 entries");
  p->f.seg_cfg = *cfg;
  p->f.seg_on = true;
  return H3D_OK;
}

// compacts the per-iteration records to the lines the reference tests, up to the first abort
static size_t collect_segments(const h3d_segment *rec, size_t n_rec, h3d_segment *segs, size_t cap, bool *aborted,
                               bool *truncated) {
  size_t m = 0;
  *aborted = false; *truncated = false;
  for (size_t k = 0; k < n_rec; ++k) {
    if (rec[k].status == kSegNone) continue;
    if (rec[k].status == kSegTruncated) { *truncated = true; break; }
    if (m < cap && segs) segs[m] = rec[k];
    ++m;
    if (rec[k].status == 3) { *aborted = true; break; }
  }
  return m;
}

extern "C" int h3d_pipeline_segments(h3d_pipeline *p, h3d_segment *segs, size_t cap, size_t *n_segs, int *rc_ref) {
  if (!p || !n_segs) return fail(H3D_ERR_ARG, "bad argument");
  if (!p->f.seg_on) return fail(H3D_ERR_ARG, "segment post-tests are not enabled on this pipeline");
  bool aborted, truncated;
  *n_segs = collect_segments(p->f.seg_host, p->f.seg_n, segs, cap, &aborted, &truncated);
  if (truncated) return fail(H3D_ERR_RANGE, "an inlier list did not fit the device index pool");
  if (rc_ref) *rc_ref = aborted ? 1 : 0;
  return H3D_OK;
}

extern "C" int h3d_pipeline_set_pose(h3d_pipeline *p, const h3d_pose *pose) {
  if (!p) return fail(H3D_ERR_ARG, "bad argument");
  p->f.seg_pose_on = pose != nullptr;
  if (pose) p->f.seg_pose = *pose;
  return H3D_OK;
}

extern "C" int h3d_segment_test(const double *y_xyz, size_t n, const double shift[3], const double a[3],
                                const double b[3], const h3d_segment_config *cfg, h3d_segment *out) {
  return h3d_segment_test_pose(y_xyz, n, shift, a, b, cfg, nullptr, out);
}

extern "C" int h3d_segment_test_pose(const double *y_xyz, size_t n, const double shift[3], const double a[3],
                                     const double b[3], const h3d_segment_config *cfg, const h3d_pose *pose,
                                     h3d_segment *out) {
  if (!y_xyz || !n || !shift || !a || !b || !out || n > 0x7FFFFFF0u) return fail(H3D_ERR_ARG, "bad argument");
  if (!segment_config_ok(cfg)) return fail(H3D_ERR_ARG, "radius_sizes needs 1..8 entries");
  double *pts = nullptr, *scr = nullptr;
  h3d_iter *rec = nullptr;
  h3d_segment *seg = nullptr;
  auto cleanup = [&]() { cudaFree(pts); cudaFree(scr); cudaFree(rec); cudaFree(seg); };
  h3d_iter r;
  memset(&r, 0, sizeof r);
  r.status = 0; r.n2 = (uint32_t)n; r.off2 = 0;
  for (int c = 0; c < 3; ++c) { r.a2[c] = a[c]; r.b2[c] = b[c]; }
  cudaError_t e = cudaMalloc(&pts, n * 3 * sizeof(double));
  if (e == cudaSuccess) e = cudaMalloc(&scr, n * 5 * sizeof(double));
  if (e == cudaSuccess) e = cudaMalloc(&rec, sizeof r);
  if (e == cudaSuccess) e = cudaMalloc(&seg, sizeof *out);
  if (e == cudaSuccess) e = cudaMemcpy(pts, y_xyz, n * 3 * sizeof(double), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(rec, &r, sizeof r, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) {
    SegArgs sa;
    sa.iters = rec; sa.n_iters = nullptr; sa.pool = nullptr; sa.pool_cap = 0;
    sa.src = pts; sa.stride = 3; sa.src_is_float = 0;
    for (int c = 0; c < 3; ++c) { sa.ns[c] = 0.0; sa.shift[c] = shift[c]; }   // y - 0.0 == y
    sa.ns_dev = nullptr;
    sa.cfg = *cfg; sa.scr = scr; sa.scr_n = n; sa.out = seg;
    sa.pose_on = pose ? 1 : 0;
    if (pose) sa.pose = *pose; else memset(&sa.pose, 0, sizeof sa.pose);
    segment_tests_kernel<<<1, kSegThreads>>>(sa);
    note_launches(1);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpy(out, seg, sizeof *out, cudaMemcpyDeviceToHost);
  cleanup();
  if (e != cudaSuccess) return fail(H3D_ERR_CUDA, cudaGetErrorString(e));
  return H3D_OK;
}

extern "C" int h3d_pipeline_run_cloud2(h3d_pipeline *p, const uint8_t *data, size_t data_bytes,
                                       const h3d_cloud2_layout *L, h3d_iter *iters, uint32_t *pool, size_t pool_cap,
                                       h3d_result *res) {
  if (!p || !L || !res) return fail(H3D_ERR_ARG, "bad argument");
  if (L->is_bigendian) return fail(H3D_ERR_ARG, "big-endian PointCloud2 is not supported");
  const size_t n = (size_t)L->width * L->height;
  if (n > p->cfg.max_raw_points) return fail(H3D_ERR_ARG, "frame larger than max_raw_points");
  const uint32_t off[3] = {L->off_x, L->off_y, L->off_z};
  for (uint32_t o : off)
    if ((o & 3u) || (size_t)o + 4 > L->point_step) return fail(H3D_ERR_ARG, "x/y/z must be 4-byte aligned FLOAT32 fields inside point_step");
  if ((L->point_step & 3u) || (L->row_step & 3u) || (size_t)L->width * L->point_step > L->row_step)
    return fail(H3D_ERR_ARG, "bad point_step / row_step");
  if (n && (!data || (size_t)L->height * L->row_step > data_bytes)) return fail(H3D_ERR_ARG, "data shorter than height * row_step");
  H3D_CUDA(cudaSetDevice(p->device));
  cudaStream_t s = p->f.hg->stream;
  const size_t bytes = (size_t)L->height * L->row_step;
  if (bytes > p->msg_cap) {
    H3D_CUDA(cudaStreamSynchronize(s));
    cudaFree(p->msg_dev); p->msg_dev = nullptr; p->msg_cap = 0;
    H3D_CUDA(cudaMalloc(&p->msg_dev, bytes + 16));
    p->msg_cap = bytes;
  }
  if (bytes) H3D_CUDA(cudaMemcpyAsync(p->msg_dev, data, bytes, cudaMemcpyHostToDevice, s));
  const bool in_place = L->off_y == L->off_x + 4 && L->off_z == L->off_x + 8 &&
                        (L->height <= 1 || L->row_step == L->width * L->point_step);
  if (in_place)   // point i starts at msg_dev + off_x + i * point_step: the kernels' (pointer, stride) form
    return h3d_pipeline_run_dev(p, reinterpret_cast<const float *>(p->msg_dev + L->off_x), n, L->point_step / 4, iters, pool,
                                pool_cap, res);
  if (n) {
    cloud2_gather_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(p->msg_dev, *L, reinterpret_cast<float4 *>(p->raw_dev));
    note_launches(1);
    H3D_CUDA(cudaGetLastError());
  }
  return h3d_pipeline_run_dev(p, p->raw_dev, n, 4, iters, pool, pool_cap, res);
}

extern "C" int h3d_pipeline_filtered(h3d_pipeline *p, float *out_xyz, size_t cap, size_t *n_out) {
  if (!p || !n_out) return fail(H3D_ERR_ARG, "bad argument");
  if (!p->cfg.prefilter) return fail(H3D_ERR_ARG, "pipeline has no pre-filter stage");
  cudaStream_t s = p->f.hg->stream;
  if (p->fast.last_was_fast) {
    H3D_CUDA(cudaStreamSynchronize(s));
    *n_out = (size_t)reinterpret_cast<const FrameDev *>(p->fast.stage_host)->n;
  } else {
    H3D_CUDA(cudaMemcpyAsync(p->pf.st_host, p->pf.st, sizeof(K0State), cudaMemcpyDeviceToHost, s));
    H3D_CUDA(cudaStreamSynchronize(s));
    *n_out = (size_t)p->pf.st_host->n_out;
  }
  const size_t n_copy = *n_out < cap ? *n_out : cap;
  if (n_copy && out_xyz) H3D_CUDA(cudaMemcpy(out_xyz, p->filt_dev, n_copy * 3 * sizeof(float), cudaMemcpyDeviceToHost));
  return H3D_OK;
}

extern "C" int h3d_pipeline_profile(h3d_pipeline *p, int iters, unsigned long long *out, size_t cap, int *n_cta,
                                    int *n_slots) {
  if (!p) return fail(H3D_ERR_ARG, "bad argument");
  FrameCtx &f = p->f;
  if (n_slots) *n_slots = kProfSlots;
  if (out) {
    if (!f.prof_dev) return fail(H3D_ERR_ARG, "profiling is not enabled");
    if (n_cta) *n_cta = f.prof_grid;
    const size_t words = (size_t)f.prof_iters * f.prof_grid * kProfSlots;
    H3D_CUDA(cudaMemcpy(out, f.prof_dev, (words < cap ? words : cap) * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    return H3D_OK;
  }
  cudaFree(f.prof_dev); f.prof_dev = nullptr; f.prof_iters = 0;
  if (iters > 0) {
    f.prof_ctas = f.hg->dev.sm_count;
    H3D_CUDA(cudaMalloc(&f.prof_dev, (size_t)iters * f.prof_ctas * kProfSlots * sizeof(unsigned long long)));
    f.prof_iters = iters;
  }
  if (n_cta) *n_cta = f.prof_ctas;
  return H3D_OK;
}

static int replay_impl(h3d_pipeline *const *pipes, size_t n_pipes, const float *const *raw, const size_t *n_points,
                       size_t stride, size_t n_frames, int frames_on_device, h3d_iter *iters, size_t max_iters,
                       uint32_t *pool, size_t pool_cap, h3d_result *results, h3d_segment *segments,
                       size_t max_segments, uint32_t *n_segments, int *rc_ref, const h3d_pose *poses);

// Batched replay (BASELINE configs[2]): n_pipes host threads, one per pipeline, pull frames from a
// shared counter.  Frame i: raw[i] (n_points[i] points, `stride` floats apart; device pointers when
// frames_on_device), results[i], iters + i*max_iters, pool + i*pool_cap.
extern "C" int h3d_replay_run(h3d_pipeline *const *pipes, size_t n_pipes, const float *const *raw,
                              const size_t *n_points, size_t stride, size_t n_frames, int frames_on_device,
                              h3d_iter *iters, size_t max_iters, uint32_t *pool, size_t pool_cap,
                              h3d_result *results) {
  return h3d_replay_run_segments(pipes, n_pipes, raw, n_points, stride, n_frames, frames_on_device, iters, max_iters,
                                 pool, pool_cap, results, nullptr, 0, nullptr, nullptr);
}

extern "C" int h3d_replay_run_segments(h3d_pipeline *const *pipes, size_t n_pipes, const float *const *raw,
                                       const size_t *n_points, size_t stride, size_t n_frames, int frames_on_device,
                                       h3d_iter *iters, size_t max_iters, uint32_t *pool, size_t pool_cap,
                                       h3d_result *results, h3d_segment *segments, size_t max_segments,
                                       uint32_t *n_segments, int *rc_ref) {
  return replay_impl(pipes, n_pipes, raw, n_points, stride, n_frames, frames_on_device, iters, max_iters, pool,
                     pool_cap, results, segments, max_segments, n_segments, rc_ref, nullptr);
}

// Replay with one sensor pose per frame and the world model fed in frame order (node.cpp:286-306:
// hough3dlines -> drone2WorldSeg -> heighSegmentCutoff -> segFiltering).  Frames run in flight on the
// pipelines; segFiltering is sequential across frames, so it runs on the collected records afterwards.
extern "C" int h3d_replay_run_world(h3d_pipeline *const *pipes, size_t n_pipes, const float *const *raw,
                                    const size_t *n_points, size_t stride, size_t n_frames, int frames_on_device,
                                    const h3d_pose *poses, h3d_result *results, h3d_segment *segments,
                                    size_t max_segments, uint32_t *n_segments, int *rc_ref, h3d_world *world) {
  if (!poses || !segments || !n_segments || !world) return fail(H3D_ERR_ARG, "bad argument");
  int rc = replay_impl(pipes, n_pipes, raw, n_points, stride, n_frames, frames_on_device, nullptr, 0, nullptr, 0,
                       results, segments, max_segments, n_segments, rc_ref, poses);
  if (rc != H3D_OK) return rc;
  for (size_t i = 0; i < n_frames; ++i) {
    if (n_segments[i] > max_segments) return fail(H3D_ERR_RANGE, "max_segments too small for a frame");
    rc = h3d_world_update_frame(world, segments + i * max_segments, n_segments[i], nullptr);
    if (rc != H3D_OK) return rc;
  }
  return H3D_OK;
}

static int replay_impl(h3d_pipeline *const *pipes, size_t n_pipes, const float *const *raw, const size_t *n_points,
                       size_t stride, size_t n_frames, int frames_on_device, h3d_iter *iters, size_t max_iters,
                       uint32_t *pool, size_t pool_cap, h3d_result *results, h3d_segment *segments,
                       size_t max_segments, uint32_t *n_segments, int *rc_ref, const h3d_pose *poses) {
  if (!pipes || n_pipes == 0 || (!raw && n_frames) || (!n_points && n_frames) || !results)
    return fail(H3D_ERR_ARG, "bad argument");
  if (segments && !n_segments) return fail(H3D_ERR_ARG, "segments needs n_segments");
  for (size_t k = 0; k < n_pipes; ++k) if (!pipes[k]) return fail(H3D_ERR_ARG, "null pipeline");
  std::atomic<size_t> next{0};
  std::atomic<int> first_rc{H3D_OK};
  std::mutex err_mu;
  std::string err_text;
  auto worker = [&](size_t t) {
    for (;;) {
      const size_t i = next.fetch_add(1);
      if (i >= n_frames || first_rc.load() != H3D_OK) break;
      h3d_iter *it = iters ? iters + i * max_iters : nullptr;
      uint32_t *pl = pool ? pool + i * pool_cap : nullptr;
      if (poses) h3d_pipeline_set_pose(pipes[t], &poses[i]);
      int rc = frames_on_device
                         ? h3d_pipeline_run_dev(pipes[t], raw[i], n_points[i], stride, it, pl, pool ? pool_cap : 0, &results[i])
                         : h3d_pipeline_run(pipes[t], raw[i], n_points[i], stride, it, pl, pool ? pool_cap : 0, &results[i]);
      if (rc == H3D_OK && segments) {
        size_t m = 0;
        int ref = 0;
        rc = h3d_pipeline_segments(pipes[t], segments + i * max_segments, max_segments, &m, &ref);
        n_segments[i] = (uint32_t)m;
        if (rc_ref) rc_ref[i] = ref;
      }
      if (rc != H3D_OK) {
        int expected = H3D_OK;
        if (first_rc.compare_exchange_strong(expected, rc)) {
          std::lock_guard<std::mutex> lock(err_mu);
          err_text = h3d_last_error();      // thread-local text of the failing worker
        }
        break;
      }
    }
  };
  // Fewer host threads than pipelines (h3d_set_replay_threads; several ranks on a box with few cores): thread t
  // drives pipelines t, t + T, ...  It submits a frame on each (the fast path enqueues a whole frame without
  // waiting), then polls their last events and completes / refills whichever is done.
  auto report = [&](int rc) {
    int expected = H3D_OK;
    if (first_rc.compare_exchange_strong(expected, rc)) {
      std::lock_guard<std::mutex> lock(err_mu);
      err_text = h3d_last_error();
    }
  };
  auto worker_mux = [&](size_t t, size_t T) {
    struct Slot { h3d_pipeline *p; size_t frame; const float *raw_dev; bool busy; };
    std::vector<Slot> slots;
    for (size_t q = t; q < n_pipes; q += T) slots.push_back(Slot{pipes[q], 0, nullptr, false});
    size_t active = 0;
    bool feed = true;
    auto finish = [&](Slot &sl, size_t i) -> int {   // what follows a frame's results
      if (!segments) return H3D_OK;
      size_t m = 0;
      int ref = 0;
      const int rc = h3d_pipeline_segments(sl.p, segments + i * max_segments, max_segments, &m, &ref);
      n_segments[i] = (uint32_t)m;
      if (rc_ref) rc_ref[i] = ref;
      return rc;
    };
    auto refill = [&](Slot &sl) -> int {
      while (feed) {
        const size_t i = next.fetch_add(1);
        if (i >= n_frames || first_rc.load() != H3D_OK) { feed = false; break; }
        h3d_iter *it = iters ? iters + i * max_iters : nullptr;
        uint32_t *pl = pool ? pool + i * pool_cap : nullptr;
        if (poses) h3d_pipeline_set_pose(sl.p, &poses[i]);
        bool async = false;
        int rc = pipeline_submit(sl.p, raw[i], frames_on_device, n_points[i], stride, pool ? pool_cap : 0, &sl.raw_dev, &async);
        if (rc != H3D_OK) return rc;
        if (async) { sl.frame = i; sl.busy = true; ++active; return H3D_OK; }
        rc = frames_on_device
                 ? h3d_pipeline_run_dev(sl.p, raw[i], n_points[i], stride, it, pl, pool ? pool_cap : 0, &results[i])
                 : h3d_pipeline_run(sl.p, raw[i], n_points[i], stride, it, pl, pool ? pool_cap : 0, &results[i]);
        if (rc == H3D_OK) rc = finish(sl, i);
        if (rc != H3D_OK) return rc;
      }
      return H3D_OK;
    };
    int rc = H3D_OK;
    for (auto &sl : slots) if ((rc = refill(sl)) != H3D_OK) break;
    const bool spin = g_wait_mode.load(std::memory_order_relaxed) == 0;
    while (rc == H3D_OK && active) {
      bool progressed = false;
      for (auto &sl : slots) {
        if (!sl.busy) continue;
        const int ready = pipeline_ready(sl.p);
        if (ready == 0) continue;
        const size_t i = sl.frame;
        sl.busy = false; --active; progressed = true;
        rc = pipeline_complete(sl.p, sl.raw_dev, n_points[i], stride, iters ? iters + i * max_iters : nullptr,
                               pool ? pool + i * pool_cap : nullptr, pool ? pool_cap : 0, &results[i], ready == 1);
        if (rc == H3D_OK) rc = finish(sl, i);
        if (rc == H3D_OK) rc = refill(sl);
        if (rc != H3D_OK) break;
      }
      if (!progressed && !spin) std::this_thread::yield();
    }
    if (rc != H3D_OK) {
      report(rc);
      for (auto &sl : slots) if (sl.busy) cudaStreamSynchronize(sl.p->f.hg->stream);   // nothing left in flight
    }
  };
  size_t T = (size_t)g_replay_threads.load(std::memory_order_relaxed);
  std::vector<std::thread> threads;
  if (T == 0 || T >= n_pipes) {
    threads.reserve(n_pipes);
    for (size_t t = 1; t < n_pipes; ++t) threads.emplace_back(worker, t);
    worker(0);
  } else {
    threads.reserve(T);
    for (size_t t = 1; t < T; ++t) threads.emplace_back(worker_mux, t, T);
    worker_mux(0, T);
  }
  for (auto &th : threads) th.join();
  if (first_rc.load() != H3D_OK) { set_error(err_text); return first_rc.load(); }
  return H3D_OK;
}

// Host threads of a replay: 0 = one per pipeline (default); n < pipelines = n threads, each driving several
// pipelines (for boxes with fewer cores than ranks x pipelines).  Returns the previous setting.
extern "C" int h3d_set_replay_threads(int n) { return g_replay_threads.exchange(n > 0 ? n : 0); }

extern "C" int h3d_set_wait_mode(int mode) {
  return g_wait_mode.exchange(mode == 1 ? 1 : mode == 2 ? 2 : 0);
}

extern "C" int h3d_pipeline_set_sm_share(h3d_pipeline *p, int n_ctas) {
  if (!p || n_ctas < 0) return fail(H3D_ERR_ARG, "bad argument");
  p->f.fused_ctas = n_ctas;
  return H3D_OK;
}

extern "C" int h3d_pipeline_timing(h3d_pipeline *p, float us[4]) {
  if (!p || !us) return fail(H3D_ERR_ARG, "bad argument");
  for (int k = 0; k < 4; ++k) us[k] = p->f.us[k];
  return H3D_OK;
}

// =====================================================================================
// PointCloud entry points on host arrays (lib/pointcloud.cpp:18-128, wrap.h:121-150): each is
// upload -> kernel -> download; the fused loop above never goes through them.
// =====================================================================================
namespace {

// scratch arena of one calling thread
struct Scratch {
  void *dev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  size_t bytes[6] = {0, 0, 0, 0, 0, 0};
  ~Scratch() { for (auto p : dev) cudaFree(p); }
};
int scratch_get(Scratch &s, int slot, size_t bytes, void **out) {
  if (bytes > s.bytes[slot]) {
    cudaFree(s.dev[slot]); s.dev[slot] = nullptr; s.bytes[slot] = 0;
    size_t cap = 1 << 16;
    while (cap < bytes) cap *= 2;
    H3D_CUDA(cudaMalloc(&s.dev[slot], cap));
    s.bytes[slot] = cap;
  }
  *out = s.dev[slot];
  return H3D_OK;
}
Scratch &scratch() { static thread_local Scratch s; return s; }

}  // namespace

// getMinMax3D on an AoS cloud (one CTA; min/max are order independent)
__global__ void __launch_bounds__(1024, 1) aos_minmax_kernel(const double *__restrict__ p, int n, double *out6) {
  __shared__ double red[6][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  double mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
  for (int i = threadIdx.x; i < n; i += 1024)
    for (int c = 0; c < 3; ++c) { const double v = p[3 * (size_t)i + c]; mn[c] = fmin(mn[c], v); mx[c] = fmax(mx[c], v); }
  for (int c = 0; c < 3; ++c) {
    for (int o = 16; o; o >>= 1) {
      mn[c] = fmin(mn[c], __shfl_xor_sync(0xFFFFFFFFu, mn[c], o));
      mx[c] = fmax(mx[c], __shfl_xor_sync(0xFFFFFFFFu, mx[c], o));
    }
    if (lane == 0) { red[c][warp] = mn[c]; red[3 + c][warp] = mx[c]; }
  }
  __syncthreads();
  if (warp == 0) {
    for (int c = 0; c < 3; ++c) {
      double a = red[c][lane], b = red[3 + c][lane];
      for (int o = 16; o; o >>= 1) {
        a = fmin(a, __shfl_xor_sync(0xFFFFFFFFu, a, o));
        b = fmax(b, __shfl_xor_sync(0xFFFFFFFFu, b, o));
      }
      if (lane == 0) { out6[c] = n ? a : 0.0; out6[3 + c] = n ? b : 0.0; }
    }
  }
}

__global__ void aos_shift_kernel(double *p, int n, double sx, double sy, double sz) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    p[3 * (size_t)i] = __dsub_rn(p[3 * (size_t)i], sx);
    p[3 * (size_t)i + 1] = __dsub_rn(p[3 * (size_t)i + 1], sy);
    p[3 * (size_t)i + 2] = __dsub_rn(p[3 * (size_t)i + 2], sz);
  }
}

__global__ void aos_to_soa_kernel(const double *__restrict__ p, int n, double *px, double *py, double *pz) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    px[i] = p[3 * (size_t)i]; py[i] = p[3 * (size_t)i + 1]; pz[i] = p[3 * (size_t)i + 2];
  }
}

__global__ void __launch_bounds__(192, 1) soa_lsq_kernel(const double *px, const double *py, const double *pz, int n,
                                                         LsqOut *out) {
  __shared__ LsqShared sh;
  __shared__ LsqOut o;
  lsq_cta([&](int k, int c) { return (c == 0 ? px : c == 1 ? py : pz)[k]; }, n, sh, &o);
  if (threadIdx.x == 0) *out = o;
}

__global__ void aos_mask_kernel(const double *__restrict__ p, int n, double a0, double a1, double a2, double b0,
                                double b1, double b2, double dx, unsigned char *flags) {
  const double a[3] = {a0, a1, a2}, b[3] = {b0, b1, b2};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    flags[i] = point_near_line(p[3 * (size_t)i], p[3 * (size_t)i + 1], p[3 * (size_t)i + 2], a, b, dx) ? 1 : 0;
}

// ordered compaction of flagged indices (one CTA, chunked); count -> *n_out
__global__ void __launch_bounds__(1024, 1) flags_to_index_kernel(const unsigned char *__restrict__ flags, int n,
                                                                 unsigned *idx, int *n_out) {
  __shared__ int warp_sum[32];
  __shared__ int base_s;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) base_s = 0;
  __syncthreads();
  for (int c0 = 0; c0 < n; c0 += 1024) {
    const int i = c0 + tid;
    const bool keep = i < n && flags[i];
    const unsigned bal = __ballot_sync(0xFFFFFFFFu, keep);
    if (lane == 0) warp_sum[warp] = __popc(bal);
    __syncthreads();
    int off = base_s;
    for (int q = 0; q < warp; ++q) off += warp_sum[q];
    if (keep) idx[off + __popc(bal & ((1u << lane) - 1u))] = (unsigned)i;
    __syncthreads();
    if (tid == 0) { int t = 0; for (int q = 0; q < 32; ++q) t += warp_sum[q]; base_s += t; }
    __syncthreads();
  }
  if (tid == 0) *n_out = base_s;
}

// PointCloud::removePoints (lib/pointcloud.cpp:109-128): greedy ordered matching with exact ==.
// One warp: for the current y_j it scans X 32 points at a time for the first equal point; every
// x passed on the way is kept and, since exactly j points were dropped before it, lands at
// i - j (in place: the target was consumed earlier).
__global__ void __launch_bounds__(32, 1) remove_points_kernel(double *x, int nx, const double *__restrict__ y, int ny,
                                                              int *n_out) {
  const int lane = threadIdx.x;
  int i = 0, j = 0;
  while (i < nx && j < ny) {
    const double y0 = y[3 * (size_t)j], y1 = y[3 * (size_t)j + 1], y2 = y[3 * (size_t)j + 2];
    const int k = i + lane;
    double v0 = 0, v1 = 0, v2 = 0;
    bool eq = false;
    if (k < nx) {
      v0 = x[3 * (size_t)k]; v1 = x[3 * (size_t)k + 1]; v2 = x[3 * (size_t)k + 2];
      eq = (v0 == y0) && (v1 == y1) && (v2 == y2);
    }
    const unsigned m = __ballot_sync(0xFFFFFFFFu, eq);
    const int first = m ? __ffs(m) - 1 : 32;          // lanes below `first` are kept
    __syncwarp();
    if (k < nx && lane < first) { const size_t d = (size_t)(k - j); x[3 * d] = v0; x[3 * d + 1] = v1; x[3 * d + 2] = v2; }
    __syncwarp();
    if (m) { i += first + 1; ++j; } else { i += 32; }
  }
  for (int k = i + lane; k < nx; k += 32) {           // rest after the end of Y, in order
    const double v0 = x[3 * (size_t)k], v1 = x[3 * (size_t)k + 1], v2 = x[3 * (size_t)k + 2];
    __syncwarp();
    const size_t d = (size_t)(k - j);
    x[3 * d] = v0; x[3 * d + 1] = v1; x[3 * d + 2] = v2;
    __syncwarp();
  }
  if (lane == 0) *n_out = nx - j;
}

static int upload_aos(const double *xyz, size_t n, int slot, double **dev) {
  void *p = nullptr;
  H3D_TRY(scratch_get(scratch(), slot, (n ? n : 1) * 3 * sizeof(double), &p));
  if (n) H3D_CUDA(cudaMemcpy(p, xyz, n * 3 * sizeof(double), cudaMemcpyHostToDevice));
  *dev = static_cast<double *>(p);
  return H3D_OK;
}

extern "C" int h3d_min_max(const double *xyz, size_t n, double mn[3], double mx[3]) {
  if ((!xyz && n) || !mn || !mx || n > 0x7FFFFFF0u) return fail(H3D_ERR_ARG, "bad argument");
  double *p = nullptr; void *o = nullptr;
  H3D_TRY(upload_aos(xyz, n, 0, &p));
  H3D_TRY(scratch_get(scratch(), 5, 64, &o));
  aos_minmax_kernel<<<1, 1024>>>(p, (int)n, static_cast<double *>(o));
  double h[6];
  H3D_CUDA(cudaMemcpy(h, o, sizeof h, cudaMemcpyDeviceToHost));
  for (int c = 0; c < 3; ++c) { mn[c] = h[c]; mx[c] = h[3 + c]; }
  return H3D_OK;
}

extern "C" int h3d_shift_to_origin(double *xyz, size_t n, double shift[3]) {
  if ((!xyz && n) || !shift || n > 0x7FFFFFF0u) return fail(H3D_ERR_ARG, "bad argument");
  double *p = nullptr; void *o = nullptr;
  H3D_TRY(upload_aos(xyz, n, 0, &p));
  H3D_TRY(scratch_get(scratch(), 5, 64, &o));
  aos_minmax_kernel<<<1, 1024>>>(p, (int)n, static_cast<double *>(o));
  double h[6];
  H3D_CUDA(cudaMemcpy(h, o, sizeof h, cudaMemcpyDeviceToHost));
  double ns[3];
  for (int c = 0; c < 3; ++c) ns[c] = (h[c] + h[3 + c]) / 2.0;   // (p1 + p2) / 2.0
  if (n) {
    aos_shift_kernel<<<(unsigned)((n + 255) / 256), 256>>>(p, (int)n, ns[0], ns[1], ns[2]);
    H3D_CUDA(cudaMemcpy(xyz, p, n * 3 * sizeof(double), cudaMemcpyDeviceToHost));
  }
  for (int c = 0; c < 3; ++c) shift[c] = shift[c] + ns[c];
  return H3D_OK;
}

static int lsq_on_host_cloud(const double *xyz, size_t n, LsqOut *res) {
  double *p = nullptr; void *soa = nullptr, *o = nullptr;
  H3D_TRY(upload_aos(xyz, n, 0, &p));
  H3D_TRY(scratch_get(scratch(), 1, (n ? n : 1) * 3 * sizeof(double), &soa));
  H3D_TRY(scratch_get(scratch(), 5, sizeof(LsqOut), &o));
  double *px = static_cast<double *>(soa), *py = px + n, *pz = py + n;
  if (n) aos_to_soa_kernel<<<(unsigned)((n + 255) / 256), 256>>>(p, (int)n, px, py, pz);
  soa_lsq_kernel<<<1, 192>>>(px, py, pz, (int)n, static_cast<LsqOut *>(o));
  H3D_CUDA(cudaMemcpy(res, o, sizeof(LsqOut), cudaMemcpyDeviceToHost));
  return H3D_OK;
}

extern "C" int h3d_mean_value(const double *xyz, size_t n, double mean[3]) {
  if ((!xyz && n) || !mean || n > 0x7FFFFFF0u) return fail(H3D_ERR_ARG, "bad argument");
  LsqOut r;
  H3D_TRY(lsq_on_host_cloud(xyz, n, &r));
  for (int c = 0; c < 3; ++c) mean[c] = r.a[c];
  return H3D_OK;
}

extern "C" int h3d_orthogonal_lsq(const double *xyz, size_t n, double a[3], double b[3], double *rc) {
  if ((!xyz && n) || !a || !b || !rc || n > 0x7FFFFFF0u) return fail(H3D_ERR_ARG, "bad argument");
  LsqOut r;
  H3D_TRY(lsq_on_host_cloud(xyz, n, &r));
  for (int c = 0; c < 3; ++c) { a[c] = r.a[c]; b[c] = r.b[c]; }
  *rc = r.rc;
  return H3D_OK;
}

extern "C" int h3d_points_close_to_line(const double *xyz, size_t n, const double a[3], const double b[3], double dx,
                                        uint32_t *idx_out, size_t *n_out) {
  if ((!xyz && n) || !a || !b || !n_out || (!idx_out && n) || n > 0x7FFFFFF0u) return fail(H3D_ERR_ARG, "bad argument");
  *n_out = 0;
  if (n == 0) return H3D_OK;
  double *p = nullptr; void *flags = nullptr, *idx = nullptr, *cnt = nullptr;
  H3D_TRY(upload_aos(xyz, n, 0, &p));
  H3D_TRY(scratch_get(scratch(), 2, n, &flags));
  H3D_TRY(scratch_get(scratch(), 3, n * sizeof(unsigned), &idx));
  H3D_TRY(scratch_get(scratch(), 5, 64, &cnt));
  aos_mask_kernel<<<(unsigned)((n + 255) / 256), 256>>>(p, (int)n, a[0], a[1], a[2], b[0], b[1], b[2], dx,
                                                        static_cast<unsigned char *>(flags));
  flags_to_index_kernel<<<1, 1024>>>(static_cast<unsigned char *>(flags), (int)n, static_cast<unsigned *>(idx),
                                     static_cast<int *>(cnt));
  int k = 0;
  H3D_CUDA(cudaMemcpy(&k, cnt, sizeof k, cudaMemcpyDeviceToHost));
  if (k) H3D_CUDA(cudaMemcpy(idx_out, idx, (size_t)k * sizeof(unsigned), cudaMemcpyDeviceToHost));
  *n_out = (size_t)k;
  return H3D_OK;
}

extern "C" int h3d_remove_points(double *x_xyz, size_t nx, const double *y_xyz, size_t ny, size_t *n_out) {
  if ((!x_xyz && nx) || (!y_xyz && ny) || !n_out || nx > 0x7FFFFFF0u || ny > 0x7FFFFFF0u)
    return fail(H3D_ERR_ARG, "bad argument");
  *n_out = nx;
  if (ny == 0 || nx == 0) return H3D_OK;   // lib/pointcloud.cpp:111
  double *x = nullptr, *y = nullptr; void *cnt = nullptr;
  H3D_TRY(upload_aos(x_xyz, nx, 0, &x));
  H3D_TRY(upload_aos(y_xyz, ny, 1, &y));
  H3D_TRY(scratch_get(scratch(), 5, 64, &cnt));
  remove_points_kernel<<<1, 32>>>(x, (int)nx, y, (int)ny, static_cast<int *>(cnt));
  int k = 0;
  H3D_CUDA(cudaMemcpy(&k, cnt, sizeof k, cudaMemcpyDeviceToHost));
  if (k) H3D_CUDA(cudaMemcpy(x_xyz, x, (size_t)k * 3 * sizeof(double), cudaMemcpyDeviceToHost));
  *n_out = (size_t)k;
  return H3D_OK;
}

extern "C" int h3d_set_fused(int on) {
  const int old = g_use_fused ? 1 : 0;
  g_use_fused = on != 0;
  return old;
}

extern "C" int h3d_set_vote_track(int mode) { return g_vote_track.exchange(mode == 0 ? 0 : mode == 1 ? 1 : 2); }

// measurement: globaltimer stamps of the last fast-path frame's K0 cluster kernel (start, after each of its six
// cluster barriers, end), nanoseconds
extern "C" int h3d_pipeline_k0_stamps(h3d_pipeline *p, unsigned long long t_ns[10]) {
  if (!p || !t_ns) return fail(H3D_ERR_ARG, "bad argument");
  if (!p->fast.on || !p->fast.last_was_fast) return fail(H3D_ERR_ARG, "the last frame did not run on the fast path");
  const FrameDev &F = *reinterpret_cast<const FrameDev *>(p->fast.stage_host);
  for (int k = 0; k < 10; ++k) t_ns[k] = F.t_ns[k];
  return H3D_OK;
}

extern "C" int h3d_set_fast_frames(int on) { return g_fast_frames.exchange(on != 0 ? 1 : 0); }

extern "C" int h3d_set_debug_poison(int on) { return g_debug_poison.exchange(on != 0 ? 1 : 0); }

extern "C" int h3d_set_stream_mode(int mode) { return g_stream_mode.exchange(mode == 1 ? 1 : mode == 2 ? 2 : 0); }

extern "C" int h3d_set_sliced_vote(int on) {
  const int old = g_sliced_vote ? 1 : 0;
  g_sliced_vote = on != 0;
  return old;
}

extern "C" int h3d_set_big_kernels(int on) {
  const int old = g_big_kernels ? 1 : 0;
  g_big_kernels = on != 0;
  return old;
}
