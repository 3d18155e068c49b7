// h3d_prefilter.cuh -- K0: the node's point-level pre-filter (sm_100a).
//   cloudFiltering()  node.cpp:386-421 = pcl::PassThrough x, y, z (inclusive limits, drops
//   non-finite fields) then pcl::VoxelGrid (leaf^3, one centroid per voxel, ascending voxel
//   index).  PCL 1.10 is not vendored by the reference: the semantics are restated from the
//   published algorithm (oracle/h3d_oracle.c h3do_prefilter is the parity target, PARITY
//   UNPINNED against PCL itself).
//
// Sort-by-voxel-key as a counting sort over the dense voxel grid of the filtered cloud's
// bounding box (<= 45 x 91 x 91 cells for the ToF window):
//   k0_bbox      window test + float bbox of the survivors (ordered-int atomics)
//   k0_setup     min_b / div_b / strides exactly as VoxelGrid computes them
//   k0_count     voxel key per survivor, histogram
//   scan (3 kernels)  exclusive scans of the counts (list offsets) and of the non-empty
//                flags (output slots): the output order is ascending voxel index
//   k0_scatter   raw point indices into their voxel's list (any order)
//   k0_centroid  per voxel: order its few indices (input order = the canonical summation
//                order), sequential float sums, divide by the float count
// Raw points are read with one 16-byte load when the stride is 4 floats (pcl::PointXYZ).
#ifndef H3D_PREFILTER_CUH_
#define H3D_PREFILTER_CUH_

#include <cuda_runtime.h>
#include <limits.h>

namespace h3d {

struct K0State {
  // bbox of the survivors as ordered ints (float bits mapped monotonically)
  int mn[3], mx[3];
  int n_in_window;
  // VoxelGrid geometry
  float inv_leaf;
  int min_b[3];
  int div_b[3];
  int mul[3];
  int n_vox;        // div_b[0]*div_b[1]*div_b[2]
  int overflow;     // 1: index would overflow int32 / exceed the grid capacity
  int n_out;        // number of non-empty voxels
};

__device__ __forceinline__ int float_to_ordered(float f) {
  const int i = __float_as_int(f);
  return i >= 0 ? i : i ^ 0x7FFFFFFF;
}
__device__ __forceinline__ float ordered_to_float(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7FFFFFFF); }

__device__ __forceinline__ float3 load_xyz(const float *__restrict__ in, int i, int stride) {
  if (stride == 4 && (reinterpret_cast<size_t>(in) & 15) == 0) {   // 16-byte aligned quadruples
    const float4 v = reinterpret_cast<const float4 *>(in)[i];
    return make_float3(v.x, v.y, v.z);
  }
  const float *p = in + (size_t)i * stride;
  return make_float3(p[0], p[1], p[2]);
}

struct K0Window { float lo[3], hi[3]; };

// PassThrough x then y then z (node.cpp:394-407) is one conjunctive, order-preserving test
__device__ __forceinline__ bool in_window(const float3 p, const K0Window &w) {
  if (!(isfinite(p.x) && isfinite(p.y) && isfinite(p.z))) return false;
  return !(p.x > w.hi[0] || p.x < w.lo[0]) && !(p.y > w.hi[1] || p.y < w.lo[1]) && !(p.z > w.hi[2] || p.z < w.lo[2]);
}

__global__ void k0_init_kernel(K0State *st) {
  for (int c = 0; c < 3; ++c) { st->mn[c] = INT_MAX; st->mx[c] = INT_MIN; }
  st->n_in_window = 0; st->overflow = 0; st->n_out = 0; st->n_vox = 0;
}

__global__ void k0_bbox_kernel(const float *__restrict__ in, int n, int stride, K0Window w, K0State *st) {
  int mn[3] = {INT_MAX, INT_MAX, INT_MAX}, mx[3] = {INT_MIN, INT_MIN, INT_MIN};
  int cnt = 0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float3 p = load_xyz(in, i, stride);
    if (!in_window(p, w)) continue;
    const int o[3] = {float_to_ordered(p.x), float_to_ordered(p.y), float_to_ordered(p.z)};
    for (int c = 0; c < 3; ++c) { mn[c] = min(mn[c], o[c]); mx[c] = max(mx[c], o[c]); }
    ++cnt;
  }
  for (int c = 0; c < 3; ++c) {
    mn[c] = __reduce_min_sync(0xFFFFFFFFu, mn[c]);
    mx[c] = __reduce_max_sync(0xFFFFFFFFu, mx[c]);
  }
  cnt = __reduce_add_sync(0xFFFFFFFFu, cnt);
  if ((threadIdx.x & 31) == 0 && cnt) {
    for (int c = 0; c < 3; ++c) { atomicMin(&st->mn[c], mn[c]); atomicMax(&st->mx[c], mx[c]); }
    atomicAdd(&st->n_in_window, cnt);
  }
}

// VoxelGrid::applyFilter geometry: min_b = floor(min*inv), div_b = max_b - min_b + 1
__global__ void k0_setup_kernel(K0State *st, float leaf, int grid_cap) {
  if (threadIdx.x || blockIdx.x) return;
  const float inv = 1.0f / leaf;
  st->inv_leaf = inv;
  if (st->n_in_window == 0) { st->n_vox = 0; return; }
  long long div[3];
  for (int c = 0; c < 3; ++c) {
    const int lo = (int)floorf(ordered_to_float(st->mn[c]) * inv);
    const int hi = (int)floorf(ordered_to_float(st->mx[c]) * inv);
    st->min_b[c] = lo;
    div[c] = (long long)hi - lo + 1;
    st->div_b[c] = (int)div[c];
  }
  const long long total = div[0] * div[1] * div[2];
  if (total > (long long)INT_MAX || total > (long long)grid_cap) { st->overflow = 1; st->n_vox = 0; return; }
  st->mul[0] = 1; st->mul[1] = (int)div[0]; st->mul[2] = (int)(div[0] * div[1]);
  st->n_vox = (int)total;
}

__device__ __forceinline__ int voxel_key(const float3 p, const K0State &s) {
  const int i0 = (int)(floorf(p.x * s.inv_leaf) - (float)s.min_b[0]);
  const int i1 = (int)(floorf(p.y * s.inv_leaf) - (float)s.min_b[1]);
  const int i2 = (int)(floorf(p.z * s.inv_leaf) - (float)s.min_b[2]);
  return i0 * s.mul[0] + i1 * s.mul[1] + i2 * s.mul[2];
}

__global__ void k0_count_kernel(const float *__restrict__ in, int n, int stride, K0Window w, const K0State *st,
                                int *count) {
  if (st->n_vox == 0) return;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float3 p = load_xyz(in, i, stride);
    if (in_window(p, w)) atomicAdd(&count[voxel_key(p, *st)], 1);
  }
}

// ---- exclusive scan of (count, count > 0) over the voxel grid: reduce / spine / apply ----
constexpr int kScanBlock = 1024, kScanItems = 4;   // 4096 cells per CTA

__global__ void __launch_bounds__(kScanBlock) k0_scan_reduce_kernel(const int *__restrict__ count, const K0State *st,
                                                                    int2 *block_sums) {
  __shared__ int2 ws[32];
  const int n = st->n_vox;
  const int base = blockIdx.x * kScanBlock * kScanItems;
  if (base >= n) return;
  int a = 0, b = 0;
  for (int k = 0; k < kScanItems; ++k) {
    const int i = base + threadIdx.x * kScanItems + k;
    const int c = i < n ? count[i] : 0;
    a += c; b += c > 0;
  }
  a = __reduce_add_sync(0xFFFFFFFFu, a); b = __reduce_add_sync(0xFFFFFFFFu, b);
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = make_int2(a, b);
  __syncthreads();
  if (threadIdx.x < 32) {
    int2 v = ws[threadIdx.x];
    v.x = __reduce_add_sync(0xFFFFFFFFu, v.x); v.y = __reduce_add_sync(0xFFFFFFFFu, v.y);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = v;
  }
}

__global__ void __launch_bounds__(1024) k0_scan_spine_kernel(int2 *block_sums, K0State *st) {
  // exclusive scan of the block sums in place (one CTA, chunked)
  __shared__ int2 ws[32];
  __shared__ int2 carry;
  const int nb = (st->n_vox + kScanBlock * kScanItems - 1) / (kScanBlock * kScanItems);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry = make_int2(0, 0);
  __syncthreads();
  for (int c0 = 0; c0 < nb; c0 += 1024) {
    const int i = c0 + threadIdx.x;
    const int2 v = i < nb ? block_sums[i] : make_int2(0, 0);
    int2 inc = v;
    for (int o = 1; o < 32; o <<= 1) {
      const int x = __shfl_up_sync(0xFFFFFFFFu, inc.x, o), y = __shfl_up_sync(0xFFFFFFFFu, inc.y, o);
      if (lane >= o) { inc.x += x; inc.y += y; }
    }
    if (lane == 31) ws[warp] = inc;
    __syncthreads();
    int2 off = carry;
    for (int q = 0; q < warp; ++q) { off.x += ws[q].x; off.y += ws[q].y; }
    if (i < nb) block_sums[i] = make_int2(off.x + inc.x - v.x, off.y + inc.y - v.y);
    __syncthreads();
    if (threadIdx.x == 1023) carry = make_int2(off.x + inc.x, off.y + inc.y);
    __syncthreads();
  }
  if (threadIdx.x == 0) st->n_out = carry.y;
}

// start[i] = exclusive sum of counts, slot[i] = exclusive count of non-empty voxels
__global__ void __launch_bounds__(kScanBlock) k0_scan_apply_kernel(const int *__restrict__ count, const K0State *st,
                                                                   const int2 *__restrict__ block_sums,
                                                                   int *start, int *slot, int *cursor) {
  __shared__ int2 ws[32];
  const int n = st->n_vox;
  const int base = blockIdx.x * kScanBlock * kScanItems;
  if (base >= n) return;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int c[kScanItems];
  int a = 0, b = 0;
  for (int k = 0; k < kScanItems; ++k) {
    const int i = base + threadIdx.x * kScanItems + k;
    c[k] = i < n ? count[i] : 0;
    a += c[k]; b += c[k] > 0;
  }
  int2 inc = make_int2(a, b);
  for (int o = 1; o < 32; o <<= 1) {
    const int x = __shfl_up_sync(0xFFFFFFFFu, inc.x, o), y = __shfl_up_sync(0xFFFFFFFFu, inc.y, o);
    if (lane >= o) { inc.x += x; inc.y += y; }
  }
  if (lane == 31) ws[warp] = inc;
  __syncthreads();
  int2 off = block_sums[blockIdx.x];
  for (int q = 0; q < warp; ++q) { off.x += ws[q].x; off.y += ws[q].y; }
  int ra = off.x + inc.x - a, rb = off.y + inc.y - b;
  for (int k = 0; k < kScanItems; ++k) {
    const int i = base + threadIdx.x * kScanItems + k;
    if (i < n) { start[i] = ra; slot[i] = rb; cursor[i] = 0; }
    ra += c[k]; rb += c[k] > 0;
  }
}

__global__ void k0_scatter_kernel(const float *__restrict__ in, int n, int stride, K0Window w, const K0State *st,
                                  const int *__restrict__ start, int *cursor, int *members) {
  if (st->n_vox == 0) return;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float3 p = load_xyz(in, i, stride);
    if (!in_window(p, w)) continue;
    const int key = voxel_key(p, *st);
    members[start[key] + atomicAdd(&cursor[key], 1)] = i;
  }
}

// one thread per voxel: put its members back in input order, sum sequentially in float
__global__ void k0_centroid_kernel(const float *__restrict__ in, int stride, const K0State *st,
                                   const int *__restrict__ count, const int *__restrict__ start,
                                   const int *__restrict__ slot, int *members, float *out_xyz, int out_cap,
                                   int out_stride) {
  const int n = st->n_vox;
  for (int v = blockIdx.x * blockDim.x + threadIdx.x; v < n; v += gridDim.x * blockDim.x) {
    const int cnt = count[v];
    if (cnt == 0) continue;
    int *m = members + start[v];
    for (int a = 1; a < cnt; ++a) {   // insertion sort: lists are a handful of points long
      const int key = m[a];
      int b = a - 1;
      while (b >= 0 && m[b] > key) { m[b + 1] = m[b]; --b; }
      m[b + 1] = key;
    }
    float cx = 0.0f, cy = 0.0f, cz = 0.0f;
    for (int a = 0; a < cnt; ++a) {
      const float3 p = load_xyz(in, m[a], stride);
      cx = __fadd_rn(cx, p.x); cy = __fadd_rn(cy, p.y); cz = __fadd_rn(cz, p.z);
    }
    const float fc = (float)cnt;
    const int o = slot[v];
    if (o < out_cap) {
      out_xyz[(size_t)o * out_stride] = __fdiv_rn(cx, fc);
      out_xyz[(size_t)o * out_stride + 1] = __fdiv_rn(cy, fc);
      out_xyz[(size_t)o * out_stride + 2] = __fdiv_rn(cz, fc);
    }
  }
}

}  // namespace h3d
#endif  // H3D_PREFILTER_CUH_
