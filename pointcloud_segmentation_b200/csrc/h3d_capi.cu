// h3d_capi.cu -- the extern "C" boundary of libh3d.so (include/h3d.h): contexts,
// device memory, kernel launches.  No CPU fallback: every compute entry point
// needs a CUDA device and fails with H3D_ERR_CUDA otherwise.
#include "../../include/h3d.h"

#include <dlfcn.h>
#include <nccl.h>   // types and enums only: NCCL is bound with dlopen, no symbol of it is linked

#include <atomic>
#include <cfloat>
#include <chrono>
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

// The library's C-ABI is one translation unit (the kernel headers hold out-of-line __device__ functions and
// file-scope state is shared), laid out in five parts:
#include "h3d_capi_core.inc"       // errors, device facts, spaces, h3d_hough, vote launchers, Hough entry points
#include "h3d_capi_loop.inc"       // hough3dlines on the device: fused / multi-kernel loop, NCCL, direction shards
#include "h3d_capi_pipeline.inc"   // pre-filter + pipeline + frame fast path + segments + PointCloud2
#include "h3d_capi_replay.inc"     // frames in flight: h3d_replay_run*, wait modes, split-phase workers
#include "h3d_capi_hostapi.inc"    // PointCloud entry points on host arrays, run-time switches
