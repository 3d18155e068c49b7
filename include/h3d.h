/*
 * h3d.h -- C-ABI of libh3d.so: the B200-native (sm_100a) iterative 3D Hough
 * line-detection path of DJacquemont/pointcloud_segmentation.
 *
 * The reference has no FFI for this path: its boundary is the C++ header API of
 * libraries/hough-3d-lines ("lib/" below), compiled straight into the ROS node
 * (CMakeLists.txt:164-168).  The drop-in C++ classes shipped under
 * pointcloud_segmentation_b200/include/ (Hough, PointCloud, Vector3d, Sphere,
 * initHoughSpace, orthogonal_LSQ) are thin host code over THIS interface; each
 * entry point cites the reference function it replaces.
 *
 * Conventions: plain pointers and sizes, HOST buffers unless a name ends in
 * `_dev`; every function returns an int status (H3D_OK == 0) and never throws;
 * h3d_last_error() gives the text of the last failure on the calling thread.
 * Points are `double xyz[3*n]` (lib/vector3d.h:15-20 layout, AoS) unless stated.
 * There is no CPU fallback: without a CUDA device every compute call fails with
 * H3D_ERR_CUDA.
 *
 * "wrap.h" = include/pointcloud_segmentation/hough_3d_lines.h of the reference,
 * "node.cpp" = src/pointcloud_segmentation_node.cpp.
 */
#ifndef H3D_H_
#define H3D_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum {
  H3D_OK = 0,
  H3D_ERR_ARG = 1,       /* bad argument / not initialised                   */
  H3D_ERR_CUDA = 2,      /* CUDA runtime failure or no device                */
  H3D_ERR_NOMEM = 3,     /* device or host allocation failed                 */
  H3D_ERR_RANGE = 4,     /* accumulator has >= 2^32 cells, num_x too large … */
  H3D_ERR_DEGENERATE = 5,/* wrap.h:202-212: identical points / dx too large  */
  H3D_ERR_NCCL = 6
};

const char *h3d_last_error(void);
/* kernels launched by the library so far in this process (bench.py: gpu_launches) */
unsigned long long h3d_launch_count(void);
/* library / device facts: sm count, compute capability major*10+minor */
int h3d_device_info(int *sm_count, int *cc, size_t *smem_per_block_optin);

/* ---- direction sphere: lib/hough.cpp:15-20 initHoughSpace,
 *      lib/sphere.cpp:18-330 Sphere::fromIcosahedron ----------------------- */
/* builds (and caches) the direction table for granularity g (0..8) and makes it
 * the current Hough space, like the reference's global `houghSpace`. */
int h3d_init_space(unsigned granularity);
int h3d_space_info(unsigned *granularity, size_t *num_b);
/* Sphere::vertices / Sphere::triangles of granularity g (does not change the
 * current space).  Pass NULL buffers to query the counts. */
int h3d_sphere(int g, double *vertices_xyz, size_t cap_vertices, size_t *n_vertices,
               unsigned *triangles, size_t cap_indices, size_t *n_indices);

/* ---- Hough accumulator: lib/hough.h:21-46, lib/hough.cpp:26-135 ---------- */
typedef struct h3d_hough h3d_hough;

/* Hough::Hough(minP, maxP, dx) (lib/hough.cpp:26-43).  Uses the current space.
 * The accumulator is NOT materialised in HBM: each direction's (x',y') plane
 * lives in shared memory while it is voted and only (max votes, first max cell)
 * per direction is kept; h3d_download_votes() rebuilds the dense array. */
int h3d_create(const double minP[3], const double maxP[3], double dx, h3d_hough **out);
/* direction-sharded variant: this instance owns directions
 * [dir_begin, dir_end) of the current space; flat indices stay global. */
int h3d_create_sharded(const double minP[3], const double maxP[3], double dx,
                       size_t dir_begin, size_t dir_end, h3d_hough **out);
void h3d_destroy(h3d_hough *h);                       /* Hough::~Hough */
int h3d_info(const h3d_hough *h, size_t *num_x, size_t *num_b, double *dx, double *max_x);
/* Hough::add (lib/hough.cpp:50-55) / Hough::subtract (lib/hough.cpp:58-63) */
int h3d_add(h3d_hough *h, const double *xyz, size_t n);
int h3d_subtract(h3d_hough *h, const double *xyz, size_t n);
/* Hough::getLine (lib/hough.cpp:103-135): returns votes, anchor a, direction b
 * and the flat accumulator index x_i*num_x*num_b + y_i*num_b + j of the winner
 * (first maximum in flat order, like the reference's strict `>` scan). */
int h3d_get_line(h3d_hough *h, double a[3], double b[3], unsigned *votes, uint64_t *flat_index);
/* the winner of this instance's shard as a packed key
 * (votes << 32) | (0xFFFFFFFF - flat_index); max over shards == global winner */
int h3d_get_line_key(h3d_hough *h, uint64_t *key);
/* inverse map of lib/hough.cpp:114-132 for a flat index */
int h3d_line_from_index(const h3d_hough *h, uint64_t flat_index, double a[3], double b[3]);
/* Hough::VotingSpace: dense num_x*num_x*num_b array in the reference layout
 * (direction fastest, lib/hough.cpp:88-90).  cap in cells. */
int h3d_download_votes(h3d_hough *h, unsigned *votes, size_t cap);

/* ---- PointCloud: lib/pointcloud.cpp:18-128 -------------------------------- */
int h3d_min_max(const double *xyz, size_t n, double mn[3], double mx[3]);    /* :41-59  */
int h3d_shift_to_origin(double *xyz, size_t n, double shift_inout[3]);       /* :18-26  */
int h3d_mean_value(const double *xyz, size_t n, double mean[3]);             /* :29-38  */
/* pointsCloseToLine (:94-105): indices of kept points, input order */
int h3d_points_close_to_line(const double *xyz, size_t n, const double a[3], const double b[3],
                             double dx, uint32_t *idx_out, size_t *n_out);
/* removePoints (:109-128): ordered set difference by exact ==, compacts x in place */
int h3d_remove_points(double *x_xyz, size_t nx, const double *y_xyz, size_t ny, size_t *n_out);
/* pointsCloseToLine WITHOUT ANY ALLOCATION.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same result as h3d_points_close_to_line (bit-exact: same expression tree, same order), but
 *   - xyz_dev (n x 3 doubles), idx_out_dev (>= n uint32) and n_out_dev (one int) are DEVICE pointers owned by
 *     the caller; a[3], b[3] are host arrays read before the call returns;
 *   - the one-byte-per-point inlier flags live in a file-scope __device__ array of H3D_NOALLOC_MAX_POINTS
 *     bytes (1 MiB, held by every GPU that loads the library until the process exits); n beyond that is
 *     rejected with H3D_ERR_RANGE and a message in h3d_last_error();
 *   - the two kernels are launched on `stream` (a cudaStream_t, NULL = default stream) and the call returns
 *     without synchronising; it calls no cudaMalloc / cudaFree / allocating library routine.  The first call
 *     on each device looks the array's address up (cudaGetSymbolAddress): make it outside stream capture.
 * Every call on a device shares the array (it keeps the last call's contents; nothing in it needs clearing,
 * each call writes flags[0, n) before reading them): calls that OVERLAP on the GPU race -- keep them on one
 * stream or order them with events. */
#define H3D_NOALLOC_MAX_POINTS 1048576
int h3d_points_close_to_line_noalloc(const double *xyz_dev, size_t n, const double a[3], const double b[3],
                                     double dx, uint32_t *idx_out_dev, int *n_out_dev, void *stream);
/* orthogonal_LSQ (wrap.h:121-150): anchor = f64 mean, direction = top eigenvector
 * of the f32 centred scatter, *rc = top eigenvalue */
int h3d_orthogonal_lsq(const double *xyz, size_t n, double a[3], double b[3], double *rc);

/* ---- pre-filter: node.cpp:386-421 (PassThrough x,y,z + VoxelGrid) --------- */
/* xyz_in: n points, `stride` floats apart, x,y,z first.  win = {xmin,xmax,ymin,
 * ymax,zmin,zmax} inclusive.  out_xyz: cap*3 floats, one centroid per voxel in
 * ascending voxel index.  *n_out may exceed cap (then only cap were written). */
int h3d_prefilter(const float *xyz_in, size_t n, size_t stride, const float win[6], float leaf,
                  float *out_xyz, size_t cap, size_t *n_out);

/* ---- the whole iterative loop on the device: wrap.h:172-259,339-347 /
 *      lib/hough3dlines.cpp:200-317 ---------------------------------------- */
typedef struct {
  uint32_t votes;      /* getLine return value                                  */
  uint32_t status;     /* 0 = line accepted; 1 = rc1==0; 2 = < minvotes; 3 = rc2==0 */
  uint64_t argmax;     /* flat accumulator index of the winner                  */
  double a_raw[3], b_raw[3];    /* getLine output (centred frame)               */
  double a1[3], b1[3], rc1;     /* first orthogonal_LSQ                         */
  double a2[3], b2[3], rc2;     /* second orthogonal_LSQ, a2 already + X.shift  */
  uint32_t n1, n2;     /* inlier counts of pass 1 / pass 2                      */
  uint32_t off1, off2; /* offsets of the two inlier lists in the index pool     */
} h3d_iter;

typedef struct {
  int rc;              /* 0 ok; 1 = degenerate cloud / dx too large             */
  uint32_t n_iters;    /* iterations recorded (accepted + the final break)      */
  uint32_t n_lines;    /* accepted lines                                        */
  uint32_t n_input;    /* points after the NaN/Inf strip                        */
  double shift[3];     /* X.shift                                               */
  double d;            /* bbox diagonal                                         */
  double dx;           /* effective dx                                          */
  double max_x;
  uint64_t num_x, num_b;
  uint64_t pool_used;
  uint64_t votes_cast; /* (point,direction) votes evaluated, for votes/s        */
} h3d_result;

/* mode 0 = ROS wrapper rules (wrap.h), 1 = standalone rules
 * (lib/hough3dlines.cpp:171-173,211-213: dx==0 -> d/64, minvotes >= 2).
 * pool receives, per iteration, the ORIGINAL indices of the pass-1 and pass-2
 * inliers.  Uses and may change the current space (h3d_init_space(granularity)). */
int h3d_hough3dlines(const double *xyz, size_t n, int granularity, double opt_dx,
                     int opt_minvotes, int opt_nlines, int mode,
                     h3d_iter *iters, size_t max_iters, uint32_t *pool, size_t pool_cap,
                     h3d_result *res);

/* ---- the same loop with the accumulator sharded by DIRECTION over the ranks of one box
 *      (BASELINE configs[3]) ---------------------------------------
 * A vote for direction j touches only plane j (lib/hough.cpp:90): rank r of `world` owns directions
 * [num_b*r/world, num_b*(r+1)/world), every rank holds all points, and Hough::getLine is the MAX over
 * ranks of the packed key votes << 32 | (0xFFFFFFFF - global flat index) -- one 8-byte ncclAllReduce per
 * iteration, enqueued on the loop's stream between the shard's argmax and the line fit, so the loop stays
 * device-resident.  Inliers, LSQ and removal are computed redundantly and deterministically by every rank:
 * every rank returns the same iteration records as h3d_hough3dlines on one GPU (votes_cast counts the
 * rank's own votes).  One process (or thread) per GPU; NCCL is opened with dlopen("libnccl.so.2") on first
 * use, libh3d.so does not link it. */
typedef struct h3d_comm h3d_comm;
#define H3D_COMM_ID_BYTES 128
/* rank 0 makes the id (ncclGetUniqueId) and hands it to the other ranks by the host's own means */
int h3d_comm_unique_id(unsigned char id[H3D_COMM_ID_BYTES]);
/* collective over all ranks; binds to the calling thread's current CUDA device */
int h3d_comm_create(const unsigned char id[H3D_COMM_ID_BYTES], int rank, int world, h3d_comm **out);
void h3d_comm_destroy(h3d_comm *c);
/* comm == NULL: all directions on this GPU (== h3d_hough3dlines).  Collective: every rank calls it with
 * the same points and options. */
int h3d_hough3dlines_sharded(const double *xyz, size_t n, int granularity, double opt_dx,
                             int opt_minvotes, int opt_nlines, int mode, h3d_comm *comm,
                             h3d_iter *iters, size_t max_iters, uint32_t *pool, size_t pool_cap,
                             h3d_result *res);
/* device times (CUDA events on the loop's stream, microseconds) of the calling thread's last
 * h3d_hough3dlines / h3d_hough3dlines_sharded call: [0] upload + NaN strip + bbox, [1] Hough::add (the
 * first vote), [2] the loop's kernels including the in-stream all-reduces, [3] the whole call with its
 * result copies.  All zero when the call returned early (rc != 0). */
int h3d_hough3dlines_timing(float us[4]);

/* ---- persistent pipeline: pre-filter + loop, buffers kept between frames --- */
typedef struct h3d_pipeline h3d_pipeline;
typedef struct {
  int granularity;
  float window[6];       /* PassThrough limits, node.cpp:394-407                */
  float leaf;            /* VoxelGrid leaf, node.cpp:410-412                    */
  double opt_dx;         /* node.cpp:243: sqrt(3)*leaf                          */
  int opt_minvotes, opt_nlines;
  size_t max_raw_points; /* capacity of one raw frame                           */
  size_t max_iters;      /* capacity of the per-frame iteration records         */
  int prefilter;         /* 1: raw f32 frame -> K0 -> loop; 0: input is already
                            the filtered f32 cloud (wrap.h:175-189 widening)    */
} h3d_pipeline_config;

int h3d_pipeline_create(const h3d_pipeline_config *cfg, h3d_pipeline **out);
void h3d_pipeline_destroy(h3d_pipeline *p);
/* one frame, host in / host out (H2D + kernels + D2H).  raw: n points of
 * `stride` floats (x,y,z first; sensor_msgs/PointCloud2 with point_step =
 * 4*stride).  Inlier indices in pool refer to the FILTERED cloud. */
int h3d_pipeline_run(h3d_pipeline *p, const float *raw, size_t n, size_t stride,
                     h3d_iter *iters, uint32_t *pool, size_t pool_cap, h3d_result *res);
/* same with the raw frame already resident in device memory */
int h3d_pipeline_run_dev(h3d_pipeline *p, const float *raw_dev, size_t n, size_t stride,
                         h3d_iter *iters, uint32_t *pool, size_t pool_cap, h3d_result *res);
/* sensor_msgs/PointCloud2 adapter (node.cpp:286-289 pcl_conversions::toPCL +
 * cloudFiltering, node.cpp:415 fromPCLPointCloud2): the message's byte buffer goes to the device
 * as it is and the frame runs from there -- no host pass over the points.  x, y, z must be
 * FLOAT32 fields (datatype 7) of a little-endian message.  Point i = row i / width, column
 * i % width, at data[row * row_step + col * point_step + off_*].  When x,y,z are adjacent and the
 * rows carry no padding the kernels read the message bytes in place; otherwise one gather kernel
 * packs them first. */
typedef struct {
  uint32_t width, height;       /* sensor_msgs/PointCloud2.width / .height      */
  uint32_t point_step, row_step;
  uint32_t off_x, off_y, off_z; /* fields[].offset of "x", "y", "z"             */
  uint32_t is_bigendian;        /* must be 0                                    */
} h3d_cloud2_layout;
int h3d_pipeline_run_cloud2(h3d_pipeline *p, const uint8_t *data, size_t data_bytes,
                            const h3d_cloud2_layout *layout, h3d_iter *iters, uint32_t *pool,
                            size_t pool_cap, h3d_result *res);
/* filtered cloud of the last frame (f32 xyz), for inspection */
int h3d_pipeline_filtered(h3d_pipeline *p, float *out_xyz, size_t cap, size_t *n_out);
/* per-phase device times of the last frame, microseconds:
 * [0] pre-filter + widening/bbox, [1] the first vote kernel alone (Hough::add),
 * [2] the loop's kernels (one persistent launch on the fused path), [3] whole
 * frame incl. the result copies and the host<->device copies of
 * h3d_pipeline_run.  All zero when the frame ended early (rc != 0). */
int h3d_pipeline_timing(h3d_pipeline *p, float us[4]);

/* ---- segment post-tests: wrap.h:261-337 with find_t (wrap.h:41-68), find_proj
 *      (wrap.h:78-85), segPCA (wrap.h:94-110) -- SURVEY 8(f) rank 1 ------------
 * The reference runs them on the host on a copy of Y.points, line by line; here
 * one launch tests every accepted line of the frame from the inlier index lists
 * that are still on the device, and only these records come back. */
#define H3D_MAX_RADIUS_SIZES 8
typedef struct {
  double diag_voxel;       /* node.cpp:242: sqrt(3) * leaf_size                  */
  double min_pca_coeff;    /* node.cpp:200                                       */
  double rad_2_leaf_ratio; /* node.cpp:210                                       */
  uint32_t n_radius;       /* entries of radius_sizes in use (1..8), node.cpp:230-239 */
  uint32_t pad_;
  double radius_sizes[H3D_MAX_RADIUS_SIZES];
} h3d_segment_config;
typedef struct {
  uint32_t iter;           /* iteration record (h3d_iter) the line came from     */
  uint32_t status;         /* 0 = pushed to computed_lines (wrap.h:324-335);
                              1 = failed the radius / gap test (wrap.h:312);
                              2 = failed the PCA / density test (wrap.h:322);
                              3 = find_t had no solution, b.x == 0: the reference
                                  returns 1 from hough3dlines here (wrap.h:276-279) */
  uint32_t n_points;       /* points_size                                        */
  int32_t min_nb_points;   /* min_nb_points_seg (wrap.h:320); 0 unless the first test passed */
  double a[3], b[3];
  double t_min, t_max;     /* t_values.front() / back()                          */
  double radius_raw;       /* max(p_radius.front(), p_radius.back()), wrap.h:295 */
  double radius;           /* closest_radius                                     */
  double max_dist;         /* largest neighbour gap of p_norm, wrap.h:286-292    */
  double pca_coeff;        /* 0 unless the first test passed                     */
  double pca_eigenvalues[3];
  double length;           /* |p(t_max) - p(t_min)|, wrap.h:317-319              */
  /* drone -> world transform + height cut-off (node.cpp:429-470), filled for status 0
   * when a pose is set (h3d_pipeline_set_pose), else zero: */
  double a_world[3], b_world[3];   /* R*a + position, R*b (node.cpp:437-438)     */
  uint32_t world_status;   /* 0 = no pose / not a kept segment; 1 = above the floor
                              (kept by heighSegmentCutoff, node.cpp:461); 2 = trimmed */
  uint32_t pad_;
} h3d_segment;
/* pose of the sensor for the following frames: orientation quaternion
 * (w, x, y, z) as Eigen::Quaterniond::toRotationMatrix reads it, position, and
 * floor_trim_height (config_pc_seg: 0.3).  NULL: no transform. */
typedef struct {
  double q[4];
  double position[3];
  double floor_trim_height;
} h3d_pose;
int h3d_pipeline_set_pose(h3d_pipeline *p, const h3d_pose *pose);
/* cfg != NULL: every following frame of the pipeline also runs the post-tests
 * (one extra launch, no extra synchronisation); NULL switches them off. */
int h3d_pipeline_enable_segments(h3d_pipeline *p, const h3d_segment_config *cfg);
/* records of the last frame in iteration order, one per line the reference would
 * have tested.  *rc_ref = what hough3dlines returns: 1 when the LAST record has
 * status 3 (the reference stops there; later lines are not reported), else the
 * frame's rc.  *n_segs may exceed cap (then only cap were written). */
int h3d_pipeline_segments(h3d_pipeline *p, h3d_segment *segs, size_t cap, size_t *n_segs, int *rc_ref);
/* the same tests on one line from host buffers: y_xyz = the line's final inliers
 * Y.points (centred frame, n points), shift = X.shift, a = anchor (+ shift), b =
 * direction.  Returns H3D_OK also when out->status != 0. */
int h3d_segment_test(const double *y_xyz, size_t n, const double shift[3], const double a[3],
                     const double b[3], const h3d_segment_config *cfg, h3d_segment *out);
/* same with the drone -> world step (pose may be NULL) */
int h3d_segment_test_pose(const double *y_xyz, size_t n, const double shift[3], const double a[3],
                          const double b[3], const h3d_segment_config *cfg, const h3d_pose *pose,
                          h3d_segment *out);

/* frame-parallel replay on ONE GPU (BASELINE configs[2]): frames are independent (one Hough per
 * frame, wrap.h:220/347), and the loop of a frame is latency-bound, so K pipelines driven by K
 * host threads can each take sm_count/K SMs for their persistent loop kernel.  n_ctas = 0: all
 * SMs (lowest latency for a single frame). */
int h3d_pipeline_set_sm_share(h3d_pipeline *p, int n_ctas);

/* Batched replay of independent frames (BASELINE configs[2]; node.cpp:286-294 hands one frame at a
 * time to hough3dlines): n_pipes host threads, one per pipeline, pull frames from a shared counter.
 * Frame i = raw[i] (n_points[i] points, `stride` floats apart; DEVICE pointers when
 * frames_on_device != 0) -> results[i], iters + i*max_iters, pool + i*pool_cap (iters / pool may be
 * NULL).  Set CUDA_DEVICE_MAX_CONNECTIONS >= n_pipes + 1 before the first CUDA call, or the
 * pipelines' streams share hardware queues and serialise. */
int h3d_replay_run(h3d_pipeline *const *pipes, size_t n_pipes, const float *const *raw,
                   const size_t *n_points, size_t stride, size_t n_frames, int frames_on_device,
                   h3d_iter *iters, size_t max_iters, uint32_t *pool, size_t pool_cap,
                   h3d_result *results);

/* how the host thread of a frame waits for the device: 0 (default) = cudaStreamSynchronize, which
 * spins -- lowest latency; 1 = a blocking event, the thread sleeps -- use it when more pipelines
 * are in flight than the host has cores (several GPUs x frames in flight); 2 = poll an event and
 * yield the core between polls.  Process-wide; returns the previous setting. */
int h3d_set_wait_mode(int mode);
/* Host threads of h3d_replay_run*: 0 (default) = one per pipeline; 0 < n < pipelines = n threads, each driving
 * several pipelines in split phase (submit a frame, poll, complete) -- for boxes with fewer host cores than
 * ranks x pipelines (node.cpp has one callback thread; this is replay plumbing).  Returns the previous value. */
int h3d_set_replay_threads(int n);

/* same, also collecting the segment records of every frame (the pipelines must have
 * h3d_pipeline_enable_segments on): segments + i*max_segments, n_segments[i] (may exceed
 * max_segments), rc_ref[i] = what the reference's hough3dlines returns for frame i (may be NULL). */
int h3d_replay_run_segments(h3d_pipeline *const *pipes, size_t n_pipes, const float *const *raw,
                            const size_t *n_points, size_t stride, size_t n_frames, int frames_on_device,
                            h3d_iter *iters, size_t max_iters, uint32_t *pool, size_t pool_cap,
                            h3d_result *results, h3d_segment *segments, size_t max_segments,
                            uint32_t *n_segments, int *rc_ref);

/* ---- world model across frames: segFiltering, checkSimilarity, checkConnections
 *      (node.cpp:479-667), saveSegmentsToFile / saveIntersectionsToFile
 *      (node.cpp:850-897) -- SURVEY 8(f) rank 4 --------------------------------
 * State kept ACROSS frames on tens of segments, O(s^2) scalar work: host C++ in
 * the reference and here (no kernel); it consumes the 200-byte records the device
 * post-tests return, so a replay reproduces segments.csv / intersections.csv
 * without ROS.  Vector arithmetic follows Eigen's fixed-size evaluation order
 * (3-term sums reduce as x0 + (x1 + x2)), no FMA. */
typedef struct h3d_world h3d_world;
typedef struct {
  double diag_voxel;       /* node.cpp:242 */
  double min_weight;       /* node.cpp:205 (config_pc_seg: /min_weight)          */
} h3d_world_config;
typedef struct {           /* `segment` of wrap.h:21-29 without the point copies */
  double a[3], b[3];
  double t_min, t_max;
  double radius;
  int32_t points_size;
  int32_t pad_;
  double pca_coeff;
  double pca_eigenvalues[3];
} h3d_world_segment;
int h3d_world_create(const h3d_world_config *cfg, h3d_world **out);
void h3d_world_destroy(h3d_world *w);
/* segFiltering(drone_segments) (node.cpp:479-542): drone = the frame's segments
 * in the world frame, in order.  Quirks kept: similarity is tested against the
 * segments as they were BEFORE the frame; a later drone segment similar to the
 * same world segment overwrites the earlier fusion; a new segment is recorded
 * under the index `size after push + i` (node.cpp:508), so its connections are
 * looked for only when that index happens to exist; entries of the intersection
 * matrix are never reset.  Defined here where the reference reads indeterminate
 * values: `segment target_seg;` (node.cpp:494) is uninitialised when
 * checkSimilarity adds to its points_size / pca_coeff / pca_eigenvalues
 * (node.cpp:651-654) -- this build starts them from the WORLD segment (the
 * "[in,out]" the doc comment announces); and points_size 0 + 0 (integer division
 * by zero, node.cpp:617) gives weight 0. */
int h3d_world_update(h3d_world *w, const h3d_world_segment *drone, size_t n);
/* same from the device records of one frame (h3d_pipeline_segments): takes those
 * with status 0 and world_status 1 (drone2WorldSeg + heighSegmentCutoff kept
 * them), a/b = a_world/b_world.  *n_used (may be NULL) = how many were taken. */
int h3d_world_update_frame(h3d_world *w, const h3d_segment *segs, size_t n, size_t *n_used);
int h3d_world_segments(const h3d_world *w, h3d_world_segment *out, size_t cap, size_t *n);
/* intersection_matrix (node.cpp:138): t_pairs[(i*n + j)*2 + {0,1}] for all i, j < n
 * ((-1, -1) = none; only j < i is ever set); cap_pairs = pairs t_pairs can hold. */
int h3d_world_intersections(const h3d_world *w, double *t_pairs, size_t cap_pairs, size_t *n);
/* checkSimilarity (node.cpp:594-667): *similar = 1 and *target = the fused segment,
 * else *target = *drone.  target_init = the value `target_seg` holds on entry. */
int h3d_check_similarity(const h3d_world_segment *drone, const h3d_world_segment *world,
                         const h3d_world_config *cfg, const h3d_world_segment *target_init,
                         h3d_world_segment *target, int *similar);
/* checkConnections (node.cpp:556-590): sol = colPivHouseholderQr solution
 * (t along s1 from t_min, t along s2 from t_min, distance); *connected as the
 * reference returns it.  sol is written also when not connected (zero if parallel). */
int h3d_check_connections(const h3d_world_segment *s1, const h3d_world_segment *s2,
                          const h3d_world_config *cfg, double sol[3], int *connected);
/* saveSegmentsToFile / saveIntersectionsToFile (node.cpp:850-897): same columns,
 * default ostream formatting (%g, 6 significant digits).  Either path may be NULL. */
int h3d_world_save_csv(const h3d_world *w, const char *segments_path, const char *intersections_path);
/* The node's verbosity levels (node.cpp:23 enum verbose {NONE, INFO, WARN}, /verbose_level): above NONE the
 * callback time of every frame is printed, above INFO also the warnings, the intersection matrix and the segment
 * list (node.cpp:296-346); ROS_INFO / ROS_WARN become "[ INFO] ..." / "[ WARN] ..." lines on stderr.
 * Returns the previous level. */
int h3d_set_verbosity(int level);
/* The tail of the node's callback for one frame (node.cpp:308-346): logs by verbosity and appends the frame's
 * callback_info {wall_time, callback_time, world_segments.size(), nblines} (microseconds).  h3d_replay_run_world
 * calls it for every frame. */
int h3d_world_note_frame(h3d_world *w, int wall_time_us, int callback_time_us, int nblines, int hough_rc);
/* saveProcessingTimeToFile (node.cpp:905-919): "wall_time,processing_time,seg_vec_size,nblines" + one row per frame */
int h3d_world_save_processing_time(const h3d_world *w, const char *path);

/* ROS-free replay of node.cpp:286-306 for n_frames frames: hough3dlines with the post-tests,
 * drone2WorldSeg + heighSegmentCutoff with poses[i] (the pipelines must have
 * h3d_pipeline_enable_segments on), then segFiltering into `world` in frame order.  Frames run in
 * flight on the pipelines; the world update is sequential and runs on the collected records.
 * segments / n_segments / rc_ref as in h3d_replay_run_segments; max_segments must hold every frame. */
int h3d_replay_run_world(h3d_pipeline *const *pipes, size_t n_pipes, const float *const *raw,
                         const size_t *n_points, size_t stride, size_t n_frames, int frames_on_device,
                         const h3d_pose *poses, h3d_result *results, h3d_segment *segments,
                         size_t max_segments, uint32_t *n_segments, int *rc_ref, h3d_world *world);

/* measurement hook: phase stamps of the fused loop kernel.  out == NULL: enable for `iters`
 * iterations (0 disables).  out != NULL: copy the stamps of the last frame,
 * [iteration][cta][n_slots] globaltimer nanoseconds (slot 8: re-voted planes + 1). */
int h3d_pipeline_profile(h3d_pipeline *p, int iters, unsigned long long *out, size_t cap, int *n_cta,
                         int *n_slots);

/* the frame loop runs as one persistent cooperative kernel when the cloud fits the register-
 * resident form (<= 6144 points); on = 0 forces the general multi-kernel loop.  Returns the
 * previous setting.  Results are identical either way. */
int h3d_set_fused(int on);

/* clouds of 32 768 points and more run K3 and removePoints with chunked multi-CTA kernels (every
 * 1024-point chunk is one CTA's work item) instead of the one-CTA kernels of the frame loop; on = 0
 * forces the one-CTA kernels.  Returns the previous setting.  Results are identical either way. */
int h3d_set_big_kernels(int on);

/* a full vote (Hough::add) over no more directions than the GPU has SMs and 65 536 points or more
 * (granularity 0-2 on large clouds, or a narrow direction shard) slices the points of a direction over
 * CTAs and merges the partial planes in global memory; on = 0 forces one work item per direction.
 * Returns the previous setting.  Results are identical either way. */
int h3d_set_sliced_vote(int on);

/* the streaming vote (clouds that do not fit the register-resident form: configs[3], the direction shards, the
 * general loop) votes up to 8 directions per tile of 2048 points held in registers.  mode 0 (default): the tile is
 * fetched with coalesced LDGs; 1: a producer warp stages the tiles in shared memory with TMA bulk copies
 * (cp.async.bulk + full/empty mbarriers) -- measured 8 % slower on B200, the kernel is bound by the shared-memory
 * pipe; 2: the round-1 kernel (2 directions per work item, four LDGs per point).  Returns the previous mode.
 * Results are identical in every mode. */
int h3d_set_stream_mode(int mode);

/* frames of a pipeline with a pre-filter and a bounded line count run on the frame fast path: K0 + widening +
 * bounding box + shift + Hough constructor in ONE thread-block-cluster launch, first vote, loop, post-tests and
 * one export kernel that writes the results into mapped pinned host memory -- the host waits once per frame.
 * on = 0 forces the general path (eleven K0/prep launches, Hough constructor on the host, three waits).
 * Returns the previous setting.  Results are identical either way. */
int h3d_set_fast_frames(int on);

/* first vote of the frame fast path: mode 1 the shared-memory increment returns the old count and every thread
 * tracks the maximum of its votes (no plane scan); mode 0 increments with the merging RED form and scans every
 * plane; mode 2 (default) picks on the device: tracking up to 3072 points, scanning above.  Returns the previous
 * mode.  Results are identical in every mode. */
int h3d_set_vote_track(int mode);
/* measurement hook: globaltimer stamps (ns) of the K0 cluster kernel of the pipeline's last fast-path frame:
 * start, after each of its seven cluster barriers (count, reduce, scan, scatter, order, centroids, bbox), end */
int h3d_pipeline_k0_stamps(h3d_pipeline *p, unsigned long long t_ns[10]);

/* debug aid: on != 0 makes every frame start from per-direction records of 0xFFFFFFFF votes, so a
 * record the first vote fails to write wins every argmax instead of silently keeping the previous
 * frame's value (the parity tests of the frames-in-flight replay run with it).  Returns the previous
 * setting.  Results are identical either way when the kernels are correct. */
int h3d_set_debug_poison(int on);

/* ---- measurement hooks (bench.py): raw kernel timings with CUDA events ----- */
/* runs the vote kernel `reps` times over all directions of h and returns the
 * mean device time per launch in microseconds.  ablate: 0 = full kernel,
 * 1 = projection math only (no shared atomics), 2 = shared atomics only;
 * bits 8..15 of `ablate` optionally give the directions per work item (1,2,4; 0 = automatic: register-resident points when they fit). */
int h3d_bench_vote(h3d_hough *h, int reps, int ablate, float *us_per_launch);
/* one orthogonal_LSQ call (wrap.h:121-150) on a host cloud: device time of its phases in microseconds
 * (us[0] to the barrier after the column means and the sequential anchor, us[1] scatter sums, us[2] eigen-solve;
 * us[3] / us[4]: when warp 0 had its float mean / warp 3 its sequential double mean, from the kernel's start),
 * globaltimer stamps */
int h3d_bench_lsq(const double *xyz, size_t n, float us[5]);

#ifdef __cplusplus
}
#endif
#endif /* H3D_H_ */
