/*
 * h3d_oracle.h -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C CPU restatement of the iterative 3D Hough hot path of
 * DJacquemont/pointcloud_segmentation (library: libraries/hough-3d-lines).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this.  The product (libh3d.so) never does.
 *
 * Parity status (see DESIGN.md "Oracle"):
 *   - sphere, Hough ctor/vote/argmax, inliers, meanValue, removePoints,
 *     bbox/shift:  PINNED -- checked bit-for-bit against the reference
 *     library itself compiled in place (oracle/_ref, oracle/ref_shim.cpp)
 *     and against golden vectors generated from it (tests/golden/).
 *   - orthogonal_LSQ direction/eigenvalue (Eigen3, un-vendored) and the
 *     PassThrough/VoxelGrid pre-filter (PCL 1.10, un-vendored):
 *     PARITY UNPINNED -- restated from the published algorithms; the
 *     reference holds no test or golden vector at that boundary.
 *
 * "lib/" below = /root/reference/libraries/hough-3d-lines/,
 * "wrap.h" = /root/reference/include/pointcloud_segmentation/hough_3d_lines.h,
 * "node.cpp" = /root/reference/src/pointcloud_segmentation_node.cpp.
 */
#ifndef H3D_ORACLE_H_
#define H3D_ORACLE_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- direction sphere (lib/sphere.cpp:18-330) ------------------------- */
/* number of directions for granularity g (0..6), 0 on error */
size_t h3do_sphere_count(int g);
/* fills n*3 doubles (x,y,z per direction); returns n, 0 on error */
size_t h3do_sphere_fill(int g, double *xyz, size_t cap);
/* surviving triangle index list after makeUnique; returns count of indices */
size_t h3do_sphere_triangles(int g, unsigned *tri, size_t cap);

/* ---- Hough accumulator (lib/hough.cpp:15-135) -------------------------- */
typedef struct h3do_hough h3do_hough;
/* directions are borrowed from the caller (n_dir*3 doubles, kept alive) */
h3do_hough *h3do_hough_create(const double *dirs, size_t n_dir,
                              const double minP[3], const double maxP[3],
                              double dx);
void   h3do_hough_destroy(h3do_hough *h);
size_t h3do_hough_num_x(const h3do_hough *h);
size_t h3do_hough_num_b(const h3do_hough *h);
double h3do_hough_dx(const h3do_hough *h);
double h3do_hough_max_x(const h3do_hough *h);
size_t h3do_hough_cells(const h3do_hough *h);
const unsigned *h3do_hough_votes(const h3do_hough *h);
void   h3do_hough_add(h3do_hough *h, const double *xyz, size_t n);
void   h3do_hough_subtract(h3do_hough *h, const double *xyz, size_t n);
/* returns votes; writes anchor a, direction b and the flat argmax index */
unsigned h3do_hough_get_line(const h3do_hough *h, double a[3], double b[3],
                             uint64_t *flat_index);
/* multi-threaded variants used only by bench.py --impl reference to give the
 * CPU arm all host threads (directions are split; results are identical) */
void   h3do_hough_add_mt(h3do_hough *h, const double *xyz, size_t n, int threads);
void   h3do_hough_subtract_mt(h3do_hough *h, const double *xyz, size_t n, int threads);
unsigned h3do_hough_get_line_mt(const h3do_hough *h, double a[3], double b[3],
                                uint64_t *flat_index, int threads);

/* ---- PointCloud (lib/pointcloud.cpp:18-128) ---------------------------- */
void   h3do_min_max(const double *xyz, size_t n, double mn[3], double mx[3]);
/* shifts in place, ACCUMULATES into shift[3] (shift = shift + newshift) */
void   h3do_shift_to_origin(double *xyz, size_t n, double shift[3]);
void   h3do_mean_value(const double *xyz, size_t n, double mean[3]);
/* writes indices of kept points (input order); returns count */
size_t h3do_points_close_to_line(const double *xyz, size_t n,
                                 const double a[3], const double b[3],
                                 double dx, uint32_t *idx_out);
/* ordered set difference X \ Y by exact ==; compacts X in place, returns new n */
size_t h3do_remove_points(double *x_xyz, size_t nx, const double *y_xyz, size_t ny);

/* ---- orthogonal_LSQ (wrap.h:121-150) -- Eigen restated, parity unpinned - */
double h3do_orthogonal_lsq(const double *xyz, size_t n, double a[3], double b[3]);
/* the 3x3 float symmetric eigen-solve alone (lower triangle used);
 * evals ascending, evecs column-major 3x3; returns 0 on convergence */
int    h3do_eig3f(const float m[9], float evals[3], float evecs[9]);
int    h3do_eig3f_fixed(const float m[9], float evals[3]);   /* SelfAdjointEigenSolver<Matrix3f>, values only */

/* segment post-tests of one accepted line: wrap.h:261-337 (+ find_t :41-68, find_proj :78-85,
 * segPCA :94-110).  Layout identical to h3d_segment_config / h3d_segment of include/h3d.h. */
typedef struct {
  double diag_voxel;          /* node.cpp:242: sqrt(3) * leaf                        */
  double min_pca_coeff;
  double rad_2_leaf_ratio;
  uint32_t n_radius;          /* 1..8 entries used                                   */
  uint32_t pad_;
  double radius_sizes[8];
} h3do_segment_config;
typedef struct {
  uint32_t iter;              /* iteration record the line came from                 */
  uint32_t status;            /* 0 pushed to computed_lines; 1 failed wrap.h:312; 2 failed wrap.h:322;
                                 3 find_t had no solution (reference returns 1)      */
  uint32_t n_points;
  int32_t min_nb_points;
  double a[3], b[3];
  double t_min, t_max;
  double radius_raw;          /* max(p_radius.front(), p_radius.back())              */
  double radius;              /* closest_radius                                      */
  double max_dist;
  double pca_coeff;
  double pca_eigenvalues[3];
  double length;
  double a_world[3], b_world[3];   /* node.cpp:437-438 */
  uint32_t world_status;           /* 0 none, 1 kept by heighSegmentCutoff (node.cpp:461), 2 trimmed */
  uint32_t pad_;
} h3do_segment;
typedef struct {
  double q[4];                /* w, x, y, z */
  double position[3];
  double floor_trim_height;
} h3do_pose;
/* drone2WorldSeg + heighSegmentCutoff (node.cpp:429-470) on a kept segment (status 0) */
void h3do_segment_to_world(const h3do_pose *pose, h3do_segment *s);
int h3do_segment_test(const double *y_xyz, size_t n, const double shift[3], const double a[3],
                      const double b[3], const h3do_segment_config *cfg, h3do_segment *out);

/* ---- pre-filter (node.cpp:386-421; PCL PassThrough + VoxelGrid) --------- */
/* xyz_in: n points with `stride` floats per point (x,y,z first).
 * win = {xmin,xmax,ymin,ymax,zmin,zmax} inclusive.  out_xyz: cap*3 floats.
 * returns number of output points, or (size_t)-1 if the voxel index
 * would overflow int32 (PCL refuses that case). */
size_t h3do_prefilter(const float *xyz_in, size_t n, size_t stride,
                      const float win[6], float leaf,
                      float *out_xyz, size_t cap);

/* ---- iterative loop (wrap.h:172-259,339-347 / lib/hough3dlines.cpp:200-317) */
typedef struct {
    uint32_t votes;        /* getLine return value                       */
    uint32_t status;       /* 0 = line accepted, 1 = rc1==0, 2 = < minvotes,
                              3 = rc2==0 (the three `break`s)             */
    uint64_t argmax;       /* flat accumulator index of the winner        */
    double a_raw[3], b_raw[3];       /* getLine output (centred frame)    */
    double a1[3], b1[3], rc1;        /* first LSQ                         */
    double a2[3], b2[3], rc2;        /* second LSQ, a2 already + X.shift  */
    uint32_t n1, n2;       /* inlier counts of pass 1 / pass 2            */
    uint32_t off1, off2;   /* offsets into the index pool (ORIGINAL point
                              numbering of the input cloud)               */
} h3do_iter;

typedef struct {
    int      rc;           /* 0 ok; 1 = degenerate/dx too large (wrap.h:202,209) */
    uint32_t n_iters;      /* iterations recorded (accepted + final break) */
    uint32_t n_lines;      /* accepted lines                               */
    uint32_t n_input;      /* points after NaN/Inf strip                   */
    double   shift[3];     /* X.shift                                      */
    double   d;            /* bbox diagonal                                */
    double   dx;           /* effective dx                                 */
    double   max_x;
    uint64_t num_x, num_b;
    uint64_t pool_used;
} h3do_result;

/* mode 0 = ROS wrapper rules (wrap.h), 1 = standalone rules
 * (lib/hough3dlines.cpp: dx==0 -> d/64, minvotes clamped to >= 2).
 * xyz is not modified.  iters: capacity max_iters; pool: capacity pool_cap. */
int h3do_hough3dlines(const double *xyz, size_t n, int granularity,
                      double opt_dx, int opt_minvotes, int opt_nlines, int mode,
                      h3do_iter *iters, size_t max_iters,
                      uint32_t *pool, size_t pool_cap,
                      h3do_result *res);

#ifdef __cplusplus
}
#endif
#endif
