// h3d_world.cpp -- the world model kept across frames:
// segFiltering / checkSimilarity / checkConnections (node.cpp:479-667) and the CSV
// dumps (node.cpp:850-897).  Host code in the reference and here: tens of segments,
// O(s^2) scalar work, state that outlives a frame.  Input = the 200-byte records of
// the device post-tests (h3d_segment), so nothing else has to leave the GPU.
//
// Arithmetic: Eigen fixed-size Vector3d expressions evaluate coefficient by
// coefficient in the written order; 3-term reductions are x0 + (x1 + x2); no FMA
// (-ffp-contract=off).  colPivHouseholderQr().solve follows Eigen 3.3's algorithm
// (norm table with the LAPACK down-date, Householder reflectors, rank from the
// pivot threshold, back substitution); its tiny inner sums run in index order
// (Eigen's own order depends on the packet alignment of the stack object: unpinned).
#include <cfloat>
#include <cmath>
#include <cstdio>
#include <atomic>
#include <cstring>
#include <new>
#include <vector>

#include "../../include/h3d.h"
#include "h3d_internal.h"

namespace {

struct V3 {
  double x, y, z;
};
inline V3 ld(const double *p) { return {p[0], p[1], p[2]}; }
inline void st(double *p, const V3 &v) { p[0] = v.x, p[1] = v.y, p[2] = v.z; }
inline V3 operator+(const V3 &u, const V3 &v) { return {u.x + v.x, u.y + v.y, u.z + v.z}; }
inline V3 operator-(const V3 &u, const V3 &v) { return {u.x - v.x, u.y - v.y, u.z - v.z}; }
inline V3 operator*(double s, const V3 &v) { return {s * v.x, s * v.y, s * v.z}; }
inline double dot(const V3 &u, const V3 &v) { return u.x * v.x + (u.y * v.y + u.z * v.z); }
inline double norm(const V3 &u) { return std::sqrt(dot(u, u)); }
inline V3 at(const h3d_world_segment &s, double t) { return t * ld(s.b) + ld(s.a); }   // t*b + a

// find_proj (wrap.h:78-85): p_A + (p_B - p_A) * ((p - p_A).(p_B - p_A)) / ((p_B - p_A).(p_B - p_A))
V3 find_proj(const V3 &a, const V3 &b, const V3 &p) {
  const V3 d = (a + b) - a;
  const double num = dot(p - a, d), den = dot(d, d);
  return {a.x + (d.x * num) / den, a.y + (d.y * num) / den, a.z + (d.z * num) / den};
}

inline double std_min(double a, double b) { return (b < a) ? b : a; }   // std::min / std::max NaN behaviour
inline double std_max(double a, double b) { return (a < b) ? b : a; }

// checkSimilarity, node.cpp:594-667.  `target` holds target_init on entry.
bool check_similarity(const h3d_world_segment &drone, const h3d_world_segment &world, const h3d_world_config &cfg,
                      h3d_world_segment &target) {
  bool similar = false;
  const V3 world_p1 = at(world, world.t_min), world_p2 = at(world, world.t_max);
  const V3 test_p1 = at(drone, drone.t_min), test_p2 = at(drone, drone.t_max);
  const V3 proj_p1 = find_proj(ld(world.a), ld(world.b), test_p1);
  const V3 proj_p2 = find_proj(ld(world.a), ld(world.b), test_p2);
  const double epsilon = (drone.radius + world.radius) + 2 * (2 * cfg.diag_voxel);
  if (norm(proj_p1 - test_p1) < epsilon && norm(proj_p2 - test_p2) < epsilon && drone.radius == world.radius) {
    // int / int (node.cpp:617); 0 / 0 defined as 0 here, the sum wraps like two's complement
    const int32_t sum = (int32_t)((uint32_t)world.points_size + (uint32_t)drone.points_size);
    double weight = 0.0;
    if (sum != 0 && !(drone.points_size == INT32_MIN && sum == -1)) weight = (double)(drone.points_size / sum);
    weight = std_max(cfg.min_weight, weight);
    const double coeff = (drone.pca_coeff * weight) / (world.pca_coeff * (1 - weight) + drone.pca_coeff * weight);
    const V3 new_a = proj_p1 + coeff * (test_p1 - proj_p1);
    const V3 new_b = (proj_p2 - proj_p1) + coeff * ((test_p2 - proj_p2) - (test_p1 - proj_p1));
    const V3 tp1 = find_proj(new_a, new_b, test_p1), tp2 = find_proj(new_a, new_b, test_p2);
    const V3 wp1 = find_proj(new_a, new_b, world_p1), wp2 = find_proj(new_a, new_b, world_p2);
    const double t[4] = {(tp1.x - new_a.x) / new_b.x, (tp2.x - new_a.x) / new_b.x, (wp1.x - new_a.x) / new_b.x,
                         (wp2.x - new_a.x) / new_b.x};
    if (!((std_min(t[0], t[1]) > std_max(t[2], t[3])) || (std_max(t[0], t[1]) < std_min(t[2], t[3])))) {
      double hi = t[0], lo = t[0];                        // std::max_element / std::min_element
      for (int k = 1; k < 4; k++) {
        if (hi < t[k]) hi = t[k];
        if (t[k] < lo) lo = t[k];
      }
      st(target.a, new_a);
      st(target.b, new_b);
      target.t_max = hi;
      target.t_min = lo;
      target.radius = drone.radius;
      target.points_size = (int32_t)((uint32_t)target.points_size + (uint32_t)drone.points_size);
      target.pca_coeff = target.pca_coeff * (1 - weight) + drone.pca_coeff * weight;
      for (int k = 0; k < 3; k++)
        target.pca_eigenvalues[k] = target.pca_eigenvalues[k] * (1 - weight) + drone.pca_eigenvalues[k] * weight;
      similar = true;
    }
  }
  if (!similar) target = drone;
  return similar;
}

// Eigen::ColPivHouseholderQR<Matrix3d>(A).solve(rhs); A column-major A[c][r].
void colpiv_qr_solve3(double A[3][3], const double rhs[3], double x[3]) {
  const int n = 3;
  double hcoef[3], norm_upd[3], norm_dir[3];
  int trans[3];
  for (int k = 0; k < n; k++) {
    double s = 0.0;
    for (int r = 0; r < n; r++) s += A[k][r] * A[k][r];
    norm_dir[k] = norm_upd[k] = std::sqrt(s);
  }
  double maxn = norm_upd[0];
  for (int k = 1; k < n; k++)
    if (norm_upd[k] > maxn) maxn = norm_upd[k];
  const double th = maxn * DBL_EPSILON;
  const double threshold_helper = (th * th) / (double)n;
  const double downdate_threshold = std::sqrt(DBL_EPSILON);
  int nonzero = n;
  for (int k = 0; k < n; k++) {
    int big = k;
    for (int j = k + 1; j < n; j++)
      if (norm_upd[j] > norm_upd[big]) big = j;             // maxCoeff: first maximum
    const double big_sq = norm_upd[big] * norm_upd[big];
    if (nonzero == n && big_sq < threshold_helper * (double)(n - k)) nonzero = k;
    trans[k] = big;
    if (big != k) {
      for (int r = 0; r < n; r++) std::swap(A[k][r], A[big][r]);
      std::swap(norm_upd[k], norm_upd[big]);
      std::swap(norm_dir[k], norm_dir[big]);
    }
    // makeHouseholderInPlace on A[k][k..]
    double tail_sq = 0.0;
    for (int r = k + 1; r < n; r++) tail_sq += A[k][r] * A[k][r];
    const double c0 = A[k][k];
    double beta, tau;
    if (tail_sq <= DBL_MIN) {
      tau = 0.0, beta = c0;
      for (int r = k + 1; r < n; r++) A[k][r] = 0.0;
    } else {
      beta = std::sqrt(c0 * c0 + tail_sq);
      if (c0 >= 0.0) beta = -beta;
      for (int r = k + 1; r < n; r++) A[k][r] = A[k][r] / (c0 - beta);
      tau = (beta - c0) / beta;
    }
    hcoef[k] = tau;
    A[k][k] = beta;
    // applyHouseholderOnTheLeft to the columns right of k
    for (int j = k + 1; j < n; j++) {
      if (n - k == 1) {
        A[j][k] *= (1.0 - tau);
      } else if (tau != 0.0) {
        double tmp = 0.0;
        for (int r = k + 1; r < n; r++) tmp += A[k][r] * A[j][r];
        tmp += A[j][k];
        A[j][k] -= tau * tmp;
        for (int r = k + 1; r < n; r++) A[j][r] -= (tau * A[k][r]) * tmp;
      }
    }
    for (int j = k + 1; j < n; j++) {
      if (norm_upd[j] != 0.0) {
        double temp = std::fabs(A[j][k]) / norm_upd[j];
        temp = (1.0 + temp) * (1.0 - temp);
        temp = temp < 0.0 ? 0.0 : temp;
        const double q = norm_upd[j] / norm_dir[j];
        const double temp2 = temp * (q * q);
        if (temp2 <= downdate_threshold) {
          double s = 0.0;
          for (int r = k + 1; r < n; r++) s += A[j][r] * A[j][r];
          norm_dir[j] = norm_upd[j] = std::sqrt(s);
        } else {
          norm_upd[j] *= std::sqrt(temp);
        }
      }
    }
  }
  int perm[3] = {0, 1, 2};
  for (int k = 0; k < n; k++) std::swap(perm[k], perm[trans[k]]);
  x[0] = x[1] = x[2] = 0.0;
  if (nonzero == 0) return;
  double c[3] = {rhs[0], rhs[1], rhs[2]};
  for (int k = 0; k < nonzero; k++) {                      // Q^T c = H_{r-1} ... H_0 c
    const double tau = hcoef[k];
    if (n - k == 1) {
      c[k] *= (1.0 - tau);
    } else if (tau != 0.0) {
      double tmp = 0.0;
      for (int r = k + 1; r < n; r++) tmp += A[k][r] * c[r];
      tmp += c[k];
      c[k] -= tau * tmp;
      for (int r = k + 1; r < n; r++) c[r] -= (tau * A[k][r]) * tmp;
    }
  }
  for (int i = nonzero - 1; i >= 0; i--) {                 // upper-triangular solve
    double s = c[i];
    for (int j = i + 1; j < nonzero; j++) s -= A[j][i] * c[j];
    c[i] = s / A[i][i];
  }
  for (int i = 0; i < nonzero; i++) x[perm[i]] = c[i];
}

// checkConnections, node.cpp:556-590 (s1 = "drone_seg", s2 = "world_seg")
bool check_connections(const h3d_world_segment &s1, const h3d_world_segment &s2, const h3d_world_config &cfg,
                       double sol[3]) {
  sol[0] = sol[1] = sol[2] = 0.0;
  const V3 p1 = at(s1, s1.t_min), q1 = at(s2, s2.t_min);
  const V3 b1 = ld(s1.b), b2 = ld(s2.b);
  V3 cr = {b2.y * b1.z - b2.z * b1.y, b2.z * b1.x - b2.x * b1.z, b2.x * b1.y - b2.y * b1.x};   // world.b x drone.b
  if (norm(cr) < 1e-2) return false;
  const double z = dot(cr, cr);                             // normalize(): /= sqrt(squaredNorm)
  if (z > 0.0) {
    const double s = std::sqrt(z);
    cr = {cr.x / s, cr.y / s, cr.z / s};
  }
  const V3 rhs = q1 - p1;
  double A[3][3] = {{b1.x, b1.y, b1.z}, {-b2.x, -b2.y, -b2.z}, {cr.x, cr.y, cr.z}};
  const double r[3] = {rhs.x, rhs.y, rhs.z};
  colpiv_qr_solve3(A, r, sol);
  const double dist = std::fabs(sol[2]);
  const double epsilon = (2 * cfg.diag_voxel + s1.radius) + s2.radius;
  return ((sol[0] + s1.t_min) >= s1.t_min) && ((sol[0] + s1.t_min) <= s1.t_max) &&
         ((sol[1] + s2.t_min) >= s2.t_min) && ((sol[1] + s2.t_min) <= s2.t_max) && dist < epsilon;
}

}  // namespace

// callback_info (node.cpp:46-51): one record per processed frame, written by saveProcessingTimeToFile
struct CallbackInfo {
  int wall_time, callback_time;   // microseconds since the start of the run / spent on this frame
  size_t seg_vec_size;            // world_segments.size() after the frame
  int nblines;                    // lines the Hough loop extracted from the frame
};

struct h3d_world {
  h3d_world_config cfg;
  std::vector<h3d_world_segment> segs;              // world_segments
  std::vector<std::vector<double>> inter;           // intersection_matrix, rows of (t1, t2) pairs
  std::vector<CallbackInfo> callbacks;              // callback_info_vec
};

// The node's three verbosity levels (node.cpp:23: NONE, INFO, WARN; /verbose_level): a level above NONE prints
// the callback time, a level above INFO also the warnings, the intersection matrix and the segment list
// (node.cpp:296-346).  ROS_INFO / ROS_WARN become lines on stderr.
static std::atomic<int> g_verbosity{0};

namespace {

// segFiltering, node.cpp:479-542
void seg_filtering(h3d_world &w, const std::vector<h3d_world_segment> &drone) {
  std::vector<size_t> new_idx, mod_idx;
  std::vector<h3d_world_segment> nw = w.segs;
  if (w.segs.empty()) {
    nw = drone;
  } else {
    for (size_t i = 0; i < drone.size(); ++i) {
      bool found = false;
      for (size_t j = 0; j < w.segs.size(); ++j) {
        h3d_world_segment target = w.segs[j];       // defined start of `segment target_seg;` (see h3d.h)
        if (check_similarity(drone[i], w.segs[j], w.cfg, target)) {
          nw[j] = target;
          mod_idx.push_back(j);
          found = true;
          break;
        }
      }
      if (!found) {
        nw.push_back(drone[i]);
        new_idx.push_back(nw.size() + i);           // node.cpp:508, kept as written
      }
    }
  }
  const size_t n = nw.size();
  w.inter.resize(n);
  for (auto &row : w.inter) row.resize(2 * n, -1.0);
  std::vector<char> is_target(n, 0);
  for (size_t k : mod_idx) is_target[k] = 1;
  for (size_t k : new_idx)
    if (k < n) is_target[k] = 1;
  for (size_t i = 0; i < n; ++i)
    for (size_t j = 0; j < i; ++j)
      if (is_target[i] || is_target[j]) {
        double sol[3];
        if (check_connections(nw[i], nw[j], w.cfg, sol)) {
          w.inter[i][2 * j] = nw[i].t_min + sol[0];
          w.inter[i][2 * j + 1] = nw[j].t_min + sol[1];
        }
      }
  w.segs.swap(nw);
}

}  // namespace

extern "C" {

int h3d_world_create(const h3d_world_config *cfg, h3d_world **out) {
  if (!cfg || !out) return h3d::set_error("h3d_world_create: null argument"), H3D_ERR_ARG;
  h3d_world *w = new (std::nothrow) h3d_world();
  if (!w) return h3d::set_error("h3d_world_create: out of memory"), H3D_ERR_NOMEM;
  w->cfg = *cfg;
  *out = w;
  return H3D_OK;
}

void h3d_world_destroy(h3d_world *w) { delete w; }

int h3d_world_update(h3d_world *w, const h3d_world_segment *drone, size_t n) {
  if (!w || (n && !drone)) return h3d::set_error("h3d_world_update: null argument"), H3D_ERR_ARG;
  try {
    seg_filtering(*w, std::vector<h3d_world_segment>(drone, drone + n));
  } catch (const std::bad_alloc &) {
    return h3d::set_error("h3d_world_update: out of memory"), H3D_ERR_NOMEM;
  }
  return H3D_OK;
}

int h3d_world_update_frame(h3d_world *w, const h3d_segment *segs, size_t n, size_t *n_used) {
  if (!w || (n && !segs)) return h3d::set_error("h3d_world_update_frame: null argument"), H3D_ERR_ARG;
  try {
    std::vector<h3d_world_segment> drone;
    for (size_t i = 0; i < n; i++) {
      const h3d_segment &s = segs[i];
      if (s.status != 0 || s.world_status != 1) continue;
      h3d_world_segment d;
      std::memset(&d, 0, sizeof d);
      for (int k = 0; k < 3; k++) d.a[k] = s.a_world[k], d.b[k] = s.b_world[k], d.pca_eigenvalues[k] = s.pca_eigenvalues[k];
      d.t_min = s.t_min, d.t_max = s.t_max, d.radius = s.radius;
      d.points_size = (int32_t)s.n_points;
      d.pca_coeff = s.pca_coeff;
      drone.push_back(d);
    }
    if (n_used) *n_used = drone.size();
    seg_filtering(*w, drone);
  } catch (const std::bad_alloc &) {
    return h3d::set_error("h3d_world_update_frame: out of memory"), H3D_ERR_NOMEM;
  }
  return H3D_OK;
}

int h3d_world_segments(const h3d_world *w, h3d_world_segment *out, size_t cap, size_t *n) {
  if (!w || !n) return h3d::set_error("h3d_world_segments: null argument"), H3D_ERR_ARG;
  *n = w->segs.size();
  if (out)
    for (size_t i = 0; i < w->segs.size() && i < cap; i++) out[i] = w->segs[i];
  return H3D_OK;
}

int h3d_world_intersections(const h3d_world *w, double *t_pairs, size_t cap_pairs, size_t *n) {
  if (!w || !n) return h3d::set_error("h3d_world_intersections: null argument"), H3D_ERR_ARG;
  const size_t m = w->inter.size();
  *n = m;
  if (!t_pairs) return H3D_OK;
  if (cap_pairs < m * m) return h3d::set_error("h3d_world_intersections: buffer too small"), H3D_ERR_ARG;
  for (size_t i = 0; i < m; i++) std::memcpy(t_pairs + 2 * i * m, w->inter[i].data(), sizeof(double) * 2 * m);
  return H3D_OK;
}

int h3d_check_similarity(const h3d_world_segment *drone, const h3d_world_segment *world, const h3d_world_config *cfg,
                         const h3d_world_segment *target_init, h3d_world_segment *target, int *similar) {
  if (!drone || !world || !cfg || !target || !similar)
    return h3d::set_error("h3d_check_similarity: null argument"), H3D_ERR_ARG;
  h3d_world_segment t = target_init ? *target_init : *world;
  *similar = check_similarity(*drone, *world, *cfg, t) ? 1 : 0;
  *target = t;
  return H3D_OK;
}

int h3d_check_connections(const h3d_world_segment *s1, const h3d_world_segment *s2, const h3d_world_config *cfg,
                          double sol[3], int *connected) {
  if (!s1 || !s2 || !cfg || !sol || !connected)
    return h3d::set_error("h3d_check_connections: null argument"), H3D_ERR_ARG;
  *connected = check_connections(*s1, *s2, *cfg, sol) ? 1 : 0;
  return H3D_OK;
}

int h3d_set_verbosity(int level) { return g_verbosity.exchange(level < 0 ? 0 : level > 2 ? 2 : level); }

// The tail of the node's callback (node.cpp:308-346) for one frame: log according to the verbosity level and keep
// the callback_info record.  hough_rc: what hough3dlines() returned for the frame.
int h3d_world_note_frame(h3d_world *w, int wall_time_us, int callback_time_us, int nblines, int hough_rc) {
  if (!w) return h3d::set_error("h3d_world_note_frame: null argument"), H3D_ERR_ARG;
  const int level = g_verbosity.load(std::memory_order_relaxed);
  if (hough_rc != 0 && level > 1) std::fprintf(stderr, "[ WARN] Unable to perform the Hough transform\n");
  if (level > 1) {
    std::fprintf(stderr, "[ INFO] Intersection matrix:\n");
    for (size_t i = 0; i < w->inter.size(); i++)
      for (size_t j = 0; j < i; j++)
        std::fprintf(stderr, "[ INFO] intersection_matrix[%zu][%zu] = (%f, %f)\n", i, j, w->inter[i][2 * j], w->inter[i][2 * j + 1]);
    std::fprintf(stderr, "[ INFO] Number of segments: %zu\n", w->segs.size());
    for (size_t i = 0; i < w->segs.size(); i++)
      std::fprintf(stderr, "[ INFO] Segment %zu: a = (%f, %f, %f), t_min = %f, t_max = %f\n", i, w->segs[i].a[0],
                   w->segs[i].a[1], w->segs[i].a[2], w->segs[i].t_min, w->segs[i].t_max);
  }
  try {
    w->callbacks.push_back(CallbackInfo{wall_time_us, callback_time_us, w->segs.size(), nblines});
  } catch (const std::bad_alloc &) {
    return h3d::set_error("h3d_world_note_frame: out of memory"), H3D_ERR_NOMEM;
  }
  if (level > 0) std::fprintf(stderr, "[ INFO] Callback execution time: %d us\n", callback_time_us);
  return H3D_OK;
}

// saveProcessingTimeToFile (node.cpp:905-919): processing_time.csv, same header and column order
int h3d_world_save_processing_time(const h3d_world *w, const char *path) {
  if (!w || !path) return h3d::set_error("h3d_world_save_processing_time: null argument"), H3D_ERR_ARG;
  FILE *f = std::fopen(path, "w");
  if (!f) return h3d::set_error(std::string("h3d_world_save_processing_time: cannot open ") + path), H3D_ERR_ARG;
  std::fprintf(f, "wall_time,processing_time,seg_vec_size,nblines\n");
  for (const CallbackInfo &c : w->callbacks) std::fprintf(f, "%d,%d,%zu,%d\n", c.wall_time, c.callback_time, c.seg_vec_size, c.nblines);
  std::fclose(f);
  return H3D_OK;
}

int h3d_world_save_csv(const h3d_world *w, const char *segments_path, const char *intersections_path) {
  if (!w) return h3d::set_error("h3d_world_save_csv: null argument"), H3D_ERR_ARG;
  if (segments_path) {
    FILE *f = std::fopen(segments_path, "w");
    if (!f) return h3d::set_error(std::string("h3d_world_save_csv: cannot open ") + segments_path), H3D_ERR_ARG;
    std::fprintf(f, "segment,a_x,a_y,a_z,b_x,b_y,b_z,t_min,t_max\n");
    for (size_t i = 0; i < w->segs.size(); i++) {
      const h3d_world_segment &s = w->segs[i];
      std::fprintf(f, "%zu,%g,%g,%g,%g,%g,%g,%g,%g\n", i, s.a[0], s.a[1], s.a[2], s.b[0], s.b[1], s.b[2], s.t_min,
                   s.t_max);
    }
    std::fclose(f);
  }
  if (intersections_path) {
    FILE *f = std::fopen(intersections_path, "w");
    if (!f) return h3d::set_error(std::string("h3d_world_save_csv: cannot open ") + intersections_path), H3D_ERR_ARG;
    std::fprintf(f, "seg1,t1,seg2,t2\n");
    for (size_t i = 0; i < w->inter.size(); i++)
      for (size_t j = 0; j < i; j++) {
        const double t1 = w->inter[i][2 * j], t2 = w->inter[i][2 * j + 1];
        if (t1 != -1.0 && t2 != -1.0) std::fprintf(f, "%zu,%g,%zu,%g\n", i, t1, j, t2);
      }
    std::fclose(f);
  }
  return H3D_OK;
}

}  // extern "C"
