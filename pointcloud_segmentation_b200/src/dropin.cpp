// dropin.cpp -- host C++ classes of the hough-3d-lines API over the C-ABI of libh3d.so.
// Mirrors libraries/hough-3d-lines/{vector3d,pointcloud,sphere,hough}.cpp: same names, argument
// meaning and error behaviour; all O(N) work goes through include/h3d.h to the GPU.
#include "../include/hough.h"
#include "../include/hough3dlines_b200.h"
#include "../../include/h3d.h"

#include <cmath>
#include <fstream>
#include <new>
#include <sstream>
#include <stdexcept>
#include <string>

namespace {
[[noreturn]] void raise(int rc) {
  if (rc == H3D_ERR_NOMEM) throw std::bad_alloc();
  throw std::runtime_error(std::string("h3d: ") + h3d_last_error());
}
inline void must(int rc) { if (rc != H3D_OK) raise(rc); }
const double *flat(const std::vector<Vector3d> &v) { return v.empty() ? nullptr : &v[0].x; }
double *flat(std::vector<Vector3d> &v) { return v.empty() ? nullptr : &v[0].x; }
static_assert(sizeof(Vector3d) == 3 * sizeof(double), "Vector3d must be three packed doubles");
}  // namespace

// ---- Vector3d (lib/vector3d.cpp) -----------------------------------------------------------
std::ostream &operator<<(std::ostream &os, const Vector3d &v) { return os << v.x << " " << v.y << " " << v.z; }
double Vector3d::norm() const { return std::sqrt((x * x) + (y * y) + (z * z)); }
Vector3d operator+(Vector3d p, Vector3d q) { return Vector3d(p.x + q.x, p.y + q.y, p.z + q.z); }
Vector3d operator-(Vector3d p, Vector3d q) { return Vector3d(p.x - q.x, p.y - q.y, p.z - q.z); }
double operator*(Vector3d p, Vector3d q) { return p.x * q.x + p.y * q.y + p.z * q.z; }
Vector3d operator*(Vector3d p, double c) { return Vector3d(c * p.x, c * p.y, c * p.z); }
Vector3d operator*(double c, Vector3d p) { return Vector3d(c * p.x, c * p.y, c * p.z); }
Vector3d operator/(Vector3d p, double c) { return Vector3d(p.x / c, p.y / c, p.z / c); }

// ---- Sphere (lib/sphere.cpp:18-24) ---------------------------------------------------------
void Sphere::fromIcosahedron(int subDivisions) {
  size_t nv = 0, nt = 0;
  must(h3d_sphere(subDivisions, nullptr, 0, &nv, nullptr, 0, &nt));
  vertices.assign(nv, Vector3d());
  std::vector<unsigned> tri(nt);
  must(h3d_sphere(subDivisions, flat(vertices), nv, &nv, tri.data(), nt, &nt));
  triangles.assign(tri.begin(), tri.end());
}

// ---- PointCloud (lib/pointcloud.cpp) -------------------------------------------------------
void PointCloud::shiftToOrigin() {
  double s[3] = {shift.x, shift.y, shift.z};
  must(h3d_shift_to_origin(flat(points), points.size(), s));
  shift = Vector3d(s[0], s[1], s[2]);
}
Vector3d PointCloud::meanValue() const {
  double m[3];
  must(h3d_mean_value(flat(points), points.size(), m));
  return Vector3d(m[0], m[1], m[2]);
}
void PointCloud::getMinMax3D(Vector3d *min_pt, Vector3d *max_pt) {
  double mn[3], mx[3];
  must(h3d_min_max(flat(points), points.size(), mn, mx));
  *min_pt = Vector3d(mn[0], mn[1], mn[2]);
  *max_pt = Vector3d(mx[0], mx[1], mx[2]);
}
// lib/pointcloud.cpp:63-91: '#' comment lines, three delimiter-separated numbers per line;
// rc 0 ok, 1 cannot open, 2 malformed line
int PointCloud::readFromFile(const char *path, char delim) {
  std::ifstream in(path);
  if (!in) return 1;
  std::string line;
  while (std::getline(in, line)) {
    if (!line.empty() && line[0] == '#') continue;
    if (line.empty()) return 2;   // sscanf finds no number on an empty line: n != 3 (lib/pointcloud.cpp:72-77)
    std::stringstream ss(line);
    std::string field;
    double v[3];
    int k = 0;
    while (k < 3 && std::getline(ss, field, delim)) {
      try { v[k] = std::stod(field); } catch (...) { return 2; }
      ++k;
    }
    if (k < 3) return 2;
    points.push_back(Vector3d(v[0], v[1], v[2]));
  }
  return 0;
}
void PointCloud::pointsCloseToLine(const Vector3d &a, const Vector3d &b, double dx, PointCloud *Y) {
  Y->points.clear();
  std::vector<uint32_t> idx(points.size());
  size_t k = 0;
  const double aa[3] = {a.x, a.y, a.z}, bb[3] = {b.x, b.y, b.z};
  must(h3d_points_close_to_line(flat(points), points.size(), aa, bb, dx, idx.data(), &k));
  Y->points.reserve(k);
  for (size_t i = 0; i < k; ++i) Y->points.push_back(points[idx[i]]);
}
void PointCloud::removePoints(const PointCloud &Y) {
  if (Y.points.empty()) return;
  size_t k = points.size();
  must(h3d_remove_points(flat(points), points.size(), flat(Y.points), Y.points.size(), &k));
  points.resize(k);
}

// ---- Hough (lib/hough.cpp) -----------------------------------------------------------------
static Sphere *g_hough_space = nullptr;

void initHoughSpace(unsigned int sphereGranularity) {
  must(h3d_init_space(sphereGranularity));
  g_hough_space = new Sphere();   // never freed, like the reference's global (lib/hough.cpp:15-20)
  g_hough_space->fromIcosahedron((int)sphereGranularity);
}

Hough::Hough(const Vector3d &minP, const Vector3d &maxP, double var_dx) : sphere(g_hough_space), impl_(nullptr) {
  const double mn[3] = {minP.x, minP.y, minP.z}, mx[3] = {maxP.x, maxP.y, maxP.z};
  must(h3d_create(mn, mx, var_dx, &impl_));
  must(h3d_info(impl_, &num_x, &num_b, &dx, &max_x));
}
Hough::~Hough() { h3d_destroy(impl_); }
void Hough::add(const PointCloud &pc) { must(h3d_add(impl_, flat(pc.points), pc.points.size())); }
void Hough::subtract(const PointCloud &pc) { must(h3d_subtract(impl_, flat(pc.points), pc.points.size())); }
unsigned int Hough::getLine(Vector3d *a, Vector3d *b) {
  double aa[3], bb[3];
  unsigned votes = 0;
  must(h3d_get_line(impl_, aa, bb, &votes, nullptr));
  *a = Vector3d(aa[0], aa[1], aa[2]);
  *b = Vector3d(bb[0], bb[1], bb[2]);
  return votes;
}
void Hough::syncVotingSpace() {
  VotingSpace.resize(num_x * num_x * num_b);
  must(h3d_download_votes(impl_, VotingSpace.data(), VotingSpace.size()));
}

double orthogonal_LSQ_b200(const PointCloud &pc, Vector3d *a, Vector3d *b) {
  double aa[3], bb[3], rc = 0.0;
  must(h3d_orthogonal_lsq(flat(pc.points), pc.points.size(), aa, bb, &rc));
  *a = Vector3d(aa[0], aa[1], aa[2]);
  *b = Vector3d(bb[0], bb[1], bb[2]);
  return rc;
}

// ---- fused loop ------------------------------------------------------------------------------
int hough3dlines_b200(const double *xyz, size_t n, std::vector<Line3d> &lines, double opt_dx, int granularity,
                      int opt_minvotes, int opt_nlines, int &nblines_extracted) {
  // opt_nlines == 0 means "until exhausted" (wrap.h:341-342): the record and index buffers start from a guess
  // and the call is repeated with the sizes it reports when they were too small -- no line is ever dropped
  size_t max_iters = opt_nlines > 0 ? (size_t)opt_nlines + 1 : 256;
  size_t pool_cap = 2 * max_iters * (n ? n : 1);
  const size_t pool_guess = (size_t)64 << 20;   // 256 MB of indices before the first answer is known
  if (pool_cap > pool_guess) pool_cap = pool_guess > 4 * n ? pool_guess : 4 * (n ? n : 1);
  std::vector<h3d_iter> iters;
  std::vector<uint32_t> pool;
  h3d_result res;
  for (int attempt = 0; attempt < 3; ++attempt) {
    iters.resize(max_iters);
    pool.resize(pool_cap);
    if (h3d_hough3dlines(xyz, n, granularity, opt_dx, opt_minvotes, opt_nlines, 0, iters.data(), max_iters,
                         pool.data(), pool.size(), &res) != H3D_OK)
      return 1;
    if (res.rc != 0) break;
    if (res.n_iters <= max_iters && res.pool_used <= pool_cap) break;
    if (res.n_iters > max_iters) max_iters = (size_t)res.n_iters + 1;
    // pool_used saturates at the capacity: when it filled up, size for the worst case of the iterations seen
    if (res.pool_used >= pool_cap) pool_cap = 2 * max_iters * (n ? n : 1);
  }
  nblines_extracted = 0;
  if (res.rc != 0) return 1;
  if (res.n_iters > max_iters) return 1;   // did not converge to a consistent size (cannot happen: sizes only grow)
  const size_t n_rec = res.n_iters < max_iters ? res.n_iters : max_iters;
  for (size_t k = 0; k < n_rec; ++k) {
    const h3d_iter &it = iters[k];
    if (it.status != 0) break;
    Line3d l;
    l.a = Vector3d(it.a2[0], it.a2[1], it.a2[2]);
    l.b = Vector3d(it.b2[0], it.b2[1], it.b2[2]);
    l.votes = it.votes;
    l.inliers.assign(pool.begin() + it.off2, pool.begin() + it.off2 + it.n2);
    lines.push_back(l);
  }
  nblines_extracted = (int)res.n_lines;
  return 0;
}
