// h3d_sphere.cpp -- host-side direction table (tessellated icosahedron).
//
// Replaces Sphere::fromIcosahedron / getIcosahedron / subDivide / makeUnique of
// the reference (libraries/hough-3d-lines/sphere.cpp:18-330).  The table must be
// bit-identical to the reference's `Sphere::vertices` (values AND order) because
// the direction index j is part of every accumulator index.
//
// Design: the reference de-duplicates midpoints with a linear search over all
// vertices created so far in the subdivision step (sphere.cpp:197-214), which is
// O(V^2): 15.6 s at granularity 6.  A midpoint's value depends only on its two
// end points (IEEE addition commutes, the normalisation is a pure function), so
// "same value" is "same edge" and an edge-keyed hash map yields the same
// first-seen numbering in O(V).  g6 builds in a few milliseconds.
//
// Compile with -ffp-contract=off: no fused multiply-add may be introduced.
#include "h3d_internal.h"

#include <cmath>
#include <cstdint>
#include <unordered_map>
#include <vector>

namespace h3d {

namespace {

struct V3 { double x, y, z; };

inline V3 unit_midpoint(const V3 &p, const V3 &q) {
  // sphere.cpp:166-192: component sums, Euclidean norm, three true divisions
  V3 m{p.x + q.x, p.y + q.y, p.z + q.z};
  const double len = std::sqrt(m.x * m.x + m.y * m.y + m.z * m.z);
  m.x /= len; m.y /= len; m.z /= len;
  return m;
}

inline uint64_t edge_key(uint32_t a, uint32_t b) {
  return a < b ? (uint64_t(b) << 32) | a : (uint64_t(a) << 32) | b;
}

}  // namespace

void build_sphere(int granularity, std::vector<double> &vertices, std::vector<unsigned> &triangles) {
  // --- icosahedron (sphere.cpp:27-145); the two constants are computed in float
  float tau = 1.61803399f;
  const float len = std::sqrt(1 + tau * tau);
  const float one = 1 / len;
  tau = tau / len;
  const double s = one, t = tau;
  std::vector<V3> vert = {{-s, t, 0}, {s, t, 0},  {0, s, -t}, {0, s, t},  {-t, 0, -s}, {t, 0, -s},
                          {-t, 0, s}, {t, 0, s},  {0, -s, -t}, {0, -s, t}, {-s, -t, 0}, {s, -t, 0}};
  std::vector<uint32_t> tri = {0, 1, 2,  0, 1, 3,  0, 2, 4,   0, 4, 6,   0, 3, 6,  1, 2, 5,  1, 3, 7,
                               1, 5, 7,  2, 4, 8,  2, 5, 8,   3, 6, 9,   3, 7, 9,  4, 8, 10, 8, 10, 11,
                               5, 8, 11, 5, 7, 11, 7, 9, 11,  9, 10, 11, 6, 9, 10, 4, 6, 10};

  // --- subdivisions (sphere.cpp:148-244)
  for (int level = 0; level < granularity; ++level) {
    const size_t n_tri = tri.size() / 3;
    std::vector<uint32_t> next;
    next.reserve(n_tri * 12);
    std::unordered_map<uint64_t, uint32_t> midpoint_of;
    midpoint_of.reserve(n_tri * 2);
    auto vertex_on = [&](uint32_t p, uint32_t q, uint32_t sum_first, uint32_t sum_second) {
      auto it = midpoint_of.find(edge_key(p, q));
      if (it != midpoint_of.end()) return it->second;
      const uint32_t id = uint32_t(vert.size());
      vert.push_back(unit_midpoint(vert[sum_first], vert[sum_second]));
      midpoint_of.emplace(edge_key(p, q), id);
      return id;
    };
    for (size_t k = 0; k < n_tri; ++k) {
      const uint32_t a = tri[3 * k], b = tri[3 * k + 1], c = tri[3 * k + 2];
      // creation order d(a,b), e(c,b), f(a,c) as in sphere.cpp:166-226
      const uint32_t d = vertex_on(a, b, a, b);
      const uint32_t e = vertex_on(b, c, c, b);
      const uint32_t f = vertex_on(c, a, a, c);
      const uint32_t four[12] = {a, d, f, d, b, e, f, e, c, f, d, e};
      next.insert(next.end(), four, four + 12);
    }
    tri.swap(next);
  }

  // --- makeUnique (sphere.cpp:247-330): keep the z>0 hemisphere plus half of the
  // equator; surviving vertices keep their relative order, triangles that lose a
  // vertex disappear.
  std::vector<uint32_t> new_id(vert.size());
  vertices.clear();
  uint32_t kept = 0;
  for (size_t i = 0; i < vert.size(); ++i) {
    const V3 &p = vert[i];
    const bool dropped = p.z < 0 || (p.z == 0 && (p.x < 0 || (p.x == 0 && p.y == -1)));
    if (dropped) { new_id[i] = UINT32_MAX; continue; }
    new_id[i] = kept++;
    vertices.push_back(p.x); vertices.push_back(p.y); vertices.push_back(p.z);
  }
  triangles.clear();
  for (size_t k = 0; k + 2 < tri.size(); k += 3) {
    const uint32_t a = new_id[tri[k]], b = new_id[tri[k + 1]], c = new_id[tri[k + 2]];
    if (a == UINT32_MAX || b == UINT32_MAX || c == UINT32_MAX) continue;
    triangles.push_back(a); triangles.push_back(b); triangles.push_back(c);
  }
}

}  // namespace h3d
