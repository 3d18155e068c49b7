// h3d_internal.h -- declarations shared by the translation units of libh3d.so.
#ifndef H3D_INTERNAL_H_
#define H3D_INTERNAL_H_

#include <cstddef>
#include <cstdint>
#include <memory>
#include <string>
#include <vector>

namespace h3d {

// h3d_sphere.cpp: Sphere::fromIcosahedron (lib/sphere.cpp:18-330)
void build_sphere(int granularity, std::vector<double> &vertices_xyz, std::vector<unsigned> &triangles);

// Per-direction projection coefficients of lib/hough.cpp:72-84, evaluated on the
// host in the reference's own expression order (they depend on b only):
//   beta = 1 / (1 + b.z)
//   c0 = 1 - beta*(b.x*b.x)      c1 = beta*(b.x*b.y)      c2 = 1 - beta*(b.y*b.y)
//   x' = ((c0*p.x) - (c1*p.y)) - (b.x*p.z)
//   y' = ((c2*p.y) - (c1*p.x)) - (b.y*p.z)      [(-c1*p.x) + (c2*p.y) reordered:
//                                                IEEE addition commutes]
struct DirCoef {
  double c0, c1, c2, bx, by;
};

// One direction table, host copy + device copy (built once per granularity).
struct Space {
  unsigned granularity = 0;
  size_t num_b = 0;
  std::vector<double> dirs;        // 3*num_b, Sphere::vertices
  std::vector<unsigned> triangles; // Sphere::triangles
  DirCoef *coef_dev = nullptr;     // num_b entries
  double *dirs_dev = nullptr;      // 3*num_b
};

// returns the cached space of granularity g (building and uploading on first use)
int get_space(unsigned g, std::shared_ptr<const Space> *out);
std::shared_ptr<const Space> current_space();

void set_error(const std::string &msg);

}  // namespace h3d

#endif  // H3D_INTERNAL_H_
