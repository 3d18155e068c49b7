// sphere.h -- drop-in for libraries/hough-3d-lines/sphere.h:19-37.  fromIcosahedron() fills the
// same public members from h3d_sphere() (O(V) edge-hash build, bit-identical table).
#ifndef SPHERE_H_
#define SPHERE_H_

#include "vector3d.h"
#include <deque>
#include <vector>

class Sphere {
public:
  std::vector<Vector3d> vertices;       // direction vectors
  std::deque<unsigned int> triangles;   // surface triangles
  void fromIcosahedron(int subDivisions = 4);
};

#endif /* SPHERE_H_ */
