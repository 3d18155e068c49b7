// vector3d.h -- drop-in for libraries/hough-3d-lines/vector3d.h:15-43 (same class, same
// members, same free operators).  Plain host value type: three doubles.
#ifndef VECTOR3D_H_
#define VECTOR3D_H_

#include <iostream>

class Vector3d {
public:
  double x;
  double y;
  double z;
  Vector3d() : x(0.0), y(0.0), z(0.0) {}
  Vector3d(double a, double b, double c) : x(a), y(b), z(c) {}
  bool operator==(const Vector3d &rhs) const { return x == rhs.x && y == rhs.y && z == rhs.z; }
  Vector3d &operator=(const Vector3d &other) = default;
  Vector3d(const Vector3d &other) = default;
  friend std::ostream &operator<<(std::ostream &os, const Vector3d &vec);
  // Euclidean norm, lib/vector3d.cpp:37-39 (same summation order)
  double norm() const;
};

Vector3d operator+(Vector3d x, Vector3d y);
Vector3d operator-(Vector3d x, Vector3d y);
double operator*(Vector3d x, Vector3d y);
Vector3d operator*(Vector3d x, double c);
Vector3d operator*(double c, Vector3d x);
Vector3d operator/(Vector3d x, double c);

#endif /* VECTOR3D_H_ */
