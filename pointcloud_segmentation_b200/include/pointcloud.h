// pointcloud.h -- drop-in for libraries/hough-3d-lines/pointcloud.h:17-38.  Same public members
// and methods; the O(N) methods run on the GPU through the C-ABI (include/h3d.h).
#ifndef POINTCLOUD_H_
#define POINTCLOUD_H_

#include "vector3d.h"
#include <string>
#include <vector>

class PointCloud {
public:
  Vector3d shift;                  // translation applied by shiftToOrigin()
  std::vector<Vector3d> points;
  void shiftToOrigin();                                        // h3d_shift_to_origin
  Vector3d meanValue() const;                                  // h3d_mean_value
  void getMinMax3D(Vector3d *min_pt, Vector3d *max_pt);        // h3d_min_max
  int readFromFile(const char *path, char delim);              // host I/O (rc 0/1/2 as the reference)
  void pointsCloseToLine(const Vector3d &a, const Vector3d &b, double dx, PointCloud *Y);   // h3d_points_close_to_line
  void removePoints(const PointCloud &Y);                      // h3d_remove_points
};

#endif /* POINTCLOUD_H_ */
