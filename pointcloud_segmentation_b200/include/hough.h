// hough.h -- drop-in for libraries/hough-3d-lines/hough.h:20-46.  Same class name, public members
// and signatures; the accumulator lives on the GPU (never materialised).
// `VotingSpace` is kept for source compatibility and filled on demand by syncVotingSpace().
#ifndef HOUGH_H_
#define HOUGH_H_

#include "pointcloud.h"
#include "sphere.h"
#include "vector3d.h"
#include <cstddef>
#include <vector>

struct h3d_hough;

void initHoughSpace(unsigned int sphereGranularity = 4);

class Hough {
public:
  std::vector<unsigned int> VotingSpace;   // dense accumulator, valid after syncVotingSpace()
  Sphere *sphere;                          // directions B (borrowed, like the reference)
  size_t num_b;
  double dx, max_x;
  size_t num_x;

  // throws std::runtime_error / std::bad_alloc on CUDA failure (caught at wrap.h:221)
  Hough(const Vector3d &minP, const Vector3d &maxP, double dx);
  ~Hough();
  Hough(const Hough &) = delete;
  Hough &operator=(const Hough &) = delete;
  unsigned int getLine(Vector3d *point, Vector3d *direction);
  void add(const PointCloud &pc);
  void subtract(const PointCloud &pc);
  // extension: download the dense accumulator (num_x*num_x*num_b cells, reference layout)
  void syncVotingSpace();

private:
  h3d_hough *impl_;
};

// wrap.h:121-150 defines orthogonal_LSQ in the consumer (Eigen); this is the accelerated
// equivalent with the same signature (h3d_orthogonal_lsq).
double orthogonal_LSQ_b200(const PointCloud &pc, Vector3d *a, Vector3d *b);

#endif /* HOUGH_H_ */
