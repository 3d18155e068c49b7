// hough3dlines_b200.h -- the fused replacement of hough3dlines() (wrap.h:167-350) for consumers
// that want the whole loop on the device: one call per frame, no per-iteration host round trip.
// Segment post-tests (wrap.h:261-337) are not part of it; it returns the fitted lines and their
// inlier sets, from which the caller runs its own tests.
#ifndef HOUGH3DLINES_B200_H_
#define HOUGH3DLINES_B200_H_

#include "vector3d.h"
#include <cstdint>
#include <vector>

struct Line3d {
  Vector3d a, b;                     // anchor (already + X.shift) and direction, wrap.h:254-257
  unsigned votes;                    // Hough::getLine return value of the iteration
  std::vector<uint32_t> inliers;     // indices (into the input cloud) of the final inlier set Y
};

// xyz: n points, 3 doubles each.  Same parameters and return convention as wrap.h:167 (0 ok,
// 1 degenerate cloud / dx too large / allocation failure).
int hough3dlines_b200(const double *xyz, size_t n, std::vector<Line3d> &lines, double opt_dx, int granularity,
                      int opt_minvotes, int opt_nlines, int &nblines_extracted);

#endif /* HOUGH3DLINES_B200_H_ */
