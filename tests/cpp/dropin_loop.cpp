// dropin_loop.cpp -- the reference's loop (wrap.h:237-259, 339-342 / lib/hough3dlines.cpp:200-317)
// written against the DROP-IN headers exactly as the consumer writes it against the reference's:
// Hough, PointCloud, Vector3d, initHoughSpace.  Prints one line per detected line; the pytest
// wrapper compares it with the oracle.  usage: dropin_loop <file.dat> <granularity> <dx> <minvotes> <nlines>
#include "hough.h"
#include "pointcloud.h"
#include "vector3d.h"

#include <cstdio>
#include <cstdlib>

int main(int argc, char **argv) {
  if (argc < 6) return 2;
  const int granularity = atoi(argv[2]);
  double opt_dx = atof(argv[3]);
  int opt_minvotes = atoi(argv[4]);
  const int opt_nlines = atoi(argv[5]);
  if (opt_minvotes < 2) opt_minvotes = 2;
  PointCloud X;
  if (X.readFromFile(argv[1], ',') != 0) return 3;
  initHoughSpace((unsigned)granularity);
  Vector3d minP, maxP, minPshifted, maxPshifted;
  X.getMinMax3D(&minP, &maxP);
  const double d = (maxP - minP).norm();
  if (d == 0.0) return 1;
  X.shiftToOrigin();
  X.getMinMax3D(&minPshifted, &maxPshifted);
  if (opt_dx == 0.0) opt_dx = d / 64.0;
  else if (opt_dx >= d) return 1;
  Hough *hough;
  try { hough = new Hough(minPshifted, maxPshifted, opt_dx); } catch (const std::exception &e) { fprintf(stderr, "%s\n", e.what()); return 1; }
  hough->add(X);
  PointCloud Y;
  double rc;
  unsigned int nvotes;
  int nlines = 0;
  do {
    Vector3d a, b;
    hough->subtract(Y);
    nvotes = hough->getLine(&a, &b);
    X.pointsCloseToLine(a, b, opt_dx, &Y);
    rc = orthogonal_LSQ_b200(Y, &a, &b);
    if (rc == 0.0) break;
    X.pointsCloseToLine(a, b, opt_dx, &Y);
    const unsigned int np = Y.points.size();
    if (np < (unsigned int)opt_minvotes) break;
    rc = orthogonal_LSQ_b200(Y, &a, &b);
    if (rc == 0.0) break;
    a = a + X.shift;
    nlines++;
    printf("votes=%u npoints=%u a=%.17g,%.17g,%.17g b=%.17g,%.17g,%.17g\n", nvotes, np, a.x, a.y, a.z, b.x, b.y, b.z);
    X.removePoints(Y);
  } while ((X.points.size() > 1) && ((opt_nlines == 0) || (opt_nlines > nlines)));
  delete hough;
  return 0;
}
