// Host-side check that the product's eig3f (h3d_fit.cuh, compiled for the host) is bit-identical
// to the oracle's h3do_eig3f on random symmetric matrices.  Test infrastructure.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include "../../pointcloud_segmentation_b200/csrc/h3d_fit.cuh"
extern "C" int h3do_eig3f(const float m_in[9], float evals[3], float evecs[9]);

static unsigned long long rng = 0x9E3779B97F4A7C15ull;
static float urand() {
  rng ^= rng << 13; rng ^= rng >> 7; rng ^= rng << 17;
  return (float)((rng >> 40) / 16777216.0) * 2.0f - 1.0f;
}

int main(int argc, char **argv) {
  const int n = argc > 1 ? atoi(argv[1]) : 200000;
  int bad = 0;
  for (int t = 0; t < n; ++t) {
    float m[9];
    const int kind = t % 5;
    if (kind < 3) {           // scatter matrix of a few points: PSD like the product's input
      float s[6] = {0, 0, 0, 0, 0, 0};
      const int k = 2 + (t % 17);
      const float sx = kind == 0 ? 1.0f : 0.01f, sy = kind == 1 ? 1.0f : 0.05f;
      for (int i = 0; i < k; ++i) {
        const float x = urand() * sx, y = urand() * sy, z = urand();
        s[0] += x * x; s[1] += x * y; s[2] += x * z; s[3] += y * y; s[4] += y * z; s[5] += z * z;
      }
      const float mm[9] = {s[0], s[1], s[2], s[1], s[3], s[4], s[2], s[4], s[5]};
      memcpy(m, mm, sizeof m);
    } else if (kind == 3) {   // general symmetric
      const float a = urand(), b = urand(), c = urand(), d = urand(), e = urand(), f = urand();
      const float mm[9] = {a, b, c, b, d, e, c, e, f};
      memcpy(m, mm, sizeof m);
    } else {                  // degenerate: diagonal / rank one / zero
      const float a = (t % 3) ? urand() : 0.0f, d = (t % 7) ? urand() : 0.0f;
      const float mm[9] = {a, 0, 0, 0, d, 0, 0, 0, (t % 11) ? a : 0.0f};
      memcpy(m, mm, sizeof m);
    }
    const h3d::Eig3f e = h3d::eig3f(m);
    float ev[3], evec[9];
    const int rc = h3do_eig3f(m, ev, evec);
    if (memcmp(e.evals, ev, sizeof ev) || memcmp(e.evecs, evec, sizeof evec) || (rc == 0) != (e.ok == 1)) {
      if (++bad < 5) printf("mismatch at %d\n", t);
    }
  }
  printf("%d matrices, %d mismatches\n", n, bad);
  return bad ? 1 : 0;
}
