"""Phase times of orthogonal_LSQ (one CTA: means incl. the pinned sequential anchor, scatter sums, eigen-solve) on
inlier sets of configs[3]'s size (~11 k points) and of a ToF frame's (~300), doubles with full mantissas and
widened floats."""
# NOTE: a globaltimer read placed right after __syncthreads() (BAR.SYNC.DEFER_BLOCKING) is NOT held back by the
# barrier: thread 0's 'to barrier 1' stamp fires when ITS warp arrives, not when the last warp does.  The two stamps
# taken by the warps themselves ('float mean done', 'double mean done') are the reliable ones.
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pointcloud_segmentation_b200 import capi

rng = np.random.default_rng(5)
for n in (300, 3000, 11000, 50000):
    t = rng.uniform(-20.0, 20.0, n)
    b = np.array([0.6, -0.3, 0.74]); b /= np.linalg.norm(b)
    pts = np.array([3.0, -7.0, 2.5]) + t[:, None] * b + rng.normal(0, 0.02, (n, 3))
    w = pts.astype(np.float32).astype(np.float64)
    for name, p in (("f64", pts), ("widened f32", w), ("f64 again", pts), ("widened+1e-9", w + 1e-9), ("f64 rounded 1e-6", np.round(pts, 6))):
        capi.bench_lsq(p)
        us = capi.bench_lsq(p)
        print(f"n={n:6d} {name:18s} to barrier 1 {us[0]:7.1f} us  to barrier 2 {us[1]:7.1f} us  eig {us[2]:6.1f} us  total {sum(us[:3]):7.1f} us | float mean done {us[3]:6.1f}  double mean done {us[4]:6.1f}")
