"""One orthogonal_LSQ call on 11 k points (full-mantissa doubles, then widened floats): the command ncu captures."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pointcloud_segmentation_b200 import capi
rng = np.random.default_rng(5)
n = 11000
t = rng.uniform(-20.0, 20.0, n)
b = np.array([0.6, -0.3, 0.74]); b /= np.linalg.norm(b)
pts = np.array([3.0, -7.0, 2.5]) + t[:, None] * b + rng.normal(0, 0.02, (n, 3))
print(capi.bench_lsq(pts))
print(capi.bench_lsq(pts.astype(np.float32).astype(np.float64)))
