#!/usr/bin/env python
"""One pass of the configs[3] loop (1 M points, g6, 50 lines) for profiling: ncu launch list source."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pointcloud_segmentation_b200 import capi as h3d, synth  # noqa: E402

pts = synth.infrastructure()
tr = h3d.hough3dlines(pts, 6, 0.0, 10, 50, mode=1)
print(tr.rc, tr.n_lines, tr.votes_cast)
