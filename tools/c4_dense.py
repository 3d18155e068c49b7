#!/usr/bin/env python
"""BASELINE.json configs[3] on one B200: 1 M-point infrastructure cloud, g6, dx = d/64, opt_nlines 50.
Prints the vote-kernel time of one full add (h3d_bench_vote, CUDA events) and the wall time of the whole
loop through the C-ABI with HOST points (upload, 50 iterations, inlier lists back)."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pointcloud_segmentation_b200 import capi as h3d, synth  # noqa: E402

pts = synth.infrastructure()
n = pts.shape[0]
h3d.init_space(6)
x, shift = h3d.shift_to_origin(pts)
mn, mx = h3d.min_max(x)
out = {"points": n, "granularity": 6, "nlines": 50}
for rep in range(3):
    t = time.perf_counter()
    tr = h3d.hough3dlines(pts, 6, 0.0, 10, 50, mode=1)
    out[f"loop_wall_ms_run{rep}"] = (time.perf_counter() - t) * 1e3
out.update(rc=tr.rc, lines=tr.n_lines, num_x=tr.num_x, num_b=tr.num_b, votes_cast=tr.votes_cast,
           first_line_votes=tr.iters[0]["votes"], inliers_removed=int(sum(len(i["inliers2"]) for i in tr.iters)))
hg = h3d.Hough(mn, mx, tr.dx)
hg.add(x)
us = hg.bench_vote(reps=5)
votes = n * hg.num_b
out.update(vote_us=us, gvotes_per_s=votes / us / 1e3, algorithmic_gbs=(8 * votes + 24 * n + 40 * hg.num_b) / us / 1e3)
peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))
out["measured_peaks"] = peak
print(json.dumps(out))
