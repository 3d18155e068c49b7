"""A/B of the streaming vote kernel forms on BASELINE configs[3] (1 M points, g6, dx = d/64) and a 38 k / 200 k
cloud: mode 1 = TMA-staged tiles (cp.async.bulk + mbarrier), 0 = same loop with LDGs, 2 = round-1 vote_kernel."""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pointcloud_segmentation_b200 import capi as h3d, synth  # noqa: E402

h3d.init_space(6)
out = []
for n in (38528, 200000, 1000000):
    pts = synth.infrastructure(n)
    x, _ = h3d.shift_to_origin(pts)
    mn, mx = h3d.min_max(x)
    dx = float(np.sqrt(((mx - mn) ** 2).sum())) / 64.0
    row = {"points": n}
    for mode, name in ((1, "tma"), (0, "ldg"), (2, "round1")):
        h3d.set_stream_mode(mode)
        hg = h3d.Hough(mn, mx, dx)
        hg.add(x)
        us = hg.bench_vote(reps=3)
        key = hg.get_line_key()
        hg.close()
        row[name + "_ms"] = us / 1e3
        row[name + "_gvotes_per_s"] = n * 20481 / us / 1e3
        row[name + "_key"] = key
    row["same_winner"] = row["tma_key"] == row["ldg_key"] == row["round1_key"]
    out.append(row)
    print(json.dumps(row), flush=True)
h3d.set_stream_mode(0)
