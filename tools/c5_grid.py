#!/usr/bin/env python
"""BASELINE.json configs[4] (SURVEY 8d "C5"): vote kernel over granularity 0..6 x N in {10k, 38k, 100k, 1M, 4M},
uniform-in-beam clouds, dx = d/64 (num_x 64).  Per point: time of one full add (h3d_bench_vote, CUDA events,
mean of `reps` launches), votes/s, the algorithmic GB/s of SURVEY 8(d) (8 B/vote + 24 B/point + 40 B/direction)
and its fraction of the measured HBM peak; ablations (projection arithmetic only / shared atomics only) at
g6 give the two secondary ceilings."""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pointcloud_segmentation_b200 import capi, synth  # noqa: E402

peak = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
print(f"# measured HBM peak {peak} GB/s (MEASURED_PEAKS.json); roofline of the dense algorithm = {peak / 8:.0f} Gvotes/s")
print(f"# {'n':>8} {'g':>2} {'num_b':>6} {'us/add':>10} {'Gvotes/s':>9} {'alg GB/s':>9} {'frac':>5}  form")
rows = []
for n in (10_000, 38_000, 100_000, 1_000_000, 4_000_000):
    pts = synth.beam_cloud(synth.tower_beams(), n, seed=3)
    mn, mx = pts.min(0), pts.max(0)
    x = pts - (mn + mx) / 2.0
    mn, mx = x.min(0), x.max(0)
    for g in range(7):
        capi.init_space(g)
        h = capi.Hough(mn, mx, 0.0)
        h.add(x)
        votes = n * h.num_b
        reps = 3 if votes > 2e10 else 5
        for ab in ((0, 1, 2) if g == 6 else (0,)):
            us = h.bench_vote(reps=reps, ablate=ab)
            gbs = (8 * votes + 24 * n + 40 * h.num_b) / us / 1e3
            form = {0: "full kernel", 1: "projection only", 2: "shared atomics only"}[ab]
            print(f"  {n:>8} {g:>2} {h.num_b:>6} {us:>10.1f} {votes / us / 1e3:>9.1f} {gbs:>9.0f} {gbs / peak:>5.2f}  {form}", flush=True)
            rows.append({"n": n, "g": g, "num_b": h.num_b, "us": us, "gvotes_s": votes / us / 1e3, "alg_gbs": gbs,
                         "frac": gbs / peak, "ablate": ab})
        h.close()
os.makedirs("gpurun_out", exist_ok=True)
json.dump(rows, open("gpurun_out/c5_grid.json", "w"), indent=1)
