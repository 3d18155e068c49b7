"""A/B of the first vote of the frame fast path on the 16 bench frames: TRACK (atomics return the old count,
maxima tracked per thread, planes only zero-filled) against SCAN (merging RED increments + plane scan)."""
import os, sys, json
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pointcloud_segmentation_b200 import capi, synth
import bench

frames = bench.make_frames(16)
# the round-1 path for comparison: eleven K0 launches, Hough ctor on the host, vote_resident_kernel<P> picked by the host
capi.set_fast_frames(False)
pipe = capi.Pipeline(6, synth.WINDOW, synth.LEAF, synth.OPT_DX, 12, 20, max_raw_points=frames.shape[1])
us = []
for rep in range(4):
    for k in range(16):
        res = pipe.run(frames[k])
        if rep:
            us.append((res.n_input, *pipe.timing_us()))
pipe.close()
a = np.array(us)
print(json.dumps({"share": 148, "path": "general (round 1)", "k0_us": a[:, 1].mean(), "first_vote_us": a[:, 2].mean(),
                  "loop_us": a[:, 3].mean(), "frame_us": a[:, 4].mean(),
                  "first_vote_us_by_frame": [round(float(x), 1) for x in a[:16, 2]]}))
capi.set_fast_frames(True)
for share in (0, 18):
    for track in (2, 1, 0):
        capi.set_vote_track(track)
        pipe = capi.Pipeline(6, synth.WINDOW, synth.LEAF, synth.OPT_DX, 12, 20, max_raw_points=frames.shape[1])
        pipe.set_sm_share(share)
        us = []
        for rep in range(4):
            for k in range(16):
                res = pipe.run(frames[k])
                if rep:
                    us.append((res.n_input, *pipe.timing_us()))
        st = np.array(pipe.k0_stamps(), dtype=np.int64)
        pipe.close()
        a = np.array(us)
        if track == 2:
            print("K0 cluster kernel phases (us) on the last frame, share", share or 148, ":",
                  dict(zip(("A_count", "B1_reduce", "B2_scan", "C_scatter", "C2_order", "D_centroid", "E_shift_ctor"),
                           [round(float(x) / 1e3, 2) for x in np.diff(st[:8])])))
        print(json.dumps({"share": share or 148, "vote": ("scan", "track", "auto")[track], "k0_us": a[:, 1].mean(),
                          "first_vote_us": a[:, 2].mean(), "loop_us": a[:, 3].mean(), "frame_us": a[:, 4].mean(),
                          "gvotes_per_s": float((a[:, 0] * 20481 / a[:, 2] / 1e3).mean()),
                          "first_vote_us_by_frame": [round(float(x), 1) for x in a[:16, 2]]}))
capi.set_vote_track(2)
