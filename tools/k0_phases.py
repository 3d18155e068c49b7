"""Per-phase time of the K0 cluster kernel (frame_prefilter_kernel) over the 16 bench frames: globaltimer
stamps written by thread 0 at the start, after each cluster barrier and at the end (h3d_pipeline_k0_stamps).
argv: [sm_share (0 = all SMs)]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pointcloud_segmentation_b200 import capi, synth

share = int(sys.argv[1]) if len(sys.argv) > 1 else 0
poses = synth.tower_trajectory(1000)
pick = np.linspace(0, 999, 16).astype(int)
frames = [np.ascontiguousarray(synth.tof_frame(pose=tuple(poses[k]), seed=1000 + int(k)), dtype=np.float32) for k in pick]
pipe = capi.Pipeline(6, synth.WINDOW, synth.LEAF, synth.OPT_DX, 12, 20, max_raw_points=frames[0].shape[0])
if share:
    pipe.set_sm_share(share)
rows = []
for rep in range(6):
    for fr in frames:
        res = pipe.run(fr)
        t = pipe.k0_stamps()
        if rep:   # first pass warms up
            rows.append([(t[k + 1] - t[k]) / 1e3 if t[k + 1] >= t[k] and t[k] else float("nan") for k in range(9)] + [res.n_input])
a = np.array(rows)
print("share", share, "frames", len(a), "filtered points mean", a[:, -1].mean())
print("stamp deltas us (mean over frames):", [round(float(v), 2) for v in np.nanmean(a[:, :9], axis=0)])
print("stamp deltas us (max):", [round(float(v), 2) for v in np.nanmax(a[:, :9], axis=0)])
print("total us mean", round(float(np.nansum(np.nanmean(a[:, :9], axis=0))), 2))
