"""Bench frame 1 (4341 filtered points, 9 slots) through the general path (host-dispatched vote_resident_kernel<9,0,true>)
and through the frame fast path in scan mode (vote_frame_kernel<0>): the command ncu captures to compare the two."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pointcloud_segmentation_b200 import capi, synth

which = int(sys.argv[1]) if len(sys.argv) > 1 else 1
poses = synth.tower_trajectory(1000)
k = int(np.linspace(0, 999, 16).astype(int)[which])
frame = synth.tof_frame(pose=tuple(poses[k]), seed=1000 + k).astype(np.float32)
capi.set_vote_track(0)
for fast in (False, True):
    capi.set_fast_frames(fast)
    pipe = capi.Pipeline(6, synth.WINDOW, synth.LEAF, synth.OPT_DX, 12, 20, max_raw_points=frame.shape[0])
    for _ in range(3):
        res = pipe.run(frame)
        print(fast, res.n_input, res.n_lines, [round(u, 1) for u in pipe.timing_us()], flush=True)
    pipe.close()
