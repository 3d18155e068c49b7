"""One bench frame (frame 7: 2848 filtered points, 19 iterations) through a pipeline: 3 frames on all SMs, then 2
frames at 9 SMs per frame (the bench share) -- the command ncu captures (tools/gpu_session5.sh).  argv[1]: bench frame number."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pointcloud_segmentation_b200 import capi, synth

which = int(sys.argv[1]) if len(sys.argv) > 1 else 7
poses = synth.tower_trajectory(1000)
k = int(np.linspace(0, 999, 16).astype(int)[which])
frame = synth.tof_frame(pose=tuple(poses[k]), seed=1000 + k).astype(np.float32)
for share, reps in ((0, 3), (9, 2)):
    pipe = capi.Pipeline(6, synth.WINDOW, synth.LEAF, synth.OPT_DX, 12, 20, max_raw_points=frame.shape[0])
    pipe.set_sm_share(share)
    for _ in range(reps):
        res = pipe.run(frame)
        print(share, res.n_input, res.n_lines, [round(u, 1) for u in pipe.timing_us()], flush=True)
    pipe.close()
