"""A few frames of the bench workload through the pipeline (for ncu launch lists / captures).
argv: [frames] [fused 0/1] [segments 0/1]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pointcloud_segmentation_b200 import capi, synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 3
if len(sys.argv) > 2:
    capi.lib().h3d_set_fused(int(sys.argv[2]))
frame = synth.tof_frame()
pipe = capi.Pipeline(6, synth.WINDOW, synth.LEAF, synth.OPT_DX, 12, 20, max_raw_points=frame.shape[0])
if len(sys.argv) > 3 and int(sys.argv[3]):
    pipe.enable_segments(capi.SegmentConfig.make(np.sqrt(3) * float(synth.LEAF), 0.995, 1.5, [0.05]))
for _ in range(n):
    res = pipe.run(frame)
    print(res.n_input, res.n_lines, [round(u, 1) for u in pipe.timing_us()])
if len(sys.argv) > 3 and int(sys.argv[3]):
    rc, segs = pipe.segments()
    print("segments:", rc, [(s.iter, s.status, s.n_points) for s in segs])
