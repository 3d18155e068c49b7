"""Which of the 16 bench frames take the frame fast path at a given SM share (the loop kernel must hold the record
tables of the CTA's directions, two planes and all points in shared memory), and what a frame costs there.
argv: [sm_share ...]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pointcloud_segmentation_b200 import capi, synth

shares = [int(a) for a in sys.argv[1:]] or [0, 18, 9]
poses = synth.tower_trajectory(1000)
pick = np.linspace(0, 999, 16).astype(int)
frames = [np.ascontiguousarray(synth.tof_frame(pose=tuple(poses[k]), seed=1000 + int(k)), dtype=np.float32) for k in pick]
for share in shares:
    pipe = capi.Pipeline(6, synth.WINDOW, synth.LEAF, synth.OPT_DX, 12, 20, max_raw_points=frames[0].shape[0])
    pipe.set_sm_share(share)
    rows = []
    for k, fr in enumerate(frames):
        for _ in range(3):
            res = pipe.run(fr)
        try:
            pipe.k0_stamps(); fast = True
        except capi.H3dError:
            fast = False
        rows.append((k, res.n_input, res.n_lines, fast, round(pipe.timing_us()[3])))
    print("share", share, "fast-path frames", sum(r[3] for r in rows), "of", len(rows))
    for r in rows:
        print("  frame %2d points %5d lines %2d fast %-5s frame us %6d" % r)
    pipe.close()
