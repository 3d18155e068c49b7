"""Fused-kernel time of the bench frames for one pipeline alone, as a function of its SM share."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from pointcloud_segmentation_b200 import capi, synth
import bench
frames = bench.make_frames(16)
dev = torch.from_numpy(frames).cuda()
n_pts, stride = frames.shape[1], frames.shape[2]
for share in (0, 74, 49, 37, 24, 18, 12):
    p = capi.Pipeline(6, synth.WINDOW, synth.LEAF, synth.OPT_DX, 12, 20, max_raw_points=n_pts)
    p.set_sm_share(share)
    tot = np.zeros(4)
    for rep in range(3):
        for k in range(16):
            p.run_dev(dev[k].data_ptr(), n_pts, stride)
            if rep: tot += np.array(p.timing_us())
    tot /= 32
    s = share or 148
    print(f"share {s:3d}: fused kernel {tot[1]:8.1f} us  (SM-time {tot[1]*s/1e3:7.1f} k SM-us)  prefilter {tot[0]:6.1f}  copies {tot[2]:5.1f}  total {tot[3]:8.1f}", flush=True)
    p.close()
