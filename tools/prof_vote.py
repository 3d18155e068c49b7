"""One Hough::add + getLine on the C2-like cloud (for ncu captures of the vote kernel)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pointcloud_segmentation_b200 import capi, synth

f = synth.tof_frame()
w = synth.WINDOW
m = np.isfinite(f[:, :3]).all(1)
m &= (f[:,0]>=w[0])&(f[:,0]<=w[1])&(f[:,1]>=w[2])&(f[:,1]<=w[3])&(f[:,2]>=w[4])&(f[:,2]<=w[5])
q = np.floor(f[m,:3] / synth.LEAF).astype(np.int64)
_, first = np.unique(q, axis=0, return_index=True)
pts = f[m,:3][np.sort(first)].astype(np.float64)
mn, mx = pts.min(0), pts.max(0)
x = pts - (mn + mx) / 2.0
capi.init_space(6)
h = capi.Hough(x.min(0), x.max(0), synth.OPT_DX)
h.add(x)
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
    h.add(x[:1]); h.subtract(x[:1])     # dirties every record -> full vote on the next getLine
    print(h.get_line()[0])
