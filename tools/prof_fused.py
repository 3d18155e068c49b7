"""Phase breakdown of the fused loop kernel (globaltimer stamps, h3d_pipeline_profile)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pointcloud_segmentation_b200 import capi, synth

frame = synth.tof_frame()
pipe = capi.Pipeline(6, synth.WINDOW, synth.LEAF, synth.OPT_DX, 12, 20, max_raw_points=frame.shape[0])
pipe.profile_enable(24)
for _ in range(3):
    res = pipe.run(frame)
print(res.n_input, res.n_lines, [round(u, 1) for u in pipe.timing_us()])
t = pipe.profile_read().astype(np.int64)
names = ["publish", "barrier", "line+inl1", "lsq1", "inl2", "lsq2", "unvote", "list", "revote"]
slots = [0, 1, 2, 3, 4, 5, 6, 7, 7, 9]
tr = pipe.trace()
for it in range(t.shape[0]):
    a = t[it]
    if a[0, 0] == 0:
        break
    c0 = a[0]
    seg = [c0[1] - c0[0], c0[2] - c0[1], c0[3] - c0[2], c0[4] - c0[3], c0[5] - c0[4], c0[6] - c0[5]]
    line = f"it {it:2d} cta0 us: " + " ".join(f"{n}={v/1e3:6.1f}" for n, v in zip(names, seg))
    if a[0, 9]:
        nl = a[:, 8] - 1
        line += f" unvote={(c0[7]-c0[6])/1e3:6.1f} revote={(c0[9]-c0[7])/1e3:6.1f}"
        rv = (a[:, 9] - a[:, 7]) / 1e3
        line += f" | nlist max {nl.max()} mean {nl.mean():.1f} sum {nl.sum()} revote max {rv.max():.1f} us; iter span {(a[:,9].max()-a[:,0].min())/1e3:.1f}"
    if it < len(tr.iters):
        line += f" | n1 {tr.iters[it]['n1'] if 'n1' in tr.iters[it] else ''}"
    print(line)
print("first stamp after kernel start: n/a; loop total us", (t[t > 1e12].max() - t[0, :, 0].min()) / 1e3)
