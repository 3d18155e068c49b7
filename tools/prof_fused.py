"""Phase breakdown of the fused loop kernel (globaltimer stamps, h3d_pipeline_profile).
argv: [sm_share] [bench frame 0..15 | -1] [cold]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pointcloud_segmentation_b200 import capi, synth

share = int(sys.argv[1]) if len(sys.argv) > 1 else 0
which = int(sys.argv[2]) if len(sys.argv) > 2 else -1        # bench frame 0..15 (bench.make_frames), -1: synth default
if which >= 0:
    poses = synth.tower_trajectory(1000)
    k = int(np.linspace(0, 999, 16).astype(int)[which])
    frame = synth.tof_frame(pose=tuple(poses[k]), seed=1000 + k).astype(np.float32)
else:
    frame = synth.tof_frame()
pipe = capi.Pipeline(6, synth.WINDOW, synth.LEAF, synth.OPT_DX, 12, 20, max_raw_points=frame.shape[0])
pipe.set_sm_share(share)
pipe.profile_enable(24)
cold = len(sys.argv) > 3 and sys.argv[3] == "cold"     # flush L2 (256 MiB write) before the measured frame, as bench.py does
for _ in range(3):
    if cold:
        import torch
        torch.empty(256 << 20, dtype=torch.uint8, device="cuda").fill_(1)
        torch.cuda.synchronize()
    res = pipe.run(frame)
print(res.n_input, res.n_lines, [round(u, 1) for u in pipe.timing_us()])
t = pipe.profile_read().astype(np.int64)
# slots: 0 start, 1 published, 2 after barrier A, 3 after inliers 1, 4 after lsq 1, 5 after inliers 2,
#        6 after lsq-2 sums, 7 after eig || un-vote, 10 after zero/re-pack, 11 after barrier B, 8 nlist+1, 9 end
names = ["publish", "barA", "inl1", "sums1", "eig1+inl2", "sums2", "eig2|unvote"]
for it in range(t.shape[0]):
    a = t[it]
    if a[0, 0] == 0:
        break
    c0 = a[0]
    seg = [c0[k + 1] - c0[k] for k in range(7)]
    line = f"it {it:2d} cta0 us: " + " ".join(f"{n}={v/1e3:5.1f}" for n, v in zip(names, seg))
    uv2 = a[:, 14][a[:, 14] > 0] - 1   # records per CTA that the un-vote handed to its warp-per-record stage
    line += f" [unvote warp-stage records/CTA max {uv2.max() if len(uv2) else 0} mean {uv2.mean() if len(uv2) else 0:.1f}; eig1+mask2 {(c0[13]-c0[4])/1e3:.1f}]"
    if a[0, 9]:
        nl = a[:, 8] - 1
        rv = (a[:, 9] - a[:, 11]) / 1e3
        line += (f" remove={(c0[10]-c0[7])/1e3:5.1f} barB={(c0[11]-c0[10])/1e3:5.1f} revote={(c0[9]-c0[11])/1e3:5.1f}"
                 f" | nlist max {nl.max()} mean {nl.mean():.1f} sum {nl.sum()} revote max {rv.max():.1f};"
                 f" span {(a[:,9].max()-a[:,0].min())/1e3:.1f}")
    print(line)
k1 = (t[0, :, 0][t[0, :, 0] > 0].max() - t[0, :, 15][t[0, :, 15] > 0].min()) / 1e3
print("first vote + prologue us", k1)
print("loop total us", (t[t > 1e12].max() - t[0, :, 0].min()) / 1e3)
