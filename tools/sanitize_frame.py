"""One ToF frame (BASELINE.json configs[1]: g6, opt_nlines 20) through a pipeline limited to 18 SMs -- the
configuration of the frames-in-flight replay -- for compute-sanitizer:

  compute-sanitizer --tool racecheck --racecheck-report all python tools/sanitize_frame.py
  compute-sanitizer --tool memcheck  python tools/sanitize_frame.py
  compute-sanitizer --tool synccheck python tools/sanitize_frame.py

Kernels exercised: the K0 chain, prep/shift, vote_resident_kernel (first vote), fused_loop_kernel (cooperative,
hand-rolled grid barrier, named barriers, warp-0 / warps-1-15 overlap), segment_tests_kernel; with H3D_SAN_GENERAL=1
also the multi-kernel loop (h3d_set_fused(0)).  The result is compared with the same frame on all SMs."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from pointcloud_segmentation_b200 import capi, synth  # noqa: E402


def run(share, frame, g, nlines):
    pipe = capi.Pipeline(g, synth.WINDOW, synth.LEAF, synth.OPT_DX, 12, nlines, max_raw_points=frame.shape[0])
    pipe.set_sm_share(share)
    pipe.enable_segments(capi.SegmentConfig.make(float(np.sqrt(3.0)) * float(synth.LEAF), 0.9, 1.5, [0.05]))
    pipe.run(frame)
    tr = pipe.trace()
    _, segs = pipe.segments()
    out = [(it["votes"], it["argmax"], it["status"], it["inliers2"].tobytes(), it["a2"].tobytes(), it["b2"].tobytes())
           for it in tr.iters], [s.as_tuple() for s in segs], tr.n_lines
    pipe.close()
    return out


def main():
    g = int(os.environ.get("H3D_SAN_G", "6"))
    nlines = int(os.environ.get("H3D_SAN_NLINES", "20"))
    poses = synth.tower_trajectory(1000)
    frame = synth.tof_frame(pose=tuple(poses[466]), seed=1466).astype(np.float32)   # bench frame 7: 2848 points
    capi.set_debug_poison(1)
    if os.environ.get("H3D_SAN_GENERAL") == "1":
        capi.set_fused(False)
    a = run(18, frame, g, nlines)
    b = run(0, frame, g, nlines)
    assert a == b, "18-SM share and all-SM runs differ"
    print(f"sanitize_frame ok: g{g}, {a[2]} lines, {len(a[1])} segment records, share 18 == all SMs")


if __name__ == "__main__":
    main()
