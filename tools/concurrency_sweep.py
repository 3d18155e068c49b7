"""Frame-parallel replay on one GPU: K host threads x K pipelines, each with sm_count/K SMs for its
persistent loop kernel.  Prints frames/s (wall clock bracketed by device synchronisation)."""
import os, sys, threading, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from pointcloud_segmentation_b200 import capi, synth
import bench

frames = bench.make_frames(16)
pinned = torch.from_numpy(frames).pin_memory()
n_pts, stride = frames.shape[1], frames.shape[2]
sm = capi.device_info()["sm_count"]
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 6
combos = [(1, 0), (4, 37), (5, 37), (6, 37), (8, 37), (5, 29), (6, 49), (4, 49), (3, 49), (8, 24)]
if len(sys.argv) > 2:
    combos = [tuple(int(v) for v in c.split(":")) for c in sys.argv[2].split(",")]
capi.set_wait_mode(2)   # poll + yield, what bench.py uses with frames in flight
for K, share in combos:
    NF = 64
    rp = capi.Replay(K, share, 6, synth.WINDOW, synth.LEAF, synth.OPT_DX, 12, 20, n_pts, NF)
    ptrs = [pinned[k % 16].data_ptr() for k in range(NF)]
    def run(n_rep):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(n_rep):
            res = rp.run(ptrs, [n_pts] * NF, stride, False)
        torch.cuda.synchronize()
        return time.perf_counter() - t0, sum(r.n_lines for r in res[:NF])
    run(2)
    dt, lines = run(reps)
    print(f"K={K} sm_share={share or sm}: {NF * reps / dt:8.1f} frames/s  ({1e3 * dt / (NF * reps):.3f} ms/frame, lines {lines})", flush=True)
    rp.close()
