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
for K, share in combos:
    if True:
        pipes = [capi.Pipeline(6, synth.WINDOW, synth.LEAF, synth.OPT_DX, 12, 20, max_raw_points=n_pts) for _ in range(K)]
        for p in pipes:
            p.set_sm_share(share)
        lines = [0] * K
        def work(t, n_rep):
            for r in range(n_rep):
                for k in range(t, 16, K):
                    res = pipes[t].run_ptr(pinned[k].data_ptr(), n_pts, stride)
                    lines[t] += res.n_lines
        def run(n_rep):
            th = [threading.Thread(target=work, args=(t, n_rep)) for t in range(K)]
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for t in th: t.start()
            for t in th: t.join()
            torch.cuda.synchronize()
            return time.perf_counter() - t0
        run(2)
        dt = run(reps)
        print(f"K={K} sm_share={share or sm}: {16 * reps / dt:8.1f} frames/s  ({1e3 * dt / (16 * reps):.3f} ms/frame, lines {sum(lines)})", flush=True)
        for p in pipes:
            p.close()
