"""Measurement: vote-kernel time vs D and ablation on the C1 / C2-like shapes (run on the GPU box)."""
import json, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from pointcloud_segmentation_b200 import capi, synth

def centred(p):
    p = np.array(p, dtype=np.float64)
    mn, mx = p.min(0), p.max(0)
    p = p - (mn + mx) / 2.0
    return p, p.min(0), p.max(0)

out = []
cases = [("C1 10k g4 nx64", synth.steel_tower(10000), 4, 0.0),
         ("C2-like 4.6k g6 dx.0577", None, 6, synth.OPT_DX),
         ("38k g6 nx64", synth.beam_cloud(synth.tower_beams(), 38000, seed=3), 6, 0.0),
         ("200k g5 nx64", synth.beam_cloud(synth.tower_beams(), 200000, seed=3), 5, 0.0)]
for name, pts, g, dx in cases:
    if pts is None:
        f = synth.tof_frame()
        w = synth.WINDOW
        m = np.isfinite(f[:, :3]).all(1)
        m &= (f[:,0]>=w[0])&(f[:,0]<=w[1])&(f[:,1]>=w[2])&(f[:,1]<=w[3])&(f[:,2]>=w[4])&(f[:,2]<=w[5])
        q = np.floor(f[m,:3] / synth.LEAF).astype(np.int64)
        _, first = np.unique(q, axis=0, return_index=True)
        pts = f[m,:3][np.sort(first)].astype(np.float64)
    x, mn, mx = centred(pts)
    capi.init_space(g)
    h = capi.Hough(mn, mx, dx)
    h.add(x)
    votes = len(x) * h.num_b
    for D in (0, 1, 2, 4):
        for ab in (0, 1, 2):
            try:
                us = h.bench_vote(reps=5, ablate=ab, D=D)
            except Exception as e:
                print(name, D, ab, "ERR", e); continue
            rec = {"case": name, "n": len(x), "num_b": h.num_b, "num_x": h.num_x, "D": D, "ablate": ab,
                   "us": round(us, 2), "gvotes_s": round(votes / us / 1e3, 2)}
            print(json.dumps(rec), flush=True)
            out.append(rec)
    h.close()
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/vote_sweep.json", "w"), indent=1)
