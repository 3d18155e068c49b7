"""CPU simulation (oracle) of the lazy per-direction records: how many planes must be re-voted per
iteration when K cells are tracked per direction.  Design experiment, not product code."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from oracle import oracle as O
from pointcloud_segmentation_b200 import synth

O.build(ref=False)
orc = O.Oracle()
pose_idx = int(sys.argv[1]) if len(sys.argv) > 1 else -1
if pose_idx < 0:
    frame = synth.tof_frame()
else:
    poses = synth.tower_trajectory(1000)
    frame = synth.tof_frame(pose=tuple(poses[pose_idx]), seed=1000 + pose_idx)
filt = orc.prefilter(frame, synth.WINDOW, float(synth.LEAF)).astype(np.float64)
tr = orc.hough3dlines(filt, 6, synth.OPT_DX, 12, 20, mode=0)
print("points", len(filt), "lines", tr.n_lines, "num_x", tr.num_x)
Xc = filt - tr.shift
mn, mx = Xc.min(0), Xc.max(0)
H = orc.hough(6, mn, mx, tr.dx)
assert H.num_x == tr.num_x
t0 = time.time()
H.add(Xc, threads=8)
nb, pc = H.num_b, H.num_x * H.num_x
def planes():
    return H.votes().reshape(pc, nb).T.astype(np.uint32)   # [dir][cell]
V = planes()
print("vote", time.time() - t0)
cellkey = (65535 - np.arange(pc, dtype=np.uint32))
def topk(Vrows, K):
    keys = (Vrows << 16) | cellkey[None, :]
    part = np.partition(keys, pc - K, axis=1)[:, pc - K:]
    part.sort(axis=1)
    return part[:, ::-1].copy()        # descending keys
Ks = [1]
CHUNKS = [0, 148, 592, 2368]   # 0 = all at once (current scheme); else lazily by descending bound, chunk by chunk
state = {}
tk0 = topk(V, 1)
for K in CHUNKS:
    state[K] = dict(cells=65535 - (tk0 & 0xFFFF), B=tk0[:, 0].copy(), rebuilt=[])
Ks = CHUNKS
alive = np.ones(len(Xc), bool)
for it, rec in enumerate(tr.iters):
    if rec.status != 0: break
    Y = Xc[rec.inliers2]
    H.subtract(Y, threads=8)
    V = planes()
    gmax = V.max()
    for K in Ks:
        s = state[K]
        cur = np.take_along_axis(V, s["cells"].astype(np.int64), axis=1)
        keys = (cur << 16) | (65535 - s["cells"]).astype(np.uint32)
        M = keys.max(axis=1)
        exact = M >= s["B"]
        lb = (M[exact] >> 16).max() if exact.any() else 0
        bound = np.where(exact, M, s["B"]) >> 16
        cand = np.nonzero((~exact) & (bound >= lb))[0]
        n = 0
        rounds = 0
        if K == 0:
            n = len(cand); rounds = 1 if n else 0
            if n:
                tk = topk(V[cand], 1)
                s["cells"][cand] = 65535 - (tk & 0xFFFF); s["B"][cand] = tk[:, 0]
        else:
            order = cand[np.argsort(-bound[cand].astype(np.int64), kind="stable")]
            while len(order):
                order = order[bound[order] >= lb]
                if not len(order): break
                chunk, order = order[:K], order[K:]
                tk = topk(V[chunk], 1)
                s["cells"][chunk] = 65535 - (tk & 0xFFFF); s["B"][chunk] = tk[:, 0]
                lb = max(lb, int((tk[:, 0] >> 16).max()))
                n += len(chunk); rounds += 1
        s["rebuilt"].append((n, rounds))
    print(f"it {it:2d} |Y| {len(Y):4d} max {gmax:3d} " + "  ".join(f"K{K}: {state[K]['rebuilt'][-1][0]:5d}/{state[K]['rebuilt'][-1][1]:5d}" for K in Ks), flush=True)
for K in Ks:
    print("K", K, "total rebuilt", sum(r[0] for r in state[K]["rebuilt"]))
