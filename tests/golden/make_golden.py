"""Generate the committed golden vectors from the UNMODIFIED reference library.

Run in the build container (where /root/reference exists):  python tests/golden/make_golden.py
It compiles the reference in place into oracle/_ref (oracle/Makefile `ref`) and records, for the
reference's own fixtures (lib/testdata.dat, lib/data/*.dat) and a few synthetic clouds:
  * sphere tables g0..g4 (vertices) and a sha256 of g5/g6,
  * Hough ctor parameters, sum of votes, sha256 of the dense accumulator, first argmax,
  * the whole loop trace (votes, argmax, inlier index lists, fitted lines).
The inputs themselves are stored too, so nothing under /root/reference is read by the tests.
orthogonal_LSQ inside the loop is the oracle's restatement (Eigen is not vendored): the
traces pin votes/argmax/inliers against the reference and (a, b) against the canonical pipeline.
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402
from pointcloud_segmentation_b200 import synth  # noqa: E402

LIB = "/root/reference/libraries/hough-3d-lines"


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def pack_trace(tr):
    out = {"rc": tr.rc, "n_lines": tr.n_lines, "n_input": tr.n_input, "shift": tr.shift, "d": tr.d, "dx": tr.dx,
           "max_x": tr.max_x, "num_x": tr.num_x, "num_b": tr.num_b, "n_iters": len(tr.iters)}
    for k, it in enumerate(tr.iters):
        for f in ("votes", "status", "argmax", "a_raw", "b_raw", "a1", "b1", "rc1", "a2", "b2", "rc2",
                  "inliers1", "inliers2"):
            out[f"it{k}_{f}"] = getattr(it, f)
    return out


def main():
    O.build(ref=True)
    ref = O.Reference()
    gold = {}
    for g in range(0, 5):
        gold[f"sphere_g{g}"] = ref.sphere(g)
    for g in (5, 6):
        gold[f"sphere_g{g}_sha256"] = np.array(sha(ref.sphere(g)))
    cases = {}
    for name, dx in (("testdata.dat", 0.4), ("data/synthetic_a.dat", 0.0), ("data/synthetic_c.dat", 0.0),
                     ("data/attpc_b.dat", 0.0)):
        cases[os.path.basename(name).replace(".dat", "")] = (O.read_dat(os.path.join(LIB, name)), 4, dx, 10, 6)
    cases["tower800"] = (synth.steel_tower(800, seed=31), 3, 0.0, 8, 5)
    for key, (pts, g, dx, minvotes, nlines) in cases.items():
        pts = np.ascontiguousarray(pts, dtype=np.float64)
        gold[f"{key}_points"] = pts
        gold[f"{key}_params"] = np.array([g, dx, minvotes, nlines], dtype=np.float64)
        x, shift = ref.shift_to_origin(pts)
        mn, mx = ref.min_max(x)
        h = ref.hough(g, mn, mx, dx)
        h.add(x)
        votes = h.votes()
        v, idx, a, b = h.get_line()
        gold[f"{key}_hough"] = np.array([h.num_x, h.num_b, h.dx, h.max_x, float(votes.sum()), v, idx], dtype=np.float64)
        gold[f"{key}_votes_sha256"] = np.array(sha(votes))
        gold[f"{key}_line0"] = np.concatenate([a, b])
        h.close()
        tr = ref.hough3dlines(pts, g, dx, minvotes, nlines, mode=1)
        for k2, v2 in pack_trace(tr).items():
            gold[f"{key}_trace_{k2}"] = np.asarray(v2)
    np.savez_compressed(os.path.join(HERE, "reference_golden.npz"), **gold)
    print("wrote", os.path.join(HERE, "reference_golden.npz"), len(gold), "arrays")


if __name__ == "__main__":
    main()
