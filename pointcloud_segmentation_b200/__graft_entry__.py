"""Driver entry points: build() compiles everything for sm_100a, smoke() runs one small frame
of the hot path on cuda:0 and checks it against the CPU oracle."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def build() -> None:
    """Compile libh3d.so (nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo, in-tree), the
    oracle's C restatement and -- where /root/reference exists -- the reference library in place
    (oracle/_ref).  Building the checker is not using it: the product never links oracle/."""
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "pointcloud_segmentation_b200")], check=True)
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "pointcloud_segmentation_b200"), "dropin"], check=False)
    from oracle import oracle as O
    O.build(ref=None)
    import pointcloud_segmentation_b200  # noqa: F401
    from pointcloud_segmentation_b200 import capi
    capi.lib()


def smoke() -> None:
    """One small invocation of the hot path on cuda:0 (pre-filter + vote + iterative loop),
    bit-compared with the oracle."""
    import numpy as np
    from oracle import oracle as O
    from pointcloud_segmentation_b200 import capi, synth

    O.build(ref=False)
    orc = O.Oracle()
    frame = synth.tof_frame()
    g, nlines = 4, 4
    pipe = capi.Pipeline(g, synth.WINDOW, synth.LEAF, synth.OPT_DX, 12, nlines, max_raw_points=frame.shape[0])
    pipe.run(frame)
    got = pipe.trace()
    filt = orc.prefilter(frame, synth.WINDOW, float(synth.LEAF))
    want = orc.hough3dlines(filt.astype(np.float64), g, synth.OPT_DX, 12, nlines, mode=0)
    assert got.rc == want.rc == 0 and got.n_input == len(filt)
    assert got.n_lines == want.n_lines and len(got.iters) == len(want.iters)
    for a, b in zip(got.iters, want.iters):
        assert (a["votes"], a["argmax"], a["status"]) == (b.votes, b.argmax, b.status)
        assert np.array_equal(a["inliers1"], b.inliers1) and np.array_equal(a["inliers2"], b.inliers2)
        assert np.array_equal(a["a2"], b.a2) and np.array_equal(a["b2"], b.b2)
    pipe.close()
    print(f"smoke ok: {got.n_input} filtered points, {got.n_lines} lines, bit-identical to the oracle")


if __name__ == "__main__":
    build()
    if len(sys.argv) > 1 and sys.argv[1] == "smoke":
        smoke()
