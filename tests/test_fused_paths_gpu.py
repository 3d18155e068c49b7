"""GPU parity of the mechanisms inside the fused loop kernel that the generic cases do not force:
the sequential fallback of the exact parallel anchor, votes at the very edge of the box (slack
cells of the BOUNDED vote form), SM shares, and the batched replay (h3d_replay_run)."""
import numpy as np
import pytest

from pointcloud_segmentation_b200 import synth
from test_loop_gpu import _compare

pytestmark = pytest.mark.gpu


def _lines_cloud(seed, n=900):
    rng = np.random.default_rng(seed)
    pts = []
    for _ in range(5):
        a, b = rng.uniform(-1, 1, 3), rng.standard_normal(3)
        b /= np.linalg.norm(b)
        t = rng.uniform(-1.5, 1.5, n // 5)
        pts.append(a + t[:, None] * b + rng.normal(0, 0.01, (n // 5, 3)))
    return np.concatenate(pts)


def test_anchor_falls_back_to_the_sequential_sum(h3d, oracle):
    """coordinates whose lowest set bits span > 53 binary digits: the int64 route is not exact,
    one lane must run the reference's sequential double sum"""
    pts = _lines_cloud(11)
    pts[::7] += 1e-13                        # fills the low mantissa bits
    pts[3::11, 1] *= (1 + 2.0 ** -50)
    pts[5] = [1e-300, 2e-300, -1e-300]       # tiny exponents (frame-level fixed point not sane)
    want = oracle.hough3dlines(pts, 4, 0.0, 10, 5, mode=1)
    got = h3d.hough3dlines(pts, 4, 0.0, 10, 5, mode=1)
    _compare(got, want)


def test_float32_derived_cloud_uses_the_exact_integer_anchor(h3d, oracle):
    """f32-valued coordinates (the ToF path): every partial sum is exact, any order gives the
    reference's sequential sum"""
    pts = _lines_cloud(12).astype(np.float32).astype(np.float64)
    want = oracle.hough3dlines(pts, 4, 0.0, 10, 5, mode=1)
    got = h3d.hough3dlines(pts, 4, 0.0, 10, 5, mode=1)
    _compare(got, want)


@pytest.mark.parametrize("dx", [0.0, 0.31, 0.5])
def test_points_on_the_box_corners(h3d, oracle, dx):
    """the bounding-box corners project to |x'| = max_x for some directions: x_i or y_i reach
    num_x (dropped by the flat bound or aliased into the next row, lib/hough.cpp:90-92)"""
    pts = _lines_cloud(13, 600)
    lo, hi = pts.min(0), pts.max(0)
    corners = np.array([[x, y, z] for x in (lo[0], hi[0]) for y in (lo[1], hi[1]) for z in (lo[2], hi[2])])
    pts = np.concatenate([pts, corners, corners])
    want = oracle.hough3dlines(pts, 3, dx, 8, 4, mode=1)
    got = h3d.hough3dlines(pts, 3, dx, 8, 4, mode=1)
    _compare(got, want)


def test_lattice_aligned_cloud(h3d, oracle):
    """every projection lands exactly on a cell boundary for axis-aligned directions: the exact
    path decides every such vote"""
    g = np.arange(-4, 5) * 0.25
    pts = np.array([[x, y, z] for x in g for y in g[::2] for z in g[::3]], dtype=np.float64)
    want = oracle.hough3dlines(pts, 2, 0.25, 3, 4, mode=0)
    got = h3d.hough3dlines(pts, 2, 0.25, 3, 4, mode=0)
    _compare(got, want)


def _trace_tuple(tr):
    return [(it["votes"], it["argmax"], it["status"], it["inliers2"].tobytes(), it["a2"].tobytes(), it["b2"].tobytes())
            for it in tr.iters]


@pytest.mark.parametrize("share", [0, 74, 18, 5])
def test_sm_share_does_not_change_the_result(h3d, oracle, share):
    frame = synth.tof_frame(pose=(-0.9, 0.4, 1.4, -0.3), seed=9)
    filt = oracle.prefilter(frame, synth.WINDOW, float(synth.LEAF))
    want = oracle.hough3dlines(filt.astype(np.float64), 5, synth.OPT_DX, 12, 8, mode=0)
    pipe = h3d.Pipeline(5, synth.WINDOW, synth.LEAF, synth.OPT_DX, 12, 8, max_raw_points=frame.shape[0])
    pipe.set_sm_share(share)
    pipe.run(frame)
    _compare(pipe.trace(), want)
    pipe.close()


def test_replay_matches_one_frame_at_a_time(h3d):
    frames = [synth.tof_frame(pose=(-1.0 + 0.05 * k, -0.25 + 0.03 * k, 1.0, 0.12), seed=40 + k) for k in range(7)]
    frames = [np.ascontiguousarray(f, dtype=np.float32) for f in frames]
    n_max = max(f.shape[0] for f in frames)
    single = h3d.Pipeline(5, synth.WINDOW, synth.LEAF, synth.OPT_DX, 12, 6, max_raw_points=n_max)
    want = []
    for f in frames:
        single.run(f)
        want.append(_trace_tuple(single.trace()))
    single.close()
    rp = h3d.Replay(3, 40, 5, synth.WINDOW, synth.LEAF, synth.OPT_DX, 12, 6, n_max, len(frames))
    try:
        for mode in (0, 1, 2):    # spin, blocking event, poll + yield (h3d_set_wait_mode): same results
            h3d.set_wait_mode(mode)
            res = rp.run([f.ctypes.data for f in frames], [f.shape[0] for f in frames], frames[0].shape[1],
                         on_device=False)
            for k in range(len(frames)):
                its = rp.iters[k * rp.max_iters:(k + 1) * rp.max_iters]
                pool = rp.pool[k * rp.pool_cap:(k + 1) * rp.pool_cap]
                assert _trace_tuple(h3d.Trace(res[k], its, pool)) == want[k], f"frame {k}, wait mode {mode}"
    finally:
        assert h3d.set_wait_mode(0) in (0, 1, 2)
    rp.close()
