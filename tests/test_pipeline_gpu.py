"""GPU parity of K0 (PassThrough + VoxelGrid) and of the persistent pipeline against the oracle."""
import numpy as np
import pytest

from pointcloud_segmentation_b200 import synth
from test_loop_gpu import _compare

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True, params=["fast", "fast_scan", "fused", "multikernel"])
def loop_form(request, h3d):
    """Every test runs on the frame fast path (one cluster launch for K0 + prep + Hough ctor, one wait), with
    the fused persistent loop kernel behind the eleven-launch K0 and the host-side ctor, and with the general
    multi-kernel loop."""
    old = h3d.set_fused(request.param != "multikernel")
    old_fast = h3d.set_fast_frames(request.param.startswith("fast"))
    old_track = h3d.set_vote_track(0 if request.param == "fast_scan" else 1)   # fast_scan: first vote scans its planes (RED form)
    yield request.param
    h3d.set_fused(old)
    h3d.set_fast_frames(old_fast)
    h3d.set_vote_track(old_track)


def _frames():
    yield "tof_default", synth.tof_frame()
    yield "tof_pose2", synth.tof_frame(pose=(-0.9, 0.4, 1.4, -0.3), seed=9)
    yield "tof_stride3", synth.tof_frame(pose=(-1.1, 0.0, 0.8, 0.0), seed=5, stride=3)


@pytest.mark.parametrize("name,frame", list(_frames()), ids=[f[0] for f in _frames()])
def test_prefilter_bit_exact(h3d, oracle, name, frame):
    want = oracle.prefilter(frame, synth.WINDOW, float(synth.LEAF))
    got = h3d.prefilter(frame, synth.WINDOW, float(synth.LEAF))
    assert got.shape == want.shape
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


def test_prefilter_edges(h3d, oracle):
    win = synth.WINDOW
    empty = np.full((100, 4), np.nan, dtype=np.float32)
    assert h3d.prefilter(empty, win, 0.05).shape == (0, 3)
    one = np.array([[0.5, 0.1, -0.2, 1.0]], dtype=np.float32)
    assert np.array_equal(h3d.prefilter(one, win, 0.05), oracle.prefilter(one, win, 0.05))
    # points exactly on the window limits are kept (inclusive), duplicates share a voxel
    edge = np.array([[0.0, -1.5, -1.5], [1.5, 1.5, 1.5], [1.5, 1.5, 1.5], [1.5000001, 0, 0]], dtype=np.float32)
    assert np.array_equal(h3d.prefilter(edge, win, 0.05), oracle.prefilter(edge, win, 0.05))
    rng = np.random.default_rng(0)
    cloud = (rng.random((20000, 3)) * 4 - 2).astype(np.float32)
    want = oracle.prefilter(cloud, win, 0.1)
    got = h3d.prefilter(cloud, win, 0.1)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


@pytest.mark.parametrize("g,nlines", [(4, 6), (6, 20)])
def test_pipeline_frame_matches_oracle(h3d, oracle, g, nlines):
    frame = synth.tof_frame()
    filt = oracle.prefilter(frame, synth.WINDOW, float(synth.LEAF))
    want = oracle.hough3dlines(filt.astype(np.float64), g, synth.OPT_DX, 12, nlines, mode=0)
    pipe = h3d.Pipeline(g, synth.WINDOW, synth.LEAF, synth.OPT_DX, 12, nlines, max_raw_points=frame.shape[0])
    for _ in range(2):   # second run reuses every buffer
        pipe.run(frame)
        got_filt, n_f = pipe.filtered(frame.shape[0])
        assert n_f == filt.shape[0]
        assert np.array_equal(got_filt.view(np.uint32), filt.view(np.uint32))
        _compare(pipe.trace(), want)
    pipe.close()


def test_pipeline_without_prefilter(h3d, oracle):
    cloud = synth.steel_tower(3000, seed=12).astype(np.float32)
    want = oracle.hough3dlines(cloud.astype(np.float64), 4, 0.06, 10, 8, mode=0)
    pipe = h3d.Pipeline(4, synth.WINDOW, synth.LEAF, 0.06, 10, 8, max_raw_points=cloud.shape[0], prefilter=False)
    pipe.run(cloud)
    _compare(pipe.trace(), want)
    pipe.close()
