"""The frame fast path (csrc/h3d_frame.cuh: K0 + widening + bounding box + shift + Hough ctor in one cluster
launch, descriptor in device memory, one wait per frame) against the general path and the oracle, on the cases
where the two take different routes: early returns of hough3dlines() (wrap.h:202-212), frames that do not fit
the register-resident kernels (fall back inside the same call), and sequences that alternate between them."""
import numpy as np
import pytest

from pointcloud_segmentation_b200 import synth
from test_loop_gpu import _compare

pytestmark = pytest.mark.gpu


def _run_both(h3d, frame, g, nlines, opt_dx=synth.OPT_DX, leaf=synth.LEAF, window=synth.WINDOW, segs=None):
    out = []
    for fast in (True, False):
        old = h3d.set_fast_frames(fast)
        try:
            pipe = h3d.Pipeline(g, window, leaf, opt_dx, 12, nlines, max_raw_points=frame.shape[0])
            if segs is not None:
                pipe.enable_segments(segs)
            pipe.run(frame)
            tr = pipe.trace()
            filt, n_f = pipe.filtered(frame.shape[0])
            s = [x.as_tuple() for x in pipe.segments()[1]] if segs is not None else None
            out.append((tr, filt, n_f, s))
            pipe.close()
        finally:
            h3d.set_fast_frames(old)
    return out


def _same_trace(a, b):
    assert (a.rc, a.n_lines, a.n_input, a.n_iters, a.num_x) == (b.rc, b.n_lines, b.n_input, b.n_iters, b.num_x)
    assert (a.d, a.dx, a.max_x) == (b.d, b.dx, b.max_x) and np.array_equal(a.shift, b.shift)
    for x, y in zip(a.iters, b.iters):
        for f in ("votes", "status", "argmax", "rc1", "rc2"):
            assert x[f] == y[f], f
        for f in ("a_raw", "b_raw", "a1", "b1", "a2", "b2", "inliers1", "inliers2"):
            assert np.array_equal(x[f], y[f]), f


def test_fast_path_equals_general_path_and_oracle(h3d, oracle):
    frame = synth.tof_frame(pose=(-0.9, 0.4, 1.4, -0.3), seed=9)
    cfg = h3d.SegmentConfig.make(float(np.sqrt(3.0)) * float(synth.LEAF), 0.9, 1.5, [0.05])
    (fa, ffilt, fn, fseg), (ga, gfilt, gn, gseg) = _run_both(h3d, frame, 6, 20, segs=cfg)
    _same_trace(fa, ga)
    assert fn == gn and np.array_equal(ffilt.view(np.uint32), gfilt.view(np.uint32))
    assert fseg == gseg and len(fseg) > 0
    filt = oracle.prefilter(frame, synth.WINDOW, float(synth.LEAF))
    _compare(fa, oracle.hough3dlines(filt.astype(np.float64), 6, synth.OPT_DX, 12, 20, mode=0))


def test_empty_and_degenerate_frames(h3d):
    n = 4096
    nan = np.full((n, 4), np.nan, dtype=np.float32)                       # nothing in the window
    one = nan.copy(); one[7, :3] = (0.5, 0.1, -0.2)                       # one voxel: d == 0 (wrap.h:202-205)
    tiny = nan.copy(); tiny[:40, :3] = np.array([0.5, 0.1, -0.2]) + np.random.default_rng(0).uniform(-0.02, 0.02, (40, 3))
    for frame in (nan, one, tiny):                                        # tiny: opt_dx >= d (wrap.h:209-212)
        (fa, _, fn, _), (ga, _, gn, _) = _run_both(h3d, frame, 4, 5)
        assert fa.rc == ga.rc == 1 and fn == gn
        assert (fa.n_input, fa.d, fa.n_lines) == (ga.n_input, ga.d, ga.n_lines)


def test_frame_too_dense_for_the_resident_kernels_falls_back(h3d, oracle):
    """> 6144 voxels: FrameDev::status 2, the same call re-runs the frame on the general path"""
    rng = np.random.default_rng(3)
    cloud = np.zeros((30000, 4), dtype=np.float32)
    cloud[:, :3] = rng.uniform([0.0, -1.5, -1.5], [1.5, 1.5, 1.5], (30000, 3))
    (fa, ffilt, fn, _), (ga, gfilt, gn, _) = _run_both(h3d, cloud, 3, 3)
    assert fn == gn > 6144
    _same_trace(fa, ga)
    assert np.array_equal(ffilt.view(np.uint32), gfilt.view(np.uint32))


def test_alternating_fast_and_fallback_frames_on_one_pipeline(h3d):
    """the voxel histogram must be zero again whenever the fast path starts (the general path leaves it dirty)"""
    rng = np.random.default_rng(4)
    dense = np.zeros((38528, 4), dtype=np.float32)
    dense[:30000, :3] = rng.uniform([0.0, -1.5, -1.5], [1.5, 1.5, 1.5], (30000, 3))
    dense[30000:] = np.nan
    tof = [synth.tof_frame(pose=(-1.0 + 0.1 * k, -0.25, 1.0, 0.12), seed=70 + k) for k in range(3)]
    seq = [tof[0], dense, tof[1], tof[1], dense, tof[2], tof[0]]
    want = []
    old = h3d.set_fast_frames(False)
    try:
        pipe = h3d.Pipeline(4, synth.WINDOW, synth.LEAF, synth.OPT_DX, 12, 6, max_raw_points=38528)
        for fr in seq:
            pipe.run(fr)
            want.append(pipe.trace())
        pipe.close()
    finally:
        h3d.set_fast_frames(old)
    pipe = h3d.Pipeline(4, synth.WINDOW, synth.LEAF, synth.OPT_DX, 12, 6, max_raw_points=38528)
    for fr, w in zip(seq, want):
        pipe.run(fr)
        _same_trace(pipe.trace(), w)
    pipe.close()


@pytest.mark.parametrize("leaf,opt_dx", [(0.02, 0.05), (0.08, 0.2), (0.0333333, 0.031)])
def test_other_leaf_sizes_and_cell_sizes(h3d, leaf, opt_dx):
    """other voxel grids (window grid != bounding-box grid) and cell sizes: same voxel order, same Hough geometry"""
    frame = synth.tof_frame(pose=(-1.05, 0.1, 1.2, -0.05), seed=21)
    (fa, ffilt, fn, _), (ga, gfilt, gn, _) = _run_both(h3d, frame, 4, 6, opt_dx=opt_dx, leaf=np.float32(leaf))
    assert fn == gn and np.array_equal(ffilt.view(np.uint32), gfilt.view(np.uint32))
    _same_trace(fa, ga)
