"""The TMA-staged streaming vote (vote_stream_kernel: cp.async.bulk point tiles + mbarrier, up to 8 planes per
work item) against the oracle's dense accumulator and against its two alternatives (same loop with LDGs, the
round-1 kernel): clouds that do not fit the register-resident form, bounded and unbounded boxes, odd point counts
(the 16-byte granule of the bulk copy), partial last tiles, and a subtract / getLine sequence on the records."""
import numpy as np
import pytest

from pointcloud_segmentation_b200 import synth

pytestmark = pytest.mark.gpu


def _centred(orc, pts):
    x, _ = orc.shift_to_origin(np.array(pts, dtype=np.float64))
    mn, mx = orc.min_max(x)
    return x, mn, mx


@pytest.mark.parametrize("n,g,dx,bounded", [(7001, 3, 0.0, True), (20480, 2, 0.3, True), (9999, 3, 0.0, False),
                                            (6145, 4, 0.0, True), (12288, 1, 0.0, False)])
def test_dense_accumulator_equals_oracle_in_every_mode(h3d, oracle, n, g, dx, bounded):
    pts = synth.steel_tower(n, seed=n % 97)
    x, mn, mx = _centred(oracle, pts)
    if not bounded:                      # a box that is not symmetric about the origin: the range-tested vote form
        mx = mx * 1.25 + 0.1
    ref = oracle.hough(g, mn, mx, dx)
    ref.add(x, threads=8)
    want = ref.votes()
    v, idx, a, b = ref.get_line(threads=8)
    h3d.init_space(g)
    for mode in (1, 0, 2):
        old = h3d.set_stream_mode(mode)
        try:
            hg = h3d.Hough(mn, mx, dx)
            hg.add(x)
            gv, ga, gb, gidx = hg.get_line()          # records from the streaming vote (no dense pass)
            assert (gv, gidx) == (v, idx), mode
            assert np.array_equal(ga, a) and np.array_equal(gb, b)
            got = hg.votes()                          # dense write-back through the same kernel
            assert np.array_equal(got, want), mode
            hg.close()
        finally:
            h3d.set_stream_mode(old)
    ref.close()


@pytest.mark.parametrize("n", [9000, 40000])
def test_subtract_and_revote_lists_through_the_streaming_kernel(h3d, oracle, n):
    """n = 40000: the re-vote lists are short and the cloud is large, so the kernel slices the POINTS over the CTAs
    and merges partial planes in global memory (stream_merge_kernel)"""
    pts = synth.steel_tower(n, seed=5)
    x, mn, mx = _centred(oracle, pts)
    g = 3
    ref = oracle.hough(g, mn, mx, 0.0)
    ref.add(x, threads=8)
    h3d.init_space(g)
    hg = h3d.Hough(mn, mx, 0.0)
    hg.add(x)
    cur = x
    for it in range(4):
        v, idx, a, b = ref.get_line(threads=8)
        gv, ga, gb, gidx = hg.get_line()
        assert (gv, gidx) == (v, idx), it
        inl = oracle.points_close_to_line(cur, a, b, hg.dx)
        y = cur[inl]
        ref.subtract(y, threads=8)
        hg.subtract(y)
        keep = np.ones(len(cur), bool); keep[inl] = False
        cur = cur[keep]
    hg.close(); ref.close()
