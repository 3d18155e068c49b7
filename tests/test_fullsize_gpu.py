"""BASELINE.json configs[3] at FULL size on one GPU: 1 M-point infrastructure cloud, granularity 6 (20 481
directions, 2.05e10 votes per add), dx = d/64, opt_nlines 50.  The oracle's voting would take minutes here,
so votes and argmax are checked through size-independent properties (conservation, the dense accumulator
against the per-direction records, linearity of add / subtract, shards against the whole) and everything
downstream of getLine -- inlier sets, anchors, directions, removal -- against the oracle itself, which is
O(N) per iteration.  Bit-exact throughout."""
import numpy as np
import pytest

from pointcloud_segmentation_b200 import synth

pytestmark = pytest.mark.gpu
G, MINVOTES, NLINES = 6, 10, 50


@pytest.fixture(scope="module")
def cloud():
    return synth.infrastructure()


@pytest.fixture(scope="module")
def trace(h3d, cloud):
    return h3d.hough3dlines(cloud, G, 0.0, MINVOTES, NLINES, mode=1)


def test_c4_loop_downstream_of_getline_matches_oracle(h3d, oracle, cloud, trace):
    n = cloud.shape[0]
    assert trace.rc == 0 and trace.n_input == n and (trace.num_x, trace.num_b) == (64, 20481)
    assert trace.n_lines == NLINES and len(trace.iters) == NLINES
    x, shift = oracle.shift_to_origin(cloud)
    assert np.array_equal(shift, trace.shift)
    alive = np.ones(n, dtype=bool)
    prev_votes = None
    for k, it in enumerate(trace.iters):
        ctx = f"iteration {k}"
        assert it["status"] == 0, ctx
        assert prev_votes is None or it["votes"] <= prev_votes, ctx     # removing points only lowers the maximum
        prev_votes = it["votes"]
        live = np.flatnonzero(alive)
        xl = x[live]
        y1 = live[oracle.points_close_to_line(xl, it["a_raw"], it["b_raw"], trace.dx)]
        assert np.array_equal(y1, it["inliers1"]), ctx
        rc1, a1, b1 = oracle.orthogonal_lsq(x[y1])
        assert np.array_equal(a1, it["a1"]) and np.array_equal(b1, it["b1"]), ctx
        y2 = live[oracle.points_close_to_line(xl, a1, b1, trace.dx)]
        assert np.array_equal(y2, it["inliers2"]), ctx
        assert len(y2) >= MINVOTES, ctx
        rc2, a2, b2 = oracle.orthogonal_lsq(x[y2])
        assert np.array_equal(a2 + shift, it["a2"]) and np.array_equal(b2, it["b2"]), ctx
        alive[y2] = False


def test_c4_votes_conservation_records_linearity_and_shards(h3d, oracle, cloud, trace):
    n = cloud.shape[0]
    x, _ = oracle.shift_to_origin(cloud)
    mn, mx = oracle.min_max(x)
    h3d.init_space(G)
    hg = h3d.Hough(mn, mx, trace.dx)
    assert (hg.num_x, hg.num_b) == (trace.num_x, trace.num_b)
    hg.add(x)
    v0, a0, b0, flat0 = hg.get_line()                                   # from the per-direction records
    it0 = trace.iters[0]
    assert (v0, flat0) == (it0["votes"], it0["argmax"])
    assert np.array_equal(a0, it0["a_raw"]) and np.array_equal(b0, it0["b_raw"])
    dense = hg.votes()                                                  # the reference's VotingSpace, 84 M cells
    assert int(dense.sum(dtype=np.uint64)) == n * hg.num_b              # every vote of a centred cloud lands
    assert int(np.argmax(dense)) == flat0 and int(dense[flat0]) == v0   # first maximum in flat order
    # direction shards against the whole: keys combine to the global winner, slices tile the accumulator
    nb = hg.num_b
    keys = []
    for r in range(3):
        j0, j1 = nb * r // 3, nb * (r + 1) // 3
        sh = h3d.Hough(mn, mx, trace.dx, shard=(j0, j1))
        sh.add(x)
        keys.append(sh.get_line_key())
        sh.close()
    key = max(keys)
    assert (key >> 32, 0xFFFFFFFF - (key & 0xFFFFFFFF)) == (v0, flat0)
    # linearity: subtracting the first line's inliers = the accumulator of the remaining points
    y = x[it0["inliers2"]]
    hg.subtract(y)
    v1, _, _, flat1 = hg.get_line()
    it1 = trace.iters[1]
    assert (v1, flat1) == (it1["votes"], it1["argmax"])
    after = hg.votes()
    hy = h3d.Hough(mn, mx, trace.dx)
    hy.add(y)
    assert np.array_equal(dense - hy.votes(), after)
    hy.close()
    # add -> subtract of everything left gives the empty accumulator back
    alive = np.ones(n, dtype=bool)
    alive[it0["inliers2"]] = False
    hg.subtract(x[alive])
    assert not hg.votes().any()
    assert hg.get_line()[0] == 0
    hg.close()


def test_big_cloud_kernels_match_one_cta_kernels_and_oracle(h3d, oracle):
    """Clouds >= 32 768 points take the chunked multi-CTA compaction / removal kernels: same records as the
    one-CTA kernels of the frame loop, and both equal to the oracle (40 k-point tower, g4; a ragged last chunk)."""
    pts = synth.steel_tower(40000 + 77, seed=8)
    # non-finite points are stripped in order (wrap.h:175-189): in the first and the last chunk, a whole run of
    # them across a chunk border, and scattered ones -- the chunked prep kernels must keep order and indices
    pts[0, 0] = np.nan; pts[1023:1030, 1] = np.inf; pts[-1, 2] = -np.inf; pts[5000:39000:977, 0] = np.nan
    want = oracle.hough3dlines(pts, 4, 0.0, 10, 14, mode=1)
    assert want.n_input == len(pts) - int((~np.isfinite(pts).all(axis=1)).sum()) < len(pts)
    runs = []
    for on in (1, 0):
        old = h3d.set_big_kernels(on)
        runs.append(h3d.hough3dlines(pts, 4, 0.0, 10, 14, mode=1))
        h3d.set_big_kernels(old)
    for got in runs:
        assert (got.rc, got.n_lines, len(got.iters)) == (want.rc, want.n_lines, len(want.iters))
        assert got.n_input == want.n_input and np.array_equal(got.shift, want.shift) and got.d == want.d
        for g, w in zip(got.iters, want.iters):
            assert (g["votes"], g["argmax"], g["status"]) == (w.votes, w.argmax, w.status)
            assert np.array_equal(g["inliers1"], w.inliers1) and np.array_equal(g["inliers2"], w.inliers2)
            for key in ("a_raw", "b_raw", "a1", "b1", "a2", "b2"):
                assert np.array_equal(g[key], getattr(w, key)), key


@pytest.mark.parametrize("g", [0, 1, 2, 3])
def test_sliced_vote_few_directions_many_points(h3d, oracle, g):
    """Granularity 0-3 on a 100 k-point cloud: the full vote slices the points of a direction over CTAs and merges
    partial planes in global memory.  Same records as one work item per direction, the loop equal to the oracle, and
    the records equal to the dense accumulator (built by the streaming form)."""
    pts = synth.beam_cloud(synth.tower_beams(), 100_003, seed=5)
    want = oracle.hough3dlines(pts, g, 0.0, 10, 6, mode=1)
    for on in (1, 0):
        old = h3d.set_sliced_vote(on)
        got = h3d.hough3dlines(pts, g, 0.0, 10, 6, mode=1)
        h3d.set_sliced_vote(old)
        assert (got.rc, got.n_lines, len(got.iters)) == (want.rc, want.n_lines, len(want.iters))
        for a, w in zip(got.iters, want.iters):
            assert (a["votes"], a["argmax"], a["status"]) == (w.votes, w.argmax, w.status)
            assert np.array_equal(a["inliers1"], w.inliers1) and np.array_equal(a["inliers2"], w.inliers2)
            for key in ("a_raw", "b_raw", "a1", "b1", "a2", "b2"):
                assert np.array_equal(a[key], getattr(w, key)), key
    x, _ = oracle.shift_to_origin(pts)
    mn, mx = oracle.min_max(x)
    h3d.init_space(g)
    hg = h3d.Hough(mn, mx, 0.0)
    hg.add(x)                                   # sliced form
    v, _, _, flat = hg.get_line()
    dense = hg.votes()                          # streaming form, materialised
    assert int(dense.sum(dtype=np.uint64)) == len(x) * hg.num_b
    assert (int(np.argmax(dense)), int(dense.max())) == (flat, v)
    hg.add(x[:70000])                           # a second add on top: 170 003 points, all re-voted
    v2, _, _, flat2 = hg.get_line()
    dense2 = hg.votes()
    assert (int(np.argmax(dense2)), int(dense2.max())) == (flat2, v2)
    hg.close()
