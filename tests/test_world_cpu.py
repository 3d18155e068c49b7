"""World model across frames: segFiltering / checkSimilarity / checkConnections
(node.cpp:479-667) and the CSV dumps (node.cpp:850-897).  Host code in the reference and in the product
(no kernel), so these tests run without a GPU: the product (libh3d.so, C-ABI) against the scalar-Python
restatement in oracle/world.py -- bit-exact -- and the restatement against numpy / known answers."""
import math
import struct

import numpy as np
import pytest

from oracle import world as OW

DIAG = math.sqrt(3) * (0.05 / 1.5)
MIN_W = 0.1


def _bits(v):
    v = float(v)
    return b"nan" if v != v else struct.pack("<d", v)          # the sign / payload of a NaN carries no meaning


def _key_bits(key):
    out = []
    for item in key:
        if isinstance(item, tuple):
            out.append(tuple(_bits(v) for v in item))
        elif isinstance(item, int):
            out.append(item)
        else:
            out.append(_bits(item))
    return tuple(out)


def _to_c(h3d, s):
    return h3d.WorldSegment.make(s.a, s.b, s.t_min, s.t_max, s.radius, s.points_size, s.pca_coeff, s.pca_eigenvalues)


def _rand_seg(rng, radius=None):
    b = rng.normal(size=3)
    b /= np.linalg.norm(b)
    t0 = rng.uniform(-2, 0)
    ev = np.sort(rng.uniform(0.001, 1.0, 3))[::-1]
    return OW.Seg(rng.uniform(-2, 2, 3), b, t0, t0 + rng.uniform(0.3, 3), radius or rng.choice([0.025, 0.05, 0.1]),
                  int(rng.integers(12, 400)), ev[0] / ev.sum(), ev)


def _observe(rng, s, noise=0.01, radius=None):
    """A noisy re-observation of part of segment s (what a later frame sees of the same beam)."""
    lo = s.t_min + rng.uniform(0, 0.4) * (s.t_max - s.t_min)
    hi = s.t_max - rng.uniform(0, 0.4) * (s.t_max - s.t_min)
    b = np.array(s.b) + rng.normal(scale=noise, size=3)
    a = np.array(s.a) + lo * np.array(s.b) + rng.normal(scale=noise, size=3)
    ev = np.sort(rng.uniform(0.001, 1.0, 3))[::-1]
    return OW.Seg(a, b, 0.0, hi - lo, radius or s.radius, int(rng.integers(12, 400)), ev[0] / ev.sum(), ev)


def test_qr_solve_matches_numpy():
    rng = np.random.default_rng(1)
    for _ in range(500):
        m = rng.normal(size=(3, 3)) * 10.0 ** rng.integers(-3, 4)
        rhs = rng.normal(size=3)
        x = OW.qr_solve3([m[:, 0], m[:, 1], m[:, 2]], rhs)
        want = np.linalg.solve(m, rhs)
        assert np.allclose(x, want, rtol=1e-9 * max(1.0, np.linalg.cond(m)), atol=1e-12 * np.abs(want).max())
    # rank-deficient: third column = first; the minimum-pivot solution puts zero on a dependent column
    m = np.array([[1.0, 0.0, 1.0], [0.0, 2.0, 0.0], [3.0, 0.0, 3.0]])
    x = OW.qr_solve3([m[:, 0], m[:, 1], m[:, 2]], m @ np.array([1.0, 1.0, 0.0]))
    assert np.allclose(m @ np.array(x), m @ np.array([1.0, 1.0, 0.0]))


def test_known_answers():
    # two perpendicular beams crossing at (1, 0, 0): t = 1 on the first, t = 2 on the second
    s1 = OW.Seg([0, 0, 0], [1, 0, 0], 0.0, 3.0, 0.05, 100, 0.99, [1, 0.01, 0.001])
    s2 = OW.Seg([1, -2, 0.03], [0, 1, 0], 0.0, 4.0, 0.05, 100, 0.99, [1, 0.01, 0.001])
    ok, sol = OW.check_connections(s1, s2, DIAG)
    assert ok and np.allclose(sol, [1.0, 2.0, 0.03], atol=1e-12) or np.allclose(sol, [1.0, 2.0, -0.03], atol=1e-12)
    # farther apart than 2*diag_voxel + the radii: not connected
    s3 = OW.Seg([1, -2, 0.5], [0, 1, 0], 0.0, 4.0, 0.05, 100, 0.99, [1, 0.01, 0.001])
    assert not OW.check_connections(s1, s3, DIAG)[0]
    # crossing point outside the first segment
    s4 = OW.Seg([5, -2, 0.0], [0, 1, 0], 0.0, 4.0, 0.05, 100, 0.99, [1, 0.01, 0.001])
    assert not OW.check_connections(s1, s4, DIAG)[0]
    # parallel
    assert OW.check_connections(s1, OW.Seg([0, 1, 0], [1, 0.001, 0], 0, 3, 0.05, 9, 0.9, [1, 0, 0]), DIAG) == \
        (False, [0.0, 0.0, 0.0])
    # collinear overlapping halves fuse to the union of their extents
    w = OW.Seg([0, 0, 0], [1, 1, 0], 0.0, 2.0, 0.05, 100, 0.99, [1, 0.01, 0.001])
    d = OW.Seg([1, 1, 0], [1, 1, 0], 0.0, 2.0, 0.05, 50, 0.98, [2, 0.02, 0.002])
    similar, t = OW.check_similarity(d, w, DIAG, MIN_W)
    assert similar
    p_lo = np.array(t.a) + t.t_min * np.array(t.b)
    p_hi = np.array(t.a) + t.t_max * np.array(t.b)
    assert np.allclose(sorted([tuple(p_lo), tuple(p_hi)]), [(0, 0, 0), (3, 3, 0)], atol=1e-12)
    assert t.points_size == 150                                   # world + drone
    assert t.pca_coeff == 0.99 * (1 - MIN_W) + 0.98 * MIN_W       # int / int = 0 -> weight = min_weight
    # different radius class: never similar, target = the drone segment
    d2 = OW.Seg([1, 1, 0], [1, 1, 0], 0.0, 2.0, 0.1, 50, 0.98, [2, 0.02, 0.002])
    similar, t = OW.check_similarity(d2, w, DIAG, MIN_W)
    assert not similar and t.key() == d2.key()
    # disjoint extents on the same axis: not similar
    d3 = OW.Seg([5, 5, 0], [1, 1, 0], 0.0, 2.0, 0.05, 50, 0.98, [2, 0.02, 0.002])
    assert not OW.check_similarity(d3, w, DIAG, MIN_W)[0]


def test_product_check_functions_bit_exact(h3d):
    rng = np.random.default_rng(7)
    cfg = h3d.WorldConfig(DIAG, MIN_W)
    n_sim = n_con = 0
    for k in range(400):
        w = _rand_seg(rng)
        if k % 4 == 0:
            d = _rand_seg(rng)
        else:
            d = _observe(rng, w, noise=0.002 * (k % 7), radius=None if k % 5 else 0.1)
        if k % 50 == 1:                                           # new_b.x == 0 / b.x == 0: NaN / inf t values
            w.a[1] = w.a[2] = 0.0
            w.b = [0.0, 1.0, 0.0]
            d = OW.Seg([w.a[0], 0.5, 0.0], [0.0, 1.0, 0.0], 0.0, 1.0, w.radius, 30, 0.9, [1, 0.1, 0.1])
        if k % 50 == 2:
            d.points_size = w.points_size = 0                     # int 0 / 0
        want_sim, want_t = OW.check_similarity(d, w, DIAG, MIN_W)
        got_sim, got_t = h3d.check_similarity(_to_c(h3d, d), _to_c(h3d, w), cfg)
        assert got_sim == want_sim
        assert _key_bits(got_t.key()) == _key_bits(want_t.key())
        n_sim += want_sim
        init = _rand_seg(rng)                                     # explicit start value of target_seg
        want_sim, want_t = OW.check_similarity(d, w, DIAG, MIN_W, target_init=init)
        got_sim, got_t = h3d.check_similarity(_to_c(h3d, d), _to_c(h3d, w), cfg, _to_c(h3d, init))
        assert got_sim == want_sim and _key_bits(got_t.key()) == _key_bits(want_t.key())
        # connections: random pairs and pairs built to cross
        s1, s2 = _rand_seg(rng), _rand_seg(rng)
        if k % 2:
            x = np.array(_rand_seg(rng).a)
            s1.a = list(x - 0.5 * (s1.t_min + s1.t_max) * np.array(s1.b) + rng.normal(scale=0.02, size=3))
            s2.a = list(x - 0.5 * (s2.t_min + s2.t_max) * np.array(s2.b))
        if k % 25 == 3:
            s2.b = list(np.array(s1.b) * 1.0000001)               # parallel
        if k % 25 == 4:
            s2.b = [0.0, 0.0, 0.0]
        want_ok, want_sol = OW.check_connections(s1, s2, DIAG)
        got_ok, got_sol = h3d.check_connections(_to_c(h3d, s1), _to_c(h3d, s2), cfg)
        assert got_ok == want_ok
        assert [_bits(v) for v in got_sol] == [_bits(v) for v in want_sol]
        n_con += want_ok
    assert n_sim > 50 and n_con > 50


def _same_world(h3d_world, want):
    got = h3d_world.segments()
    assert len(got) == len(want.segments)
    for g, w in zip(got, want.segments):
        assert _key_bits(g.key()) == _key_bits(w.key())
    inter = h3d_world.intersections()
    n = len(want.segments)
    assert inter.shape == (n, n, 2)
    for i in range(n):
        for j in range(n):
            assert (_bits(inter[i, j, 0]), _bits(inter[i, j, 1])) == tuple(_bits(v) for v in want.inter[i][j])


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_world_over_many_frames_bit_exact(h3d, tmp_path, seed):
    """A tower of beams seen again and again with noise: fused segments, new segments, the intersection
    matrix and both CSV files after every frame."""
    rng = np.random.default_rng(100 + seed)
    beams = []
    for k in range(4):                                            # verticals
        beams.append(OW.Seg([k % 2, k // 2, 0], [0.02, 0.01, 1], 0, 3, 0.05, 200, 0.99, [1, 0.01, 0.01]))
    for k in range(4):                                            # horizontals joining them at z = 1 and 2
        z = 1.0 + (k % 2)
        beams.append(OW.Seg([0, k // 2, z], [1, 0.015, 0.01], 0, 1, 0.025 if k == 3 else 0.05, 120, 0.98,
                            [1, 0.02, 0.01]))
    beams.append(OW.Seg([0, 0, 0.5], [1, 1, 0.7], 0, 1.4, 0.05, 90, 0.97, [1, 0.03, 0.01]))   # a diagonal
    want = OW.World(DIAG, MIN_W)
    got = h3d.World(DIAG, MIN_W)
    n_inter = 0
    for frame in range(25):
        seen = [beams[k] for k in rng.permutation(len(beams))[: rng.integers(0, 6)]]
        drone = [_observe(rng, s, noise=0.004) for s in seen]
        if frame % 6 == 5:
            drone.append(_rand_seg(rng))                          # clutter
        want.update(drone)
        got.update([_to_c(h3d, s) for s in drone])
        _same_world(got, want)
        n_inter = sum(1 for i, row in enumerate(want.inter) for j in range(i) if row[j] != (-1.0, -1.0))
    assert len(want.segments) >= len(beams) - 1
    assert n_inter > 0
    seg_csv, int_csv = tmp_path / "segments.csv", tmp_path / "intersections.csv"
    got.save_csv(str(seg_csv), str(int_csv))
    want_seg, want_int = want.csv()
    assert seg_csv.read_text() == want_seg
    assert int_csv.read_text() == want_int
    got.close()


def test_world_quirks(h3d):
    """What segFiltering does as written (node.cpp:486-511): an empty world takes the frame unfiltered and
    looks for no connection; a new segment is recorded under `size after push + i`."""
    cfg = (DIAG, MIN_W)
    s1 = OW.Seg([0, 0, 0], [1, 0, 0], 0.0, 3.0, 0.05, 100, 0.99, [1, 0.01, 0.001])
    s2 = OW.Seg([1, -2, 0.01], [0, 1, 0], 0.0, 4.0, 0.05, 100, 0.99, [1, 0.01, 0.001])
    s3 = OW.Seg([2, -2, -0.01], [0, 1, 0.001], 0.0, 4.0, 0.05, 100, 0.99, [1, 0.01, 0.001])
    for World, conv in ((OW.World, lambda s: s), (h3d.World, lambda s: _to_c(h3d, s))):
        w = World(*cfg)
        w.update([conv(s1), conv(s2)])                            # crossing, but the world was empty
        inter = np.array(w.inter if World is OW.World else w.intersections()).reshape(2, 2, 2)
        assert (inter == -1.0).all()
        w.update([conv(s3)])                                      # new: index 3 + 0 recorded, segment sits at 2
        inter = np.array(w.inter if World is OW.World else w.intersections()).reshape(3, 3, 2)
        assert (inter == -1.0).all()
        w.update([conv(s1)])                                      # s1 seen again: modified -> its pairs are tested
        inter = np.array(w.inter if World is OW.World else w.intersections()).reshape(3, 3, 2)
        assert inter[1, 0, 0] != -1.0 and inter[2, 0, 0] != -1.0 and (inter[2, 1] == -1.0).all()
        w.update([])                                              # an empty frame changes nothing
        assert len(w.segments if World is OW.World else w.segments()) == 3


def test_update_frame_takes_kept_world_segments(h3d):
    """h3d_world_update_frame: only records with status 0 and world_status 1, a/b from a_world/b_world."""
    recs = []
    for k, (status, ws) in enumerate(((0, 1), (1, 0), (0, 2), (0, 1), (2, 0), (0, 0))):
        r = h3d.Segment()
        r.iter, r.status, r.world_status, r.n_points = k, status, ws, 40 + k
        for c in range(3):
            r.a[c], r.b[c] = 100.0, 100.0                         # sensor frame: must not be used
            r.a_world[c], r.b_world[c] = float(k + c), float(c == k % 3)
            r.pca_eigenvalues[c] = 1.0 / (c + 1)
        r.t_min, r.t_max, r.radius, r.pca_coeff = -1.0, 2.0, 0.05, 0.9
        recs.append(r)
    w = h3d.World(DIAG, MIN_W)
    assert w.update_frame(recs) == 2
    segs = w.segments()
    assert [tuple(s.a) for s in segs] == [(0.0, 1.0, 2.0), (3.0, 4.0, 5.0)]
    assert [s.points_size for s in segs] == [40, 43]
    assert tuple(segs[1].b) == (1.0, 0.0, 0.0) and segs[0].radius == 0.05 and segs[0].t_max == 2.0
    assert tuple(segs[0].pca_eigenvalues) == (1.0, 0.5, 1.0 / 3)
    assert w.update_frame([]) == 0
    w.close()


def test_verbosity_levels_and_processing_time_csv(h3d, tmp_path, capfd):
    """node.cpp:23, 296-346 (what each verbosity level prints) and saveProcessingTimeToFile (node.cpp:905-919):
    header, column order, one row per frame with world_segments.size() after the frame"""
    w = h3d.World(DIAG, MIN_W)
    segs = [h3d.WorldSegment.make((0, 0, 1.0 + k), (1, 0, 0), t_min=-1.0, t_max=1.0, radius=0.05, points_size=50,
                                  pca_coeff=0.99, pca_eigenvalues=(1.0, 0.01, 0.01)) for k in range(2)]
    old = h3d.set_verbosity(0)
    try:
        w.update(segs[:1]); w.note_frame(1000, 400, 3)                       # NONE: nothing on stderr
        assert capfd.readouterr().err == ""
        h3d.set_verbosity(1)
        w.update(segs); w.note_frame(2500, 600, 5, hough_rc=1)               # INFO: the callback time only
        err = capfd.readouterr().err
        assert err == "[ INFO] Callback execution time: 600 us\n"
        h3d.set_verbosity(2)
        w.note_frame(4000, 700, 0, hough_rc=1)                               # WARN: warning, matrix, segments, time
        err = capfd.readouterr().err.splitlines()
        assert err[0] == "[ WARN] Unable to perform the Hough transform"
        assert err[1] == "[ INFO] Intersection matrix:"
        n = len(w.segments())
        assert f"[ INFO] Number of segments: {n}" in err
        assert sum(l.startswith("[ INFO] Segment ") for l in err) == n
        assert sum(l.startswith("[ INFO] intersection_matrix[") for l in err) == n * (n - 1) // 2
        assert err[-1] == "[ INFO] Callback execution time: 700 us"
    finally:
        h3d.set_verbosity(old)
    path = tmp_path / "processing_time.csv"
    w.save_processing_time(str(path))
    rows = path.read_text().splitlines()
    assert rows[0] == "wall_time,processing_time,seg_vec_size,nblines"
    assert rows[1] == "1000,400,1,3"
    assert rows[2].split(",")[:2] == ["2500", "600"] and rows[2].split(",")[3] == "5"
    assert rows[3] == f"4000,700,{len(w.segments())},0" and len(rows) == 4
    w.close()
