"""GPU parity of the Hough class path (add / subtract / getLine / VotingSpace) against the
CPU oracle, through the C-ABI.  Bit-exact: vote counts, argmax cells, flat indices."""
import os

import numpy as np
import pytest

from pointcloud_segmentation_b200 import synth

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _centred(orc, pts):
    x = np.array(pts, dtype=np.float64)
    x, shift = orc.shift_to_origin(x)
    mn, mx = orc.min_max(x)
    return x, mn, mx


def _clouds():
    rng = np.random.default_rng(7)
    yield "tower2k_g3", synth.steel_tower(2000, seed=3), 3, 0.0
    yield "tower1k_g4_dx", synth.steel_tower(1000, seed=5), 4, 0.11
    yield "random300_g2", rng.standard_normal((300, 3)), 2, 0.0
    # grid-aligned points: projections fall exactly on cell boundaries (exact-division path)
    g = np.stack(np.meshgrid(np.arange(-4, 5), np.arange(-4, 5), np.arange(-2, 3)), -1).reshape(-1, 3) * 0.5
    yield "lattice_g3", g.astype(np.float64), 3, 0.25
    yield "two_points_g1", np.array([[0.0, 0.0, 0.0], [1.0, 2.0, 3.0]]), 1, 0.0


@pytest.mark.parametrize("name,pts,g,dx", list(_clouds()), ids=[c[0] for c in _clouds()])
def test_accumulator_and_argmax_bit_exact(h3d, oracle, name, pts, g, dx):
    x, mn, mx = _centred(oracle, pts)
    ref = oracle.hough(g, mn, mx, dx)
    ref.add(x)
    h3d.init_space(g)
    hg = h3d.Hough(mn, mx, dx)
    assert (hg.num_x, hg.num_b) == (ref.num_x, ref.num_b)
    assert hg.dx == ref.dx and hg.max_x == ref.max_x
    hg.add(x)
    want = ref.votes()
    got = hg.votes()
    assert got.sum() == want.sum()
    assert np.array_equal(got, want)
    v, idx, a, b = ref.get_line()
    gv, ga, gb, gidx = hg.get_line()
    assert (gv, gidx) == (v, idx)
    assert np.array_equal(ga, a) and np.array_equal(gb, b)
    hg.close(); ref.close()


def test_subtract_then_get_line_matches_oracle(h3d, oracle):
    """Iterated subtract/getLine, the reference loop's use of the accumulator (wrap.h:241-243)."""
    pts = synth.steel_tower(3000, seed=11)
    x, mn, mx = _centred(oracle, pts)
    g = 4
    ref = oracle.hough(g, mn, mx, 0.0)
    h3d.init_space(g)
    hg = h3d.Hough(mn, mx, 0.0)
    ref.add(x); hg.add(x)
    alive = np.ones(len(x), dtype=bool)
    for it in range(6):
        v, idx, a, b = ref.get_line()
        gv, ga, gb, gidx = hg.get_line()
        assert (gv, gidx) == (v, idx), f"iteration {it}"
        inl = oracle.points_close_to_line(x[alive], a, b, ref.dx)
        sel = np.flatnonzero(alive)[inl]
        y = x[sel]
        ref.subtract(y); hg.subtract(y)
        alive[sel] = False
    assert np.array_equal(hg.votes(), ref.votes())
    hg.close(); ref.close()


def test_subtract_of_never_added_point_wraps_like_reference(h3d, oracle):
    pts = synth.steel_tower(200, seed=2)
    x, mn, mx = _centred(oracle, pts)
    ref = oracle.hough(2, mn, mx, 0.0)
    h3d.init_space(2)
    hg = h3d.Hough(mn, mx, 0.0)
    ref.add(x); hg.add(x)
    ghost = np.array([[0.01, 0.02, 0.03]])
    ref.subtract(ghost); hg.subtract(ghost)
    v, idx, a, b = ref.get_line()
    gv, ga, gb, gidx = hg.get_line()
    assert (gv, gidx) == (v, idx)
    assert np.array_equal(hg.votes(), ref.votes())


def test_sharded_keys_combine_to_global_winner(h3d, oracle):
    pts = synth.steel_tower(1500, seed=9)
    x, mn, mx = _centred(oracle, pts)
    g = 4
    ref = oracle.hough(g, mn, mx, 0.0)
    ref.add(x)
    v, idx, a, b = ref.get_line()
    h3d.init_space(g)
    nb = ref.num_b
    keys = []
    for r in range(3):
        sh = h3d.Hough(mn, mx, 0.0, shard=(nb * r // 3, nb * (r + 1) // 3))
        sh.add(x)
        keys.append(sh.get_line_key())
        sh.close()
    k = max(keys)
    assert (k >> 32, 0xFFFFFFFF - (k & 0xFFFFFFFF)) == (v, idx)
