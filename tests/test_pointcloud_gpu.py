"""GPU parity of the PointCloud / orthogonal_LSQ entry points (host arrays through the C-ABI)."""
import numpy as np
import pytest

from pointcloud_segmentation_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def cloud():
    return synth.steel_tower(5000, seed=21)


def test_min_max_shift_mean(h3d, oracle, cloud):
    mn, mx = h3d.min_max(cloud)
    omn, omx = oracle.min_max(cloud)
    assert np.array_equal(mn, omn) and np.array_equal(mx, omx)
    got, sh = h3d.shift_to_origin(cloud, shift=(0.5, -1.0, 2.0))
    want, osh = oracle.shift_to_origin(cloud, shift=(0.5, -1.0, 2.0))
    assert np.array_equal(got, want) and np.array_equal(sh, osh)
    assert np.array_equal(h3d.mean_value(cloud), oracle.mean_value(cloud))
    assert np.array_equal(h3d.mean_value(np.zeros((0, 3))), np.zeros(3))
    assert np.array_equal(h3d.mean_value(cloud[:1]), cloud[0])


@pytest.mark.parametrize("n", [0, 1, 2, 31, 32, 33, 500, 5000])
def test_orthogonal_lsq_bit_exact(h3d, oracle, cloud, n):
    rc, a, b = h3d.orthogonal_lsq(cloud[:n])
    orc, oa, ob = oracle.orthogonal_lsq(cloud[:n])
    assert (rc == orc) or (np.isnan(rc) and np.isnan(orc))
    assert np.array_equal(a, oa)
    assert np.array_equal(b, ob, equal_nan=True)


def test_orthogonal_lsq_degenerate(h3d, oracle):
    same = np.tile(np.array([[1.0, 2.0, 3.0]]), (10, 1))
    rc, a, b = h3d.orthogonal_lsq(same)
    orc, oa, ob = oracle.orthogonal_lsq(same)
    assert rc == orc == 0.0 and np.array_equal(a, oa) and np.array_equal(b, ob)
    line = np.outer(np.linspace(-1, 1, 50), [1.0, 1.0, 0.0])
    rc, a, b = h3d.orthogonal_lsq(line)
    orc, oa, ob = oracle.orthogonal_lsq(line)
    assert rc == orc and np.array_equal(a, oa) and np.array_equal(b, ob)


def test_points_close_to_line_and_remove(h3d, oracle, cloud):
    x, _ = oracle.shift_to_origin(cloud)
    a = np.array([0.5, 0.5, 0.0]); b = np.array([0.0, 0.0, 1.0])
    for dx in (0.0, 0.03, 0.1, 10.0):
        got = h3d.points_close_to_line(x, a, b, dx)
        want = oracle.points_close_to_line(x, a, b, dx)
        assert np.array_equal(got, want)
    idx = oracle.points_close_to_line(x, a, b, 0.1)
    y = x[idx]
    assert np.array_equal(h3d.remove_points(x, y), oracle.remove_points(x, y))
    # duplicates are matched one for one, unmatched y stalls the matching like the reference
    xd = np.vstack([x[:100], x[:100]])
    yd = np.vstack([x[10:20], x[10:12]])
    assert np.array_equal(h3d.remove_points(xd, yd), oracle.remove_points(xd, yd))
    ghost = np.vstack([x[5:6], [[9.0, 9.0, 9.0]], x[50:51]])
    assert np.array_equal(h3d.remove_points(x[:200], ghost), oracle.remove_points(x[:200], ghost))
    assert np.array_equal(h3d.remove_points(x[:50], np.zeros((0, 3))), x[:50])
