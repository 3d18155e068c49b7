"""PointCloud::meanValue (lib/pointcloud.cpp:29-38) is a SEQUENTIAL double sum; the device evaluates it exactly in
parallel while the running sum stays inside one binade and falls back to the chain otherwise
(mean_axis_warp, csrc/h3d_fit.cuh).  Adversarial inputs for that logic, bit-compared with the oracle's C loop:
sums that hover around zero (a binade crossing every few elements), exact ties of the rounding, powers of two,
huge dynamic range, sizes around the batch (32) and group (256) boundaries, non-finite values."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SIZES = [1, 2, 31, 32, 33, 255, 256, 257, 511, 513, 1000, 4097, 20000]


def _columns(n, seed):
    rng = np.random.default_rng(seed)
    k = np.arange(n)
    cols = {
        "zero-mean normal": rng.normal(0.0, 1.0, n),
        "drifting": rng.normal(0.3, 1.0, n),
        "alternating +-1 + noise": np.where(k % 2 == 0, 1.0, -1.0) + rng.normal(0, 1e-3, n),
        "ties at 2^-43 on a sum near 1024": np.concatenate([[1024.0], (2 * rng.integers(0, 8, max(n - 1, 0)) + 1) * 2.0 ** -43])[:n],
        "ties then crossing": np.concatenate([[2048.0 - 2.0 ** -40], np.full(max(n - 1, 0), 2.0 ** -42)])[:n],
        "powers of two": 2.0 ** rng.integers(-30, 30, n) * rng.choice([-1.0, 1.0], n),
        "huge range": rng.normal(0, 1, n) * 10.0 ** rng.integers(-12, 13, n),
        "cancel big": np.where(k % 3 == 0, 1e16, np.where(k % 3 == 1, 1.0, -1e16)),
        "metres (like a cloud)": rng.uniform(-25.0, 25.0, n),
        "exact halves": rng.integers(-1000, 1000, n) * 0.5,
        "zeros and signed zeros": np.where(k % 2 == 0, 0.0, -0.0),
    }
    return cols


@pytest.mark.parametrize("n", SIZES)
def test_sequential_mean_bit_exact_on_adversarial_columns(h3d, oracle, n):
    cols = _columns(n, seed=n)
    names = list(cols)
    for g in range(0, len(names), 3):          # three columns per call (x, y, z)
        trio = (names[g:g + 3] + names[:3])[:3]
        cloud = np.ascontiguousarray(np.stack([cols[c] for c in trio], axis=1), dtype=np.float64)
        got, want = h3d.mean_value(cloud), oracle.mean_value(cloud)
        assert got.tobytes() == want.tobytes(), (n, trio, got, want)
        rc, a, _ = h3d.orthogonal_lsq(cloud)   # the anchor of orthogonal_LSQ is the same sum
        _, oa, _ = oracle.orthogonal_lsq(cloud)
        assert a.tobytes() == oa.tobytes(), (n, trio)


def test_sequential_mean_with_non_finite_values(h3d, oracle):
    rng = np.random.default_rng(3)
    cloud = rng.normal(0, 1, (3000, 3))
    cloud[1500, 0] = np.inf
    cloud[700, 1] = np.nan
    cloud[2999, 2] = -np.inf
    got, want = h3d.mean_value(cloud), oracle.mean_value(cloud)
    assert np.array_equal(got, want, equal_nan=True) and np.isnan(got[1]) and np.isinf(got[0]) and np.isinf(got[2])
    cloud[10, 0] = -np.inf                     # inf - inf -> NaN from that element on
    got, want = h3d.mean_value(cloud), oracle.mean_value(cloud)
    assert np.array_equal(got, want, equal_nan=True) and np.isnan(got[0])
