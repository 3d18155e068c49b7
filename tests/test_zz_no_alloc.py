"""h3d_points_close_to_line_noalloc (include/h3d.h): the allocation-free variant of pointsCloseToLine whose
inlier flags live in a fixed __device__ array.  It has no counterpart in the reference; the checker is the
existing entry point h3d_points_close_to_line (itself bit-compared with the oracle in tests/test_capi_gpu.py),
and the result must be IDENTICAL (integer index lists: tolerance zero, as everywhere in this suite).

Collected last (file name).  Passes alone in a fresh process and after the
other tests (it shares no state with them: its scratch array belongs to this entry point only).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _cloud(n, seed):
    rng = np.random.default_rng(seed)
    t = rng.uniform(-4.0, 4.0, n)
    a = np.array([0.3, -0.2, 1.1])
    b = np.array([1.0, 2.0, -0.5]); b /= np.linalg.norm(b)
    pts = a + t[:, None] * b + rng.normal(0.0, 0.08, (n, 3))     # about half of the points within dx
    pts[::7] += rng.uniform(-3.0, 3.0, (len(pts[::7]), 3))       # and clear outliers
    return np.ascontiguousarray(pts, dtype=np.float64), a, b


def _run_noalloc(h3d, torch, pts, a, b, dx, idx_dev, cnt_dev, stream):
    xyz_dev = torch.from_numpy(pts).cuda()
    torch.cuda.synchronize()
    h3d.points_close_to_line_noalloc(xyz_dev.data_ptr(), pts.shape[0], a, b, dx, idx_dev.data_ptr(),
                                     cnt_dev.data_ptr(), stream)
    torch.cuda.synchronize()
    k = int(cnt_dev.item())
    return idx_dev[:k].cpu().numpy().astype(np.uint32)


def test_noalloc_matches_points_close_to_line_on_two_calls_in_a_row(h3d):
    import torch
    dx = 0.1
    idx_dev = torch.empty(h3d.NOALLOC_MAX_POINTS, dtype=torch.int32, device="cuda")
    cnt_dev = torch.zeros(1, dtype=torch.int32, device="cuda")
    side = torch.cuda.Stream()
    # first a LARGE cloud, then a small one, then a mid-sized one: the shared array keeps the earlier call's
    # flags beyond the later call's n, and they must not leak into its result
    for k, (n, stream) in enumerate([(200_000, 0), (3_001, 0), (65_537, side.cuda_stream), (0, 0), (1, 0)]):
        pts, a, b = _cloud(max(n, 1), seed=100 + k)
        pts = pts[:n]
        want = h3d.points_close_to_line(pts, a, b, dx)
        idx_dev.fill_(-1)
        got = _run_noalloc(h3d, torch, pts, a, b, dx, idx_dev, cnt_dev, stream)
        assert got.dtype == want.dtype and np.array_equal(got, want), (k, n, len(got), len(want))
        if n > 1000:
            assert 0.2 * n < len(want) < 0.8 * n      # the case separates inliers from outliers
        assert int(idx_dev[len(got):].max().item()) == -1      # nothing written past the count


def test_noalloc_allocates_nothing(h3d):
    """the free-memory figure of the device does not move across 50 calls (no cudaMalloc, no growth)"""
    import torch
    pts, a, b = _cloud(50_000, seed=7)
    xyz_dev = torch.from_numpy(pts).cuda()
    idx_dev = torch.empty(len(pts), dtype=torch.int32, device="cuda")
    cnt_dev = torch.zeros(1, dtype=torch.int32, device="cuda")
    h3d.points_close_to_line_noalloc(xyz_dev.data_ptr(), len(pts), a, b, 0.1, idx_dev.data_ptr(), cnt_dev.data_ptr())
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    first = int(cnt_dev.item())
    for _ in range(50):
        h3d.points_close_to_line_noalloc(xyz_dev.data_ptr(), len(pts), a, b, 0.1, idx_dev.data_ptr(), cnt_dev.data_ptr())
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0 and int(cnt_dev.item()) == first == len(h3d.points_close_to_line(pts, a, b, 0.1))


def test_noalloc_rejects_clouds_beyond_the_fixed_scratch(h3d):
    import torch
    cnt_dev = torch.zeros(1, dtype=torch.int32, device="cuda")
    dummy = torch.zeros(8, dtype=torch.float64, device="cuda")
    with pytest.raises(h3d.H3dError) as err:
        h3d.points_close_to_line_noalloc(dummy.data_ptr(), h3d.NOALLOC_MAX_POINTS + 1, (0, 0, 0), (0, 0, 1), 0.1,
                                         dummy.data_ptr(), cnt_dev.data_ptr())
    assert "H3D_NOALLOC_MAX_POINTS" in str(err.value) and err.value.code == 4      # H3D_ERR_RANGE
    with pytest.raises(h3d.H3dError):
        h3d.points_close_to_line_noalloc(0, 10, (0, 0, 0), (0, 0, 1), 0.1, dummy.data_ptr(), cnt_dev.data_ptr())
