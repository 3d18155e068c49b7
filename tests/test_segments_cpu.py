"""CPU checks of the oracle's segment post-tests (wrap.h:261-337 restated in oracle/h3d_oracle.c)
against independent numpy arithmetic, and of the C-ABI structure layouts."""
import ctypes as C

import numpy as np

from pointcloud_segmentation_b200 import capi, synth


def _cfg(mod, radii=(0.05,), min_pca=0.995):
    return mod.SegmentConfig.make(np.sqrt(3) * float(synth.LEAF), min_pca, 1.5, list(radii))


def test_layouts_agree(oracle_mod):
    assert C.sizeof(capi.Segment) == C.sizeof(oracle_mod.Segment) == 200
    assert C.sizeof(capi.SegmentConfig) == C.sizeof(oracle_mod.SegmentConfig) == 96
    for a, b in zip(capi.Segment._fields_, oracle_mod.Segment._fields_):
        assert a[0] == b[0] and getattr(capi.Segment, a[0]).offset == getattr(oracle_mod.Segment, b[0]).offset


def test_oracle_segment_against_numpy(oracle, oracle_mod):
    rng = np.random.default_rng(11)
    shift = np.array([0.4, 0.2, -0.6])
    b = np.array([0.48, 0.6, 0.64])
    a = np.array([0.1, -0.2, 0.3]) + shift
    t = rng.uniform(-1.2, 1.2, 400)
    y = (a - shift) + t[:, None] * b + rng.normal(0, 0.015, (400, 3))
    rc, s = oracle.segment_test(y, shift, a, b, _cfg(oracle_mod, min_pca=0.9))
    assert rc == 0
    p = y + shift
    proj = a + np.outer((p - a) @ b / (b @ b), b)
    tt = (proj[:, 0] - a[0]) / b[0]
    order = np.argsort(tt, kind="stable")
    assert np.isclose(s.t_min, tt.min(), rtol=1e-12) and np.isclose(s.t_max, tt.max(), rtol=1e-12)
    norms = np.linalg.norm(a + tt[order, None] * b, axis=1)
    assert np.isclose(s.max_dist, np.abs(np.diff(norms)).max(), rtol=1e-9)
    rad = np.linalg.norm(proj - p, axis=1)
    assert np.isclose(s.radius_raw, max(rad[0], rad[-1]), rtol=1e-9)
    assert s.radius == 0.05 and s.n_points == 400
    ev = np.sort(np.linalg.eigvalsh(np.cov(p.T)))[::-1]
    assert np.allclose(list(s.pca_eigenvalues), ev, rtol=2e-4, atol=1e-7)      # float PCA vs f64
    assert np.isclose(s.pca_coeff, ev[0] / ev.sum(), rtol=1e-4)
    assert np.isclose(s.length, np.linalg.norm((tt.max() - tt.min()) * b), rtol=1e-9)
    diag = np.sqrt(3) * float(synth.LEAF)
    assert s.min_nb_points == int(2.0 * 0.05 * s.length / (1.5 * 2 * diag * 2 * diag))
    assert s.status == (0 if (s.pca_coeff > 0.9 and 400 > s.min_nb_points) else 2)


def test_oracle_segment_quirks(oracle, oracle_mod):
    shift = np.zeros(3)
    a = np.array([0.0, 0.0, 0.0]); b = np.array([0.6, 0.0, 0.8])
    y = np.linspace(-1, 1, 50)[:, None] * b + np.array([0.0, 0.04, 0.0])
    # only the largest radius size can pass `max_radius <= closest_radius` (wrap.h:296-312)
    rc, s = oracle.segment_test(y, shift, a, b, _cfg(oracle_mod, (0.04, 0.1), 0.5))
    assert rc == 0 and s.radius == 0.04 and s.status == 1
    rc, s = oracle.segment_test(y, shift, a, b, _cfg(oracle_mod, (0.04,), 0.5))
    assert rc == 0 and s.status == 0 and abs(s.radius_raw - 0.04) < 1e-12
    # equal t values keep insertion order: t_min / t_max unaffected, gap 0 between duplicates
    rc, s2 = oracle.segment_test(np.repeat(y, 3, axis=0), shift, a, b, _cfg(oracle_mod, (0.04,), 0.5))
    assert rc == 0 and s2.t_min == s.t_min and s2.t_max == s.t_max and s2.max_dist == s.max_dist
    # b.x == 0: find_t has no solution, hough3dlines returns 1 (wrap.h:43-45, 276-279)
    rc, s3 = oracle.segment_test(y, shift, a, np.array([0.0, 0.6, 0.8]), _cfg(oracle_mod))
    assert rc == 1 and s3.status == 3


def test_oracle_world_transform_against_scipy(oracle, oracle_mod):
    """drone2WorldSeg + heighSegmentCutoff (node.cpp:429-470)."""
    from scipy.spatial.transform import Rotation
    rng = np.random.default_rng(2)
    shift = np.zeros(3)
    a = np.array([0.2, -0.1, 0.4]); b = np.array([0.6, 0.0, 0.8])
    y = a + np.linspace(-1, 1, 80)[:, None] * b + rng.normal(0, 0.01, (80, 3))
    kept = trimmed = 0
    for k in range(40):
        q = rng.normal(size=4); q /= np.linalg.norm(q)
        pos = rng.uniform(-2, 2, 3)
        floor = rng.uniform(-1.5, 1.5)
        rc, s = oracle.segment_test(y, shift, a, b, _cfg(oracle_mod, (0.05,), 0.5))
        assert rc == 0 and s.status == 0
        pose = oracle_mod.Pose.make(q, pos, floor)
        oracle.lib.h3do_segment_to_world(C.byref(pose), C.byref(s))
        R = Rotation.from_quat([q[1], q[2], q[3], q[0]]).as_matrix()          # scipy: x, y, z, w
        assert np.allclose(list(s.a_world), R @ a + pos, atol=1e-12)
        assert np.allclose(list(s.b_world), R @ b, atol=1e-12)
        z1 = s.t_min * s.b_world[2] + s.a_world[2]; z2 = s.t_max * s.b_world[2] + s.a_world[2]
        assert s.world_status == (1 if (z1 > floor or z2 > floor) else 2)
        kept += s.world_status == 1; trimmed += s.world_status == 2
    assert kept and trimmed
    # a rejected line is not transformed
    rc, s = oracle.segment_test(y, shift, a, b, _cfg(oracle_mod, (0.05,), 0.99999))
    oracle.lib.h3do_segment_to_world(C.byref(oracle_mod.Pose.make([1, 0, 0, 0], [0, 0, 0], 0.0)), C.byref(s))
    assert s.status == 2 and s.world_status == 0 and list(s.a_world) == [0.0, 0.0, 0.0]


def test_oracle_fixed_eigensolver(oracle):
    rng = np.random.default_rng(3)
    for _ in range(200):
        m = rng.normal(size=(3, 3)).astype(np.float32)
        m = (m @ m.T).astype(np.float32)
        rc, ev = oracle.eig3f_fixed(m)
        assert rc == 0
        ref = np.linalg.eigvalsh(m.astype(np.float64))
        assert np.allclose(ev, ref, rtol=5e-5, atol=5e-6 * abs(ref).max())
    rc, ev = oracle.eig3f_fixed(np.diag([3.0, 1.0, 2.0]).astype(np.float32))
    assert rc == 0 and list(ev) == [1.0, 2.0, 3.0]


def test_library_exports_segment_entry_points():
    lib = C.CDLL(capi.LIB_PATH)
    for name in ("h3d_pipeline_enable_segments", "h3d_pipeline_segments", "h3d_segment_test", "h3d_segment_test_pose",
                 "h3d_pipeline_set_pose", "h3d_pipeline_run_cloud2"):
        assert hasattr(lib, name)
