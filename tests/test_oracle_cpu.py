"""CPU suite (no GPU): the oracle against the golden vectors recorded from the UNMODIFIED
reference library, the host-side direction table of libh3d.so, and the C-ABI surface."""
import ctypes
import os
import re

import numpy as np
import pytest

import golden_util as G

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def gold():
    return G.load()


@pytest.mark.parametrize("g", range(0, 7))
def test_oracle_sphere_matches_reference(oracle, gold, g):
    v = oracle.sphere(g)
    assert v.shape[0] == (6, 21, 81, 321, 1281, 5121, 20481)[g]
    if g <= 4:
        assert np.array_equal(v, gold[f"sphere_g{g}"])
    else:
        assert G.sha(v) == str(gold[f"sphere_g{g}_sha256"])


@pytest.mark.parametrize("g", range(0, 7))
def test_product_sphere_matches_reference(h3d, oracle, gold, g):
    """h3d_sphere is pure host code (edge-keyed hash dedup): no device needed."""
    v, tri = h3d.sphere(g)
    if g <= 4:
        assert np.array_equal(v, gold[f"sphere_g{g}"])
    else:
        assert G.sha(v) == str(gold[f"sphere_g{g}_sha256"])
    assert np.array_equal(tri, oracle.sphere_triangles(g))


@pytest.mark.parametrize("key", G.CASES)
def test_oracle_accumulator_matches_reference(oracle, gold, key):
    pts = gold[f"{key}_points"]
    g, dx, _, _ = G.case_params(gold, key)
    x, _ = oracle.shift_to_origin(pts)
    mn, mx = oracle.min_max(x)
    h = oracle.hough(g, mn, mx, dx)
    h.add(x)
    votes = h.votes()
    v, idx, a, b = h.get_line()
    num_x, num_b, hdx, max_x, total, gv, gidx = gold[f"{key}_hough"]
    assert (h.num_x, h.num_b, h.dx, h.max_x) == (int(num_x), int(num_b), hdx, max_x)
    assert int(votes.sum()) == int(total) and (v, idx) == (int(gv), int(gidx))
    assert G.sha(votes) == str(gold[f"{key}_votes_sha256"])
    assert np.array_equal(np.concatenate([a, b]), gold[f"{key}_line0"])
    # multi-threaded variants (used by bench.py --impl reference) give the same answers
    h2 = oracle.hough(g, mn, mx, dx)
    h2.add(x, threads=4)
    assert np.array_equal(h2.votes(), votes)
    assert h2.get_line(threads=4)[:2] == (v, idx)


@pytest.mark.parametrize("key", G.CASES)
def test_oracle_loop_matches_reference(oracle, gold, key):
    pts = gold[f"{key}_points"]
    g, dx, minvotes, nlines = G.case_params(gold, key)
    tr = oracle.hough3dlines(pts, g, dx, minvotes, nlines, mode=1)
    header = {f: getattr(tr, f) for f in ("rc", "n_lines", "n_input", "num_x", "num_b", "d", "dx", "max_x", "shift")}
    G.check_trace(gold, key, header, [vars(it) for it in tr.iters])


def test_testdata_known_answers(gold):
    """values captured from the reference on lib/testdata.dat, dx 0.4, g4."""
    num_x, num_b, dx, max_x, total, votes, idx = gold["testdata_hough"]
    assert (int(num_x), int(num_b), int(total), int(votes)) == (32, 1281, 192150, 44)
    assert abs(max_x - 6.32481065) < 1e-8
    assert np.allclose(gold["testdata_line0"], [-2.32234245, 0.908403379, 1.19663145, 0.54064998, 0.581951963, 0.607477993])


def test_oracle_eig3f_against_numpy(oracle):
    rng = np.random.default_rng(0)
    for _ in range(50):
        a = rng.standard_normal((30, 3)).astype(np.float32) * rng.uniform(0.1, 5)
        m = (a.T @ a).astype(np.float32)
        rc, ev, evec = oracle.eig3f(m)
        assert rc == 0
        w, v = np.linalg.eigh(m.astype(np.float64))
        assert np.allclose(ev, w, rtol=2e-5, atol=2e-5 * abs(w).max())
        assert abs(abs(np.dot(evec[:, 2], v[:, 2])) - 1) < 1e-4


GOLD56 = os.path.join(ROOT, "tests", "golden", "reference_golden_g56.npz")


@pytest.mark.parametrize("key", ["synthetic_b_g5", "testdata_g6", "attpc_c_g6"])
def test_oracle_loop_matches_reference_at_g5_g6(oracle, key):
    """the reference's own fixtures at granularity 5 and 6 (traces recorded from oracle/_ref)"""
    gold = np.load(GOLD56)
    pts = gold[f"{key}_points"]
    g, dx, minvotes, nlines = G.case_params(gold, key)
    tr = oracle.hough3dlines(pts, g, dx, minvotes, nlines, mode=1)
    header = {f: getattr(tr, f) for f in ("rc", "n_lines", "n_input", "num_x", "num_b", "d", "dx", "max_x", "shift")}
    G.check_trace(gold, key, header, [vars(it) for it in tr.iters])


def test_oracle_tof_frame_matches_reference_at_g6(oracle):
    """one of the 16 bench frames (the smallest) through the C restatement: pre-filter + g6 loop, 20 lines"""
    from pointcloud_segmentation_b200 import synth
    gold = np.load(GOLD56)
    count, g, minvotes, nlines = (int(v) for v in gold["tof_params"])
    k = count - 1
    poses = synth.tower_trajectory(1000)
    pick = np.linspace(0, 999, count).astype(int)
    frame = synth.tof_frame(pose=tuple(poses[pick[k]]), seed=1000 + int(pick[k])).astype(np.float32)
    if G.sha(frame) != str(gold[f"tof{k}_raw_sha256"]):
        pytest.skip("this machine's libm generates other frame bits than the recorded ones")
    filt = oracle.prefilter(frame, synth.WINDOW, float(synth.LEAF))
    assert G.sha(filt) == str(gold[f"tof{k}_filtered_sha256"])
    tr = oracle.hough3dlines(filt.astype(np.float64), g, synth.OPT_DX, minvotes, nlines, mode=0)
    assert (tr.n_lines, len(tr.iters), tr.n_input) == tuple(int(gold[f"tof{k}_{f}"]) for f in ("n_lines", "n_iters", "n_input"))
    assert np.array_equal(np.array([it.votes for it in tr.iters]), gold[f"tof{k}_votes"])
    assert np.array_equal(np.array([it.argmax for it in tr.iters], dtype=np.uint64), gold[f"tof{k}_argmax"])
    assert [G.sha(np.asarray(it.inliers2, dtype=np.uint32)) for it in tr.iters] == [str(v) for v in gold[f"tof{k}_inl2"]]
    assert np.array_equal(np.array([it.a2 for it in tr.iters]), gold[f"tof{k}_a2"])
    assert np.array_equal(np.array([it.b2 for it in tr.iters]), gold[f"tof{k}_b2"])


def test_lsq_direction_within_north_star_tolerance_on_loop_inliers(oracle):
    """BASELINE.json north_star: fitted anchor/direction within 1e-5 relative.  The float pipeline of
    wrap.h:129-147 (restated: TREE32 sums + Eigen's tridiagonal QR) against float64 numpy.linalg.eigh on
    the well-conditioned inlier sets the loop itself fits (golden fixtures, every accepted line)."""
    gold = np.load(G.GOLD)
    checked = 0
    for key in G.CASES:
        pts = gold[f"{key}_points"]
        centred = pts - gold[f"{key}_trace_shift"]
        for k in range(int(gold[f"{key}_trace_n_iters"])):
            if int(gold[f"{key}_trace_it{k}_status"]) != 0:
                continue
            y = centred[gold[f"{key}_trace_it{k}_inliers2"]]
            rc, a, b = oracle.orthogonal_lsq(y)
            y32 = y.astype(np.float32).astype(np.float64)
            c = y32 - y32.mean(0)
            w, v = np.linalg.eigh(c.T @ c)
            if w[2] < 50 * w[1]:
                continue                      # an ill-conditioned top eigenvector is not a tolerance statement
            ref = v[:, 2] * np.sign(np.dot(v[:, 2], b))
            assert np.linalg.norm(b - ref) <= 1e-5 * np.linalg.norm(ref), (key, k, np.linalg.norm(b - ref))
            assert abs(rc - w[2]) <= 1e-5 * w[2], (key, k)
            assert np.linalg.norm(a - y.mean(0)) <= 1e-5 * max(1.0, np.linalg.norm(a)), (key, k)
            checked += 1
    assert checked >= 10


def test_oracle_prefilter_properties(oracle):
    from pointcloud_segmentation_b200 import synth
    f = synth.tof_frame()
    out = oracle.prefilter(f, synth.WINDOW, float(synth.LEAF))
    assert 1000 < len(out) < 10000
    w = synth.WINDOW
    assert (out[:, 0] >= w[0]).all() and (out[:, 0] <= w[1]).all()
    keys = np.floor(out * (np.float32(1.0) / np.float32(synth.LEAF))).astype(np.int64)
    # one centroid per voxel (a centroid may round onto a voxel face: allow a handful of ties)
    assert len(out) - len(np.unique(keys, axis=0)) <= 3
    again = oracle.prefilter(np.pad(out, ((0, 0), (0, 1))), w, float(synth.LEAF))
    assert len(again) == len(out)                               # idempotent in count


def test_capi_exports_every_declared_symbol(h3d):
    hdr = open(os.path.join(ROOT, "include", "h3d.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = sorted(set(re.findall(r"\b(h3d_[a-z0-9_]+)\s*\(", hdr)))
    assert len(names) >= 30
    lib = h3d.lib()
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/h3d.h but not exported by libh3d.so"


def test_capi_argument_errors_without_device(h3d):
    lib = h3d.lib()
    assert lib.h3d_sphere(ctypes.c_int(99), None, ctypes.c_size_t(0), None, None, ctypes.c_size_t(0), None) == 1
    assert b"granularity" in lib.h3d_last_error()
    assert lib.h3d_get_line(None, None, None, None, None) == 1
    assert lib.h3d_prefilter(None, ctypes.c_size_t(5), ctypes.c_size_t(3), None, ctypes.c_float(0.1), None,
                             ctypes.c_size_t(0), None) == 1


def test_product_fails_loudly_without_gpu(h3d):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(h3d.H3dError):
        h3d.init_space(2)
    with pytest.raises(h3d.H3dError):
        h3d.hough3dlines(np.random.default_rng(0).standard_normal((10, 3)), 2, 0.1, 2, 2)


def test_noalloc_entry_point_checks_its_arguments_before_touching_cuda(h3d):
    """h3d_points_close_to_line_noalloc: a cloud beyond the fixed scratch is refused with H3D_ERR_RANGE and a
    message that names the limit; null pointers with H3D_ERR_ARG -- both without a device (no compute call is made)"""
    dummy = 4096                                       # never dereferenced: the checks come first
    with pytest.raises(h3d.H3dError) as err:
        h3d.points_close_to_line_noalloc(dummy, h3d.NOALLOC_MAX_POINTS + 1, (0, 0, 0), (0, 0, 1), 0.1, dummy, dummy)
    assert err.value.code == 4 and "H3D_NOALLOC_MAX_POINTS" in str(err.value)
    with pytest.raises(h3d.H3dError) as err:
        h3d.points_close_to_line_noalloc(0, 10, (0, 0, 0), (0, 0, 1), 0.1, dummy, dummy)
    assert err.value.code == 1


def test_product_eigensolver_matches_oracle_on_host(tmp_path):
    """eig3f of the product (csrc/h3d_fit.cuh, its __host__ side) vs the oracle's h3do_eig3f."""
    import shutil
    import subprocess
    if not shutil.which("nvcc"):
        pytest.skip("nvcc not available")
    exe = str(tmp_path / "eig_check")
    subprocess.run(["nvcc", "-O2", "-std=c++17", "-fmad=false", "-gencode", "arch=compute_100a,code=sm_100a",
                    "-Xcompiler", "-ffp-contract=off", "-o", exe, os.path.join(ROOT, "tests", "cpp", "eig_host_check.cu"),
                    os.path.join(ROOT, "oracle", "h3d_oracle.c"), "-lm", "-lpthread"], check=True,
                   capture_output=True)
    r = subprocess.run([exe, "100000"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout
